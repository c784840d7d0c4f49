"""Micro-benchmark of the fused IRLS step: a short binomial path on n x p synthetic data, per-launch
times from ATE_PROFILE_CSV.  ATE_STEP_DEBUG toggles parts of the kernel (profiling experiments)."""
import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from ate_replication_causalml_b200.engine import Engine
n, p = int(sys.argv[1]), int(sys.argv[2])
rng = np.random.default_rng(0)
x = rng.standard_normal((n, p)); beta = np.zeros(p); beta[:12] = rng.choice([-0.3, 0.3], 12)
w = (rng.random(n) < 1/(1+np.exp(-(x @ beta - 0.4)))).astype(float)
eng = Engine(0)
eng.glmnet(x[:10000], w[:10000], "binomial", nlambda=5)   # warm
eng.profile_enable(True)
r = eng.glmnet(x, w, "binomial", nlambda=int(sys.argv[3]) if len(sys.argv) > 3 else 12)
os.environ["ATE_PROFILE_CSV"] = "gpurun_out/micro.csv"
print(eng.profile_read(), "lmu", r["lmu"], "nnz last", int((r["beta"][:, -1] != 0).sum()))
