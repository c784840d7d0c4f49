"""First-contact probe of the step kernel on a GPU box: binomial paths of growing size against the
oracle, printing what differs instead of asserting (tools/gpu_session.sh runs it before the tests)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402

from ate_replication_causalml_b200 import synth  # noqa: E402
from ate_replication_causalml_b200.engine import Engine  # noqa: E402
from oracle import oracle as O  # noqa: E402
from test_gpu_wide_binomial import unique_product_design  # noqa: E402

cases = [("n200 p5", 200, 5, None), ("n2000 p5", 2000, 5, None), ("n5000 p21", 5000, 21, None), ("n20000 p40", 20000, 40, None),
         ("uniq p100", 4000, 100, 1), ("uniq p256", 6000, 256, 1), ("uniq p344", 5000, 344, 1), ("uniq p512", 6000, 512, 1)]
only = os.environ.get("PROBE_ONLY")
eng = Engine(0)
bad = 0
for name, n, p, uniq in cases:
    if only and only not in name:
        continue
    if uniq:
        X, y = unique_product_design(n, p, seed=300 + p)
        kw = dict(nlambda=40)
    else:
        d = synth.make_dataset(n, p)
        X, _, _ = O.scale(d["x_raw"], cols=np.arange(d["n_continuous"]))
        y = d["w"]
        kw = {}
    t = time.time()
    try:
        g = eng.glmnet(X, y, "binomial", **kw)
    except Exception as e:  # noqa: BLE001
        print(f"{name}: ENGINE FAILED: {e}", flush=True)
        bad += 1
        break
    tg = time.time() - t
    o = O.glmnet(X, y, "binomial", **kw)
    L = min(g["lmu"], o["lmu"])
    db = np.abs(g["beta"][:, :L] - o["beta"][:, :L]).max() if L else float("nan")
    da = np.abs(g["a0"][:L] - o["a0"][:L]).max() if L else float("nan")
    first = next((l for l in range(L) if np.abs(g["beta"][:, l] - o["beta"][:, l]).max() > 1e-8), -1)
    okk = g["lmu"] == o["lmu"] and g["npasses"] == o["npasses"] and db < 1e-8
    bad += 0 if okk else 1
    print(f"{name}: lmu {g['lmu']}/{o['lmu']} npasses {g['npasses']}/{o['npasses']} jerr {g['jerr']}/{o['jerr']} "
          f"max|dbeta| {db:.2e} max|da0| {da:.2e} first bad lambda {first} ({tg:.2f}s) {'OK' if okk else 'MISMATCH'}", flush=True)
eng.close()
sys.exit(1 if bad else 0)
