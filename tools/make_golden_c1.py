"""Mint tests/golden/c1_design_crossfit.npz: the CPU oracle's cross-fitted AIPW result on bench.py's
C1 DESIGN (p0 = 16 -> the 256 ordered degree-2 products of ate_functions.R:290-296, 120 twin pairs)
at n = 8000 with thresh = 1e-13, where both the oracle and the engine sit on the unique optimum of
the fitted values.  Takes ~2 minutes on 4 threads, which is why the GPU test reads the fixture
instead of running the oracle on the GPU box (tests/test_gpu_wide_binomial.py).

    python tools/make_golden_c1.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import make_folds  # noqa: E402
from oracle import oracle as O  # noqa: E402
from ate_replication_causalml_b200 import synth  # noqa: E402

N, P0, THRESH = 8000, 16, 1e-13
d = synth.make_dataset(N, P0)
fo, ips, iout = make_folds(N)
xs, _, _ = O.scale(d["x_raw"], cols=np.arange(d["n_continuous"]))
xd = O.expand_deg2(xs, False)
o = O.crossfit_aipw(xd, d["w"], d["y"], fo, ips, iout, thresh=THRESH)
assert o["rc"] == 0
out = os.path.join(ROOT, "tests", "golden", "c1_design_crossfit.npz")
np.savez_compressed(out, n=N, p0=P0, thresh=THRESH, tau=o["tau"], se=o["se"], lambda_ps=o["lambda_ps"],
                    lambda_out=o["lambda_out"], phat=o["phat"], m0=o["m0"], m1=o["m1"])
print("written", out, os.path.getsize(out), "bytes; tau", o["tau"], "se", o["se"])
