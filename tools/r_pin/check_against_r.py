"""Step 3 of tools/r_pin/README.md: compare oracle/ate_oracle.c with the outputs of real R + glmnet."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402

d = sys.argv[1] if len(sys.argv) > 1 else "r_pin_out"
rd = lambda f: np.loadtxt(os.path.join(d, f), delimiter=",", ndmin=2)  # noqa: E731
rs = lambda f: np.array(open(os.path.join(d, f)).read().split(), dtype=float)  # noqa: E731
print(open(os.path.join(d, "R_versions.txt")).read().strip())
worst = 0.0


def report(what, a, b, scale=None):
    global worst
    a, b = np.asarray(a, float), np.asarray(b, float)
    if a.shape != b.shape:
        print(f"  {what:34s} SHAPE {a.shape} vs {b.shape}")
        worst = np.inf
        return
    err = np.abs(a - b).max() / (scale if scale else max(1.0, np.abs(b).max()))
    worst = max(worst, err)
    print(f"  {what:34s} max rel dev {err:.3e}")


for tag in ("a", "b"):
    print("fixture", tag)
    x_raw = rd(f"{tag}_x_raw.csv")
    w = np.loadtxt(os.path.join(d, f"{tag}_w.csv")); y = np.loadtxt(os.path.join(d, f"{tag}_y.csv"))
    fid = np.loadtxt(os.path.join(d, f"{tag}_foldid.csv")).astype(np.int32)
    ncont = int(open(os.path.join(d, f"{tag}_meta.txt")).read().split()[1])
    xs, _, _ = O.scale(x_raw, cols=np.arange(ncont))
    report("scale()", xs, rd(f"{tag}_R_scaled.csv"))
    p0 = xs.shape[1]
    xw = np.column_stack([xs, w]); vp = np.r_[np.ones(p0), 0.0]
    for fam, X, Y, kw, f in (("gaussian", xw, y, dict(penalty_factor=vp), "gauss"), ("binomial", xs, w, {}, "binom")):
        R = rd(f"{tag}_R_{f}_path.csv"); L = R.shape[1]
        o = O.glmnet(X, Y, fam, **kw)
        print(f"  {fam}: lmu oracle {o['lmu']} R {L}; npasses oracle {o['npasses']} R {int(rs(f'{tag}_R_{f}_scalars.txt')[0])}")
        m = min(L, o["lmu"])
        report(f"{fam} lambda", o["lambda_"][:m], R[0, :m])
        report(f"{fam} a0", o["a0"][:m], R[1, :m])
        report(f"{fam} beta", o["beta"][:, :m], R[2:-1, :m])
        report(f"{fam} dev.ratio", o["dev_ratio"][:m], R[-1, :m])
        C = rd(f"{tag}_R_cv_{f}.csv"); sel = rs(f"{tag}_R_cv_{f}_sel.txt")
        c = O.cv_glmnet(X, Y, fid, fam, **kw); m = min(C.shape[1], c["lmu"])
        report(f"cv {fam} cvm", c["cvm"][:m], C[1, :m]); report(f"cv {fam} cvsd", c["cvsd"][:m], C[2, :m])
        report(f"cv {fam} lambda.min/1se", [c["lambda_min"], c["lambda_1se"]], sel)
    gm = O.glm_binomial(np.column_stack([xs, w]), y); gp = O.glm_binomial(xs, w)
    report("glm outcome coef", gm["coef"], rd(f"{tag}_R_glm_outcome.csv").ravel())
    report("glm propensity coef", gp["coef"], rd(f"{tag}_R_glm_ps.csv").ravel())
    sc = rs(f"{tag}_R_glm_scalars.txt")
    print(f"  glm iterations oracle {gm['iters']},{gp['iters']} R {int(sc[0])},{int(sc[2])}")
    lm = O.lm(np.column_stack([xs, w]), y); Rl = rd(f"{tag}_R_lm.csv")
    report("lm Estimate", lm["coef"], Rl[:, 0]); report("lm Std. Error", lm["std_err"], Rl[:, 1])
S = rd("R_sample_streams.csv").astype(int)
for kind, row in ((1, 0), (0, 1)):
    O.set_seed(1991)
    print("  sample() kind", kind, "bit-exact:", bool((O.sample_replace(50000, 20, kind) == S[row]).all()))
print("worst relative deviation:", worst)
