import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from ate_replication_causalml_b200.engine import Engine
from ate_replication_causalml_b200 import synth
from oracle import oracle as O
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from conftest import make_folds
eng = Engine(0)
n, p0, K = 3000, 6, 5
d = synth.make_dataset(n, p0)
fo, ips, iout = make_folds(n)
r = eng.crossfit_aipw(d["x_raw"], d["w"], d["y"], foldid=fo, inner_ps=ips, inner_out=iout, K=K, n_continuous=d["n_continuous"], expand=2, B=0)
xs, _, _ = O.scale(d["x_raw"], cols=np.arange(d["n_continuous"]))
xd = O.expand_deg2(xs, True)
o = O.crossfit_aipw(xd, d["w"], d["y"], fo, ips, iout)
print("tau", r["tau"], o["tau"])
print("lambda_ps gpu", r["lambda_ps"]); print("lambda_ps orc", o["lambda_ps"])
print("lambda_out gpu", r["lambda_out"]); print("lambda_out orc", o["lambda_out"])
for k in range(1, K+1):
    m = fo == k
    print(k, "phat", np.abs(r["phat"][m]-o["phat"][m]).max(), "m0", np.abs(r["m0"][m]-o["m0"][m]).max(), "m1", np.abs(r["m1"][m]-o["m1"][m]).max())
# per outer fold: compare cv_glmnet on training rows
for k in range(1, K+1):
    tr = fo != k
    nt = tr.sum()
    xt = xd[tr]; wt = d["w"][tr]; yt = d["y"][tr]
    for fam, (xx, yy, pf, fid) in {"binomial": (xt, wt, None, ips[k-1, :nt]), "gaussian": (np.column_stack([xt, wt]), yt, np.r_[np.ones(xd.shape[1]), 0.0], iout[k-1, :nt])}.items():
        g = eng.cv_glmnet(xx, yy, fam, penalty_factor=pf, foldid=fid)
        oc = O.cv_glmnet(xx, yy, fid, fam, penalty_factor=pf)
        L = min(g["lmu"], oc["lmu"])
        print(f"k={k} {fam}: lmu {g['lmu']}/{oc['lmu']} imin {g['idx_min']}/{oc['idx_min']} i1se {g['idx_1se']}/{oc['idx_1se']} beta {np.abs(g['beta'][:, :L]-oc['beta'][:, :L]).max():.2e} cvm {np.abs(g['cvm'][:L]-oc['cvm'][:L]).max():.2e} fitted(1se) {np.abs(xx@g['beta'][:, g['idx_1se']]+g['a0'][g['idx_1se']] - xx@oc['beta'][:, oc['idx_1se']]-oc['a0'][oc['idx_1se']]).max():.2e}")
        if np.abs(g['beta'][:, :L]-oc['beta'][:, :L]).max() > 1e-9:
            db = np.abs(g['beta'][:, :L]-oc['beta'][:, :L]).max(0)
            first = int(np.argmax(db > 1e-9))
            print("    first differing lambda idx", first, "cols", np.nonzero(np.abs(g['beta'][:, first]-oc['beta'][:, first]) > 1e-9)[0], "gpu", g['beta'][:, first][np.abs(g['beta'][:, first]-oc['beta'][:, first]) > 1e-9], "orc", oc['beta'][:, first][np.abs(g['beta'][:, first]-oc['beta'][:, first]) > 1e-9])
