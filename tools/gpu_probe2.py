"""GPU probe 2: glmnet paths / cv.glmnet / crossfit against the CPU oracle."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from ate_replication_causalml_b200.engine import Engine
from ate_replication_causalml_b200 import synth
from oracle import oracle as O

eng = Engine(0)
def rel(a, b): return float(np.abs(np.asarray(a)-np.asarray(b)).max()/max(1e-300, np.abs(np.asarray(b)).max()))
for (n, p0) in [(2000, 5), (5000, 21)]:
    d = synth.make_dataset(n, p0)
    x, _, _ = O.scale(d["x_raw"], cols=np.arange(d["n_continuous"]))
    w, y = d["w"], d["y"]
    # gaussian with pf: [x, w] -> y
    xw = np.column_stack([x, w]); vp = np.ones(p0+1); vp[-1] = 0
    t = time.time(); g = eng.glmnet(xw, y, "gaussian", penalty_factor=vp); tg = time.time()-t
    o = O.glmnet(xw, y, "gaussian", penalty_factor=vp)
    print(f"[gauss n={n} p={p0+1}] lmu {g['lmu']} vs {o['lmu']} jerr {g['jerr']} passes {g['npasses']} vs {o['npasses']} t={tg:.3f}s")
    L = min(g['lmu'], o['lmu'])
    print("   lambda rel", rel(g['lambda_'][:L], o['lambda_'][:L]), "beta maxabs", np.abs(g['beta'][:, :L]-o['beta'][:, :L]).max(), "a0", np.abs(g['a0'][:L]-o['a0'][:L]).max(), "dev", np.abs(g['dev_ratio'][:L]-o['dev_ratio'][:L]).max(), "nulldev rel", rel(g['nulldev'], o['nulldev']))
    # binomial: x -> w
    t = time.time(); g = eng.glmnet(x, w, "binomial"); tg = time.time()-t
    o = O.glmnet(x, w, "binomial")
    print(f"[binom n={n} p={p0}] lmu {g['lmu']} vs {o['lmu']} jerr {g['jerr']} passes {g['npasses']} vs {o['npasses']} t={tg:.3f}s")
    L = min(g['lmu'], o['lmu'])
    print("   lambda rel", rel(g['lambda_'][:L], o['lambda_'][:L]), "beta maxabs", np.abs(g['beta'][:, :L]-o['beta'][:, :L]).max(), "a0", np.abs(g['a0'][:L]-o['a0'][:L]).max(), "dev", np.abs(g['dev_ratio'][:L]-o['dev_ratio'][:L]).max(), "nulldev rel", rel(g['nulldev'], o['nulldev']))
    # cv
    fid = synth.fold_ids(n, 10, 7)
    for fam, (xx, yy, pf) in {"gaussian": (xw, y, vp), "binomial": (x, w, None)}.items():
        t = time.time(); g = eng.cv_glmnet(xx, yy, fam, penalty_factor=pf, foldid=fid); tg = time.time()-t
        o = O.cv_glmnet(xx, yy, fid, fam, penalty_factor=pf)
        L = min(g['lmu'], o['lmu'])
        print(f"[cv {fam}] lmu {g['lmu']}/{o['lmu']} idx_min {g['idx_min']}/{o['idx_min']} idx_1se {g['idx_1se']}/{o['idx_1se']} cvm rel {rel(g['cvm'][:L], o['cvm'][:L]):.2e} cvsd rel {rel(g['cvsd'][:L], o['cvsd'][:L]):.2e} beta {np.abs(g['beta'][:, :L]-o['beta'][:, :L]).max():.2e} t={tg:.3f}s")
    # predict
    pr = eng.predict(x, g['a0'][3], g['beta'][:, 3], True); po = O.predict(x, g['a0'][3], g['beta'][:, 3], True)
    print("   predict maxabs", np.abs(pr-po).max())
    # crossfit
    K = 5
    fo = synth.fold_ids(n, K, 11)
    ips = np.zeros((K, n), np.int32); iout = np.zeros((K, n), np.int32)
    for k in range(K):
        nt = int((fo != k+1).sum())
        ips[k, :nt] = synth.fold_ids(nt, 10, 100+k); iout[k, :nt] = synth.fold_ids(nt, 10, 200+k)
    for expand in (0, 2) if p0 <= 5 else (0,):
        xd = x if expand == 0 else O.expand_deg2(x, True)
        t = time.time()
        r = eng.crossfit_aipw(d["x_raw"], w, y, foldid=fo, inner_ps=ips, inner_out=iout, K=K, n_continuous=d["n_continuous"], expand=expand, B=50, boot_mode=1)
        tg = time.time()-t
        t = time.time(); oc = O.crossfit_aipw(xd, w, y, fo, ips, iout); to = time.time()-t
        print(f"[crossfit n={n} p0={p0} expand={expand}] tau {r['tau']:.8f} vs {oc['tau']:.8f} rel {abs(r['tau']-oc['tau'])/abs(oc['tau']):.2e}; se {r['se_sandwich']:.8f} vs {oc['se']:.8f} rel {abs(r['se_sandwich']-oc['se'])/oc['se']:.2e}; phat {np.abs(r['phat']-oc['phat']).max():.2e} m0 {np.abs(r['m0']-oc['m0']).max():.2e} m1 {np.abs(r['m1']-oc['m1']).max():.2e}")
        print(f"   lambda_ps rel {rel(r['lambda_ps'], oc['lambda_ps']):.2e} lambda_out rel {rel(r['lambda_out'], oc['lambda_out']):.2e} se_boot {r['se_bootstrap']:.6f} rounds {r['rounds_irls']} gpu {tg:.2f}s (design {r['ms_design']:.0f} out {r['ms_outcome']:.0f} ps {r['ms_propensity']:.0f} score {r['ms_score']:.1f} boot {r['ms_bootstrap']:.1f} ms) oracle {to:.2f}s")
print("launches", eng.launches)
