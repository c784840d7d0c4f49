"""First GPU probe: correctness of the early kernels against numpy + Gram timing + DGEMM peak."""
import ctypes as C, sys, time, json, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from ate_replication_causalml_b200 import _lib as L

lib = C.CDLL(L.SO_PATH)
def bind(name):
    res, args = L.SIGNATURES[name]; fn = getattr(lib, name); fn.restype = res; fn.argtypes = args; return fn
ctx = C.c_void_p()
rc = bind("ate_ctx_create")(0, C.byref(ctx)); print("ctx", rc)
err = bind("ate_last_error")
def chk(rc, what):
    if rc != 0: print(what, "FAILED rc", rc, err(ctx)); sys.exit(1)
rng = np.random.default_rng(1)
# --- score
n = 100003
w = (rng.random(n) < 0.4).astype(float); y = (rng.random(n) < 0.3).astype(float)
p = rng.uniform(0.05, 0.95, n); m0 = rng.uniform(0.1, 0.5, n); m1 = m0 + rng.uniform(-0.1, 0.3, n)
tau = C.c_double(); se = C.c_double(); e1 = np.empty(n); e2 = np.empty(n)
chk(bind("ate_aipw_score")(ctx, n, L.ptr(w), L.ptr(y), L.ptr(p), L.ptr(m0), L.ptr(m1), C.byref(tau), C.byref(se), L.ptr(e1), L.ptr(e2), 0), "score")
est1 = w*(y-m1)/p + (1-w)*(y-m0)/(1-p); est2 = m1-m0; t_ref = est1.mean()+est2.mean()
Ii = (w*y)/p - m1*(w-p)/p - (((1-w)*y/(1-p)) + (m0*(w-p)/(1-p))) - t_ref
se_ref = np.sqrt((Ii**2).sum())/n
print("score tau rel", abs(tau.value-t_ref)/abs(t_ref), "se rel", abs(se.value-se_ref)/se_ref, "e1", np.abs(e1-est1).max(), np.abs(e2-est2).max())
# --- bootstrap index mode
B = 64
idx = rng.integers(1, n+1, size=(B, n)).astype(np.int32)
boot = np.empty(B); sd = C.c_double()
chk(bind("ate_bootstrap")(ctx, n, L.ptr(w), L.ptr(y), L.ptr(p), L.ptr(m0), L.ptr(m1), B, 0, L.ptr(idx), 0, 0, L.ptr(boot), C.byref(sd), 0), "boot")
ref = np.array([est1[idx[b]-1].mean()+est2[idx[b]-1].mean() for b in range(B)])
print("boot idx max rel", np.abs(boot-ref).max()/np.abs(ref).max(), "sd rel", abs(sd.value-ref.std(ddof=1))/ref.std(ddof=1))
boot2 = np.empty(B)
chk(bind("ate_bootstrap")(ctx, n, L.ptr(w), L.ptr(y), L.ptr(p), L.ptr(m0), L.ptr(m1), B, 1, None, 1234, 0, L.ptr(boot2), C.byref(sd), 0), "boot philox")
print("boot philox mean", boot2.mean(), "sd", sd.value, "(index-mode sd", ref.std(ddof=1), ")")
# --- standardize / expand
n2, p0 = 5001, 7
x = np.asfortranarray(rng.standard_normal((n2, p0))*3+1)
xs = x.copy(order="F"); cen = np.empty(4); scl = np.empty(4); cols = np.array([0,2,3,5], dtype=np.int32)
chk(bind("ate_standardize")(ctx, L.ptr(xs), n2, p0, L.ptr(cols), 4, L.ptr(cen), L.ptr(scl)), "std")
ref = x.copy(); ref[:, cols] = (x[:, cols]-x[:, cols].mean(0))/x[:, cols].std(0, ddof=1)
print("standardize max abs", np.abs(xs-ref).max())
out = np.empty((n2, p0*p0+p0), order="F")
chk(bind("ate_expand_deg2")(ctx, L.ptr(x), n2, p0, 1, L.ptr(out)), "expand")
refe = np.concatenate([x] + [ (x[:, [c1]]*x) for c1 in range(p0)], axis=1)
print("expand max abs", np.abs(out-refe).max())
# --- gram small (host API), with rows + weights
for (ng, pg) in [(1000, 40), (4097, 256), (3000, 70)]:
    xg = np.asfortranarray(rng.standard_normal((ng, pg)))
    rows = np.sort(rng.choice(ng, ng*3//4, replace=False)).astype(np.int32)
    v = rng.uniform(0.1, 1, rows.size)
    G = np.empty((pg, pg))
    chk(bind("ate_weighted_gram")(ctx, ng, pg, L.ptr(xg), 0, L.ptr(v), L.ptr(rows), rows.size, L.ptr(G), 0), "gram")
    Gr = (xg[rows]*v[:, None]).T @ xg[rows]
    print("gram", ng, pg, "max rel err", np.abs(G-Gr).max()/np.abs(Gr).max())
# --- gram timing on device-resident row-major design
dev = torch.device("cuda:0")
res = {}
for (ng, pg) in [(1_000_000, 256), (800_000, 256), (1_000_000, 128)]:
    ld = (pg+31)//32*32
    X = torch.randn(ng, ld, dtype=torch.float64, device=dev)
    v = torch.rand(ng, dtype=torch.float64, device=dev)
    G = torch.empty(pg, pg, dtype=torch.float64, device=dev)
    f = bind("ate_weighted_gram")
    for _ in range(3): chk(f(ctx, ng, pg, X.data_ptr(), ld, v.data_ptr(), None, ng, G.data_ptr(), 1), "gram dev")
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(10): f(ctx, ng, pg, X.data_ptr(), ld, v.data_ptr(), None, ng, G.data_ptr(), 1)
    torch.cuda.synchronize(); dt = (time.perf_counter()-t)/10
    Gr = (X[:, :pg]*v[:, None]).T @ X[:, :pg]
    errg = ((G-Gr).abs().max()/Gr.abs().max()).item()
    fl = ng*pg*(pg+1)
    print(f"gram dev n={ng} p={pg}: {dt*1e3:.3f} ms  {fl/dt/1e12:.2f} TFLOP/s(alg)  {ng*(pg+1)*8/dt/1e9:.0f} GB/s  err {errg:.2e}")
    res[f"gram_{ng}_{pg}"] = dict(ms=dt*1e3, tflops_alg=fl/dt/1e12, err=errg)
    del X, v, G, Gr
# --- DGEMM peak via cuBLAS
for N in (4096, 8192):
    a = torch.randn(N, N, dtype=torch.float64, device=dev); b = torch.randn(N, N, dtype=torch.float64, device=dev)
    for _ in range(2): (a@b)
    torch.cuda.synchronize(); best = 1e9
    for _ in range(5):
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        s.record(); c = a@b; e.record(); torch.cuda.synchronize(); best = min(best, s.elapsed_time(e))
    print(f"cuBLAS DGEMM {N}^3: {2*N**3/best/1e9:.2f} TFLOP/s"); res[f"dgemm_{N}"] = 2*N**3/best/1e9
# syrk-like shape: (256 x 1M) @ (1M x 256)
a = torch.randn(1_000_000, 256, dtype=torch.float64, device=dev)
for _ in range(2): (a.T@a)
torch.cuda.synchronize(); best = 1e9
for _ in range(5):
    s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    s.record(); c = a.T@a; e.record(); torch.cuda.synchronize(); best = min(best, s.elapsed_time(e))
print(f"cuBLAS X^T X (1M x 256): {best:.3f} ms -> {2*1e6*256*256/best/1e9:.2f} TFLOP/s full-square"); res["dgemm_xtx_ms"] = best
json.dump(res, open("gpurun_out/probe1.json", "w"), indent=1)
bind("ate_ctx_destroy")(ctx)
