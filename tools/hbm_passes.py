"""Drive every HBM-bound pass of the path at sizes where bandwidth is what is measured (for the ncu
captures under profiles/; also prints CUDA-event timings of the device-resident ones).
  python tools/hbm_passes.py            # timings
  ncu --set full -k regex:'col_|build_design|expand_col|aipw_score|bootstrap_kernel|gather_rows|cvdev|predict_rows' ...
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from ate_replication_causalml_b200 import synth  # noqa: E402
from ate_replication_causalml_b200.engine import Engine  # noqa: E402

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from conftest import make_folds  # noqa: E402

dev = torch.device("cuda:0")
eng = Engine(0)
stream = torch.cuda.ExternalStream(eng.stream_ptr, device=dev)
peak = 6549.1


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        fn()
    e1.record(stream)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


rng = np.random.default_rng(0)
# a7/a8: score on device vectors (5 n 8 B read + 3 n 8 B written)
n = 32_000_000
g = torch.Generator(device=dev); g.manual_seed(1)
w = (torch.rand(n, device=dev, generator=g, dtype=torch.float64) < 0.4).double()
y = (torch.rand(n, device=dev, generator=g, dtype=torch.float64) < 0.3).double()
p = torch.rand(n, device=dev, generator=g, dtype=torch.float64) * 0.8 + 0.1
m0 = torch.rand(n, device=dev, generator=g, dtype=torch.float64) * 0.5
m1 = m0 + 0.1
ms = timed(lambda: eng.aipw_score_device(n, w.data_ptr(), y.data_ptr(), p.data_ptr(), m0.data_ptr(), m1.data_ptr()))
print(f"aipw_score_kernel n={n}: {ms:.3f} ms incl. result readback, {5 * n * 8 / ms / 1e6:.0f} GB/s algorithmic ({5 * n * 8 / ms / 1e6 / peak:.2f} of measured HBM)", flush=True)
del w, y, p, m0, m1
# a9/a10 validation mode: index matrix on the device
n, Bi = 1_000_000, 1500
w = (torch.rand(n, device=dev, generator=g, dtype=torch.float64) < 0.4).double()
y = (torch.rand(n, device=dev, generator=g, dtype=torch.float64) < 0.3).double()
p = torch.rand(n, device=dev, generator=g, dtype=torch.float64) * 0.8 + 0.1
m0 = torch.rand(n, device=dev, generator=g, dtype=torch.float64) * 0.5
m1 = m0 + 0.1
idx = torch.randint(1, n + 1, (Bi, n), device=dev, dtype=torch.int32, generator=g)
ptrs = [t.data_ptr() for t in (w, y, p, m0, m1)]
ms = timed(lambda: eng.bootstrap(*ptrs, B=Bi, idx=idx.data_ptr(), on_device=True, n=n), reps=3)
print(f"bootstrap_kernel<index> B={Bi} n={n}: {ms:.3f} ms incl. score + readback, {Bi * n * 4 / ms / 1e6:.0f} GB/s of index stream ({Bi * n * 4 / ms / 1e6 / peak:.2f} of measured HBM)", flush=True)
ms = timed(lambda: eng.bootstrap(*ptrs, B=2000, seed=5, on_device=True, n=n), reps=3)
print(f"bootstrap_kernel<philox> B=2000 n={n}: {ms:.3f} ms = {2000 * n / ms / 1e6:.0f} Gdraws/s", flush=True)
del idx, w, y, p, m0, m1
torch.cuda.empty_cache()
# a1: scale() (host API: the copies are outside what ncu attributes to the kernels)
x = rng.standard_normal((4_000_000, 16))
t = time.time(); eng.standardize(x); print(f"ate_standardize n=4M p=16 (host buffers): {1e3 * (time.time() - t):.0f} ms wall", flush=True)
# a2: expansion to column-major (public entry)
x = rng.standard_normal((1_000_000, 16))
t = time.time(); eng.expand_deg2(x, include_main=False); print(f"ate_expand_deg2 n=1M p0=16 (host buffers): {1e3 * (time.time() - t):.0f} ms wall", flush=True)
# the composite at n = 400k: build_design, fold lists, gather_rows, predict, cvdev
n, p0 = 400_000, 16
d = synth.make_dataset(n, p0)
fo, ips, iout = make_folds(n)
t = time.time()
r = eng.crossfit_aipw(d["x_raw"], d["w"], d["y"], foldid=fo, inner_ps=ips, inner_out=iout, n_continuous=d["n_continuous"], expand=1, B=200, boot_mode=1)
print(f"crossfit n={n}: {1e3 * (time.time() - t):.0f} ms wall, tau {r['tau']:.6f}; stages {r['ms_design']:.2f} / {r['ms_folds']:.2f} / {r['ms_outcome']:.1f} / {r['ms_propensity']:.1f} / {r['ms_score']:.2f} / {r['ms_bootstrap']:.2f} ms", flush=True)
eng.close()
