"""ate_aipw_score_static(): the AIPW score entry point that performs no device allocation.

It exists for callers that cannot call cudaMalloc -- code running inside a memory pool that
already owns the whole GPU, or on a latency-critical path where cudaMalloc's implicit device
synchronisation is not acceptable.  Its block partials live in a fixed 48 KB array inside the
library instead of the context's growable scratch, so it accepts at most 2^28 rows, and all calls
on a device share that array (they must not overlap).

The test drives it through the C ABI with device buffers it owns, twice in a row -- the second
call finds the first call's leftovers in the shared array -- and checks both against the
allocating entry point ate_aipw_score().  Collected last on purpose (file name), so a run that
stops at the first failure has already run every other test; it also passes on its own in a
fresh process, where its first call is the one that performs the per-device set-up.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _inputs(n, seed):
    rng = np.random.default_rng(seed)
    w = (rng.random(n) < 0.4).astype(np.float64)
    y = (rng.random(n) < 0.3).astype(np.float64)
    p = rng.uniform(0.03, 0.97, n)
    m0 = rng.uniform(0.05, 0.6, n)
    m1 = np.clip(m0 + rng.uniform(-0.1, 0.3, n), 0.01, 0.99)
    return w, y, p, m0, m1


def _static(engine, n, ptrs):
    from ate_replication_causalml_b200 import _lib as L
    tau, se = C.c_double(), C.c_double()
    rc = engine._lib.ate_aipw_score_static(engine._h, n, *[L.ptr(int(q)) for q in ptrs], C.byref(tau), C.byref(se))
    return rc, tau.value, se.value


@pytest.mark.parametrize("n", [1, 1000, 300_001, 2_000_000])
def test_static_scratch_score_matches_allocating_entry_point_twice(engine, n):
    import torch
    dev = torch.device("cuda:0")
    for call, seed in enumerate((11, 12)):            # two calls in a row, different data
        host = _inputs(n, seed)
        bufs = [torch.from_numpy(a).to(dev) for a in host]
        ptrs = [b.data_ptr() for b in bufs]
        rc, tau, se = _static(engine, n, ptrs)
        assert rc == 0, engine._lib.ate_last_error(engine._h)
        ref_tau, ref_se = engine.aipw_score_device(n, *ptrs)
        # same tolerances as tests/test_gpu_kernels.py::test_aipw_score_matches_oracle
        assert abs(tau - ref_tau) <= 1e-12 * max(1.0, abs(ref_tau)), (call, tau, ref_tau)
        assert abs(se - ref_se) <= 1e-10 * ref_se + 1e-14, (call, se, ref_se)


def test_static_scratch_score_rejects_oversized_input(engine):
    import torch
    dev = torch.device("cuda:0")
    bufs = [torch.zeros(8, dtype=torch.float64, device=dev) + 0.5 for _ in range(5)]
    ptrs = [b.data_ptr() for b in bufs]
    rc, _, _ = _static(engine, (1 << 28) + 1, ptrs)    # beyond the stated maximum: refused, nothing launched
    assert rc < 0
    msg = engine._lib.ate_last_error(engine._h).decode()
    assert "static-scratch limit" in msg
    rc, tau, _ = _static(engine, 8, ptrs)              # and the entry point is still usable afterwards
    assert rc == 0 and np.isfinite(tau)
