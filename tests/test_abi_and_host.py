"""CPU-side checks: the C-ABI library loads and exports every symbol include/ate_b200.h
declares, fails loudly without a GPU, and the host logic (sharding, sd, interpolation,
2-rank gloo exchange) behaves.  No compute calls are made here."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "ate_b200.h")).read()
    return sorted(set(re.findall(r"ATE_API\s+[\w\s\*]+?\b(ate_\w+)\s*\(", src)))


def test_header_declares_expected_surface():
    syms = _declared_symbols()
    for must in ("ate_ctx_create", "ate_standardize", "ate_expand_deg2", "ate_glmnet", "ate_cv_glmnet",
                 "ate_predict", "ate_aipw_score", "ate_bootstrap", "ate_tau_hat_dr_est", "ate_crossfit_aipw",
                 "ate_weighted_gram", "ate_r_sample", "ate_r_foldid"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    from ate_replication_causalml_b200 import build as b
    from ate_replication_causalml_b200 import _lib
    b.build()
    lib = C.CDLL(_lib.SO_PATH)
    for name in _declared_symbols():
        assert hasattr(lib, name), f"{name} declared in include/ate_b200.h but not exported"
        assert name in _lib.SIGNATURES, f"{name} has no ctypes signature"
    assert set(_lib.SIGNATURES) == set(_declared_symbols())
    _lib.load()
    assert lib.ate_version() >= 100


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from ate_replication_causalml_b200.engine import Engine
    from ate_replication_causalml_b200._lib import AteError
    with pytest.raises(AteError):
        Engine(0)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "ate_replication_causalml_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
                assert "libate_oracle" not in txt and "ate_oracle.c" not in txt, f


def test_replicate_shards_tile_the_range():
    from ate_replication_causalml_b200.parallel import replicate_shard
    for B in (1, 7, 2000, 10000):
        for world in (1, 2, 3, 4, 8):
            spans = [replicate_shard(B, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == B
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def test_r_sd_matches_numpy():
    from ate_replication_causalml_b200.parallel import r_sd
    x = np.random.default_rng(0).standard_normal(1000) * 3 + 100
    assert abs(r_sd(x) - x.std(ddof=1)) < 1e-12


def test_coef_interpolation_at_and_between_lambdas():
    from ate_replication_causalml_b200.ate_functions import _coef_at
    m = dict(lambda_=np.array([1.0, 0.5, 0.25]), beta=np.array([[0.0, 1.0, 3.0]]))
    assert _coef_at(m, 0.5)[0] == 1.0
    assert _coef_at(m, 2.0)[0] == 0.0 and _coef_at(m, 0.1)[0] == 3.0
    assert abs(_coef_at(m, 0.75)[0] - 0.5) < 1e-12


def test_synthetic_inputs_are_deterministic_and_shaped():
    from ate_replication_causalml_b200 import synth
    a = synth.make_dataset(500, 21)
    b = synth.make_dataset(500, 21)
    assert a["n_continuous"] == 15 and a["x_raw"].shape == (500, 21)
    assert (a["x_raw"] == b["x_raw"]).all() and (a["w"] == b["w"]).all() and (a["y"] == b["y"]).all()
    assert set(np.unique(a["w"])) <= {0.0, 1.0} and set(np.unique(a["y"])) <= {0.0, 1.0}
    assert set(np.unique(a["x_raw"][:, 15:])) <= {0.0, 1.0}
    f = synth.fold_ids(503, 10, 1)
    c = np.bincount(f, minlength=11)[1:]
    assert c.max() - c.min() <= 1


_WORKER = r"""
import os, sys
sys.path.insert(0, sys.argv[1])
import numpy as np
import torch.distributed as dist
import ctypes as C
from ate_replication_causalml_b200.parallel import TorchComm, replicate_shard, r_sd, install_allreduce
W = int(sys.argv[4])
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%s" % sys.argv[2], rank=int(sys.argv[3]), world_size=W)
comm = TorchComm(dist, None)
# 1. the hook's contract: each rank fills the entries it owns, zeros elsewhere, sum = union
buf = np.zeros(10)
buf[comm.rank::W] = np.arange(10)[comm.rank::W] + 0.5
comm.allreduce_sum_host(buf)
assert np.array_equal(buf, np.arange(10) + 0.5), buf
# 2. replicate shards gathered in replicate order
B = 7
full = np.arange(B) * 1.25
lo, hi = replicate_shard(B, comm.rank, W)
got = comm.allgather_host(full[lo:hi].copy(), [replicate_shard(B, r, W)[1] - replicate_shard(B, r, W)[0] for r in range(W)])
assert np.array_equal(got, full), got
assert abs(r_sd(got) - full.std(ddof=1)) < 1e-14
# 3. the C callback the library calls (ate_allreduce_fn): records of the fits a rank solved, zeros for the
#    others -- the exchange of the full-data fits and of the outcome models (5 fits dealt round-robin)
class FakeLib:
    def ate_ctx_set_allreduce(self, h, cb, user):
        return 0
class FakeEngine:
    _lib = FakeLib(); _h = None
eng = FakeEngine()
install_allreduce(eng, comm)
rec, K = 6, 5
want = np.arange(rec * K, dtype=np.float64).reshape(K, rec) + 1.0
mine = np.zeros_like(want)
for k in range(K):
    if k % W == comm.rank:
        mine[k] = want[k]
rc = eng._allreduce_cb(None, mine.ctypes.data_as(C.POINTER(C.c_double)), mine.size)
assert rc == 0 and np.array_equal(mine, want), mine
# 4. the 128-byte NCCL id travels from rank 0 to every rank unchanged (parallel.init_nccl)
ident = bytes((7 * i + 3) % 256 for i in range(128))
got_id = comm.broadcast_bytes(ident if comm.rank == 0 else None, 128)
assert got_id == ident, got_id
dist.destroy_process_group()
print("ok", comm.rank)
"""


@pytest.mark.parametrize("world", [2, 4])
def test_gloo_exchange_between_ranks(tmp_path, world):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    port = str(29500 + (os.getpid() + 7 * world) % 2000)
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, port, str(r), str(world)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(world)]
    outs = [p.communicate(timeout=180)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o


def _check_round_plan(lib, widths, staged):
    """Every 8x8 tile of every fit's lower triangle is owned by exactly one MMA warp of one CTA kind; a CTA
    kind serves at most four fits and each fit is evaluated (led) by exactly one kind."""
    widths = np.asarray(widths, dtype=np.int32)
    desc = (C.c_uint * (24 * 128))()
    ng = lib.ate_round_plan_describe(len(widths), widths.ctypes.data_as(C.c_void_p), staged, 128, desc)
    assert ng >= 1, (widths, ng)
    owned = {}
    leads = {}
    for g in range(ng):
        d = desc[24 * g: 24 * g + 24]
        nfit = d[0]
        assert 1 <= nfit <= 4
        fits = [d[1 + s] for s in range(nfit)]
        for s in range(nfit):
            if (d[5] >> s) & 1:
                leads[fits[s]] = leads.get(fits[s], 0) + 1
        for w in range(8):
            tiles = 0
            for word in (d[8 + 2 * w], d[9 + 2 * w]):          # two halves per MMA warp
                slot, seg, cnt = word & 3, (word >> 2) & 63, (word >> 8) & 3
                if cnt == 0:
                    continue
                assert slot < nfit and 1 <= cnt <= 2
                f = fits[slot]
                ntr = (int(widths[f]) + 7) // 8
                for u in range(cnt):
                    r = (word >> (10 + 8 * u)) & 255
                    assert 4 * seg <= r < ntr
                    for n in range(min(4, r - 4 * seg + 1)):
                        key = (f, r, 4 * seg + n)
                        assert key not in owned, (widths, key)
                        owned[key] = (g, w)
                        tiles += 1
            assert tiles <= 16
    want = {(f, r, c) for f in range(len(widths)) for r in range((int(widths[f]) + 7) // 8) for c in range(r + 1)}
    assert set(owned) == want, widths
    assert all(leads.get(f, 0) == 1 for f in range(len(widths))), (widths, leads)
    return ng


def test_step_kernel_work_plan_covers_every_tile_once():
    """The MMA warps of the IRLS step kernel compute the lower triangle of each fit's list Gram by a
    host-built plan (csrc/irls_round.cu); checked here without a device for single fits of every
    width, for the ten cv-fold fits of a problem, and for random mixes."""
    from ate_replication_causalml_b200 import _lib
    lib = _lib.load()
    for cols in list(range(0, 130)) + [160, 200, 255, 256, 300, 412, 513, 1024]:
        _check_round_plan(lib, [cols], 264)
    assert _check_round_plan(lib, [64] * 10, 264) == 5           # 36 tiles in six halves per fit: two fits fill a CTA's MMA warps
    assert _check_round_plan(lib, [32] * 10, 264) == 3           # 10 tiles per fit: four fits per CTA
    assert _check_round_plan(lib, [40] * 10, 264) == 3           # lists up to 48 columns still share a CTA four ways
    assert _check_round_plan(lib, [64] * 10, 520) == 10          # wide rows (C4): one fit per CTA
    assert _check_round_plan(lib, [200], 264) >= 2                # a list too wide for one CTA is split
    rng = np.random.default_rng(3)
    for _ in range(60):
        k = int(rng.integers(1, 12))
        _check_round_plan(lib, rng.integers(0, 180, k), 264)
    desc = (C.c_uint * 24)()
    bad = np.array([3000], dtype=np.int32)
    assert lib.ate_round_plan_describe(1, bad.ctypes.data_as(C.c_void_p), 264, 1, desc) == -1


def test_cv_select_is_cv_glmnets_summary():
    """ate_cv_select = cvstats + getmin of cv.glmnet (grouped = TRUE): weighted fold mean, sd =
    sqrt(weighted var / (folds - 1)), lambda.min = LARGEST lambda at the minimum, lambda.1se = largest
    lambda within one sd of it; folds whose path stopped early leave the later columns."""
    from ate_replication_causalml_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(8)
    F, L = 10, 37
    lam = np.exp(np.linspace(0, -6, L))
    base = (np.linspace(-1, 1, L) ** 2)[None, :] + 1.0
    cvraw = np.ascontiguousarray(base + 0.05 * rng.standard_normal((F, L)))
    wts = rng.integers(80, 120, F).astype(np.float64)
    nl = np.full(F, L, np.int32); nl[3] = 30; nl[7] = 35
    cvm = np.empty(L); cvsd = np.empty(L); imin = C.c_int(); i1 = C.c_int()
    ptr = _lib.ptr
    assert lib.ate_cv_select(F, L, ptr(cvraw), ptr(wts), ptr(nl), ptr(lam), ptr(cvm), ptr(cvsd), C.byref(imin), C.byref(i1)) == 0
    for l in range(L):
        use = nl > l
        m = np.average(cvraw[use, l], weights=wts[use])
        v = np.average((cvraw[use, l] - m) ** 2, weights=wts[use]) / (use.sum() - 1)
        assert abs(cvm[l] - m) < 1e-14 and abs(cvsd[l] - np.sqrt(v)) < 1e-14
    assert imin.value == int(np.argmin(cvm))
    assert i1.value == int(np.nonzero(cvm <= cvm[imin.value] + cvsd[imin.value])[0][0])    # lambda is decreasing
    assert i1.value <= imin.value
    # ties go to the larger lambda
    flat = np.ones((F, 4)); lam4 = np.array([4.0, 3.0, 2.0, 1.0]); c4 = np.empty(4); s4 = np.empty(4)
    assert lib.ate_cv_select(F, 4, ptr(flat), ptr(wts), None, ptr(lam4), ptr(c4), ptr(s4), C.byref(imin), C.byref(i1)) == 0
    assert (imin.value, i1.value) == (0, 0) and np.allclose(c4, 1.0) and np.allclose(s4, 0.0)
    assert lib.ate_cv_select(1, 4, ptr(flat), ptr(wts), None, ptr(lam4), ptr(c4), ptr(s4), C.byref(imin), C.byref(i1)) == -1


def test_docs_name_kernels_and_entry_points_that_exist():
    """DESIGN.md / INTEGRATION.md must not drift from the code: every `*_kernel` they name is defined in
    csrc/, every `ate_*` entry point they name is declared in the header, and every declared entry
    point is either tied to reference lines or marked as having no counterpart."""
    csrc = os.path.join(ROOT, "ate_replication_causalml_b200", "csrc")
    code = "".join(open(os.path.join(csrc, f)).read() for f in os.listdir(csrc) if f.endswith((".cu", ".cuh")))
    kernels = set(re.findall(r"\b(\w+_kernel)\b\s*(?:<[^>]*>)?\s*\(", code))
    docs = open(os.path.join(ROOT, "DESIGN.md")).read() + open(os.path.join(ROOT, "INTEGRATION.md")).read()
    for k in set(re.findall(r"`(\w+_kernel)`", docs)):
        assert k in kernels, f"{k} is named in the docs but not defined in csrc/"
    declared = set(_declared_symbols())
    skip = {"ate_b200", "ate_ctx", "ate_crossfit_opts", "ate_crossfit_result", "ate_allreduce_fn", "ate_functions",
            "ate_replication", "ate_oracle", "ate_host_cv_summary", "ate_elnet_path", "ate_lognet_path"}
    for name in set(re.findall(r"`(ate_[a-z0-9_]+)`", docs)):
        if name in skip or name.startswith(("ate_i_", "ate_replication_", "ate_functions")):
            continue
        if name in ("ate_condmean_lasso", "ate_lasso", "ate_condmean_ols"):      # R closures, not C entry points
            continue
        assert name in declared, f"{name} is named in the docs but not declared in include/ate_b200.h"
    hdr = open(os.path.join(ROOT, "include", "ate_b200.h")).read()
    assert re.search(r"ate_functions\.R:\d+", hdr) and re.search(r"ate_replication\.Rmd:\d+", hdr)
    assert hdr.count("counterpart") >= 3          # ctx / hooks / static score / introspection say so explicitly
