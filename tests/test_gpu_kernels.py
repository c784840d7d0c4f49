"""GPU parity tests (through the C ABI) of the HBM-bound stages and the DMMA Gram:
a1 standardise, a2 expansion, a7/a8 score, a9/a10 bootstrap, f1 RNG, weighted Gram."""
import numpy as np
import pytest

from ate_replication_causalml_b200 import synth

pytestmark = pytest.mark.gpu


def _nuisances(n, seed=3):
    rng = np.random.default_rng(seed)
    w = (rng.random(n) < 0.4).astype(float)
    y = (rng.random(n) < 0.3).astype(float)
    p = rng.uniform(0.03, 0.97, n)
    m0 = rng.uniform(0.05, 0.6, n)
    m1 = np.clip(m0 + rng.uniform(-0.1, 0.3, n), 0.01, 0.99)
    return w, y, p, m0, m1


# ---------------------------------------------------------------- a1 / a2
@pytest.mark.parametrize("n,p0", [(2, 1), (37, 3), (5001, 21), (100_000, 16)])
def test_standardize_matches_scale(engine, oracle, n, p0):
    d = synth.make_dataset(max(n, 12), max(p0, 3))
    x = d["x_raw"][:n, :p0]
    cols = np.arange(0, p0, 2)
    got, cen, scl = engine.standardize(x, cols)
    ref, rc, rs = oracle.scale(x, cols)
    if n == 2:  # sd of two points is still defined
        assert np.isfinite(scl).all()
    assert np.allclose(cen, rc, rtol=1e-13, atol=1e-13) and np.allclose(scl, rs, rtol=1e-12)
    assert np.abs(got - ref).max() < 1e-11
    untouched = np.setdiff1d(np.arange(p0), cols)
    assert (got[:, untouched] == x[:, untouched]).all()


@pytest.mark.parametrize("n,p0,main", [(1, 1, True), (129, 5, True), (4099, 21, True), (50_000, 16, False)])
def test_expand_deg2_is_bit_exact(engine, oracle, n, p0, main):
    x = synth.make_dataset(max(n, 12), max(p0, 3))["x_raw"][:n, :p0]
    got = engine.expand_deg2(x, main)
    ref = oracle.expand_deg2(x, main)
    assert got.shape == ref.shape and (got == ref).all()


# ---------------------------------------------------------------- a7 / a8
@pytest.mark.parametrize("n", [1, 7, 1000, 250_001])
def test_aipw_score_matches_oracle(engine, oracle, n):
    w, y, p, m0, m1 = _nuisances(n)
    tau, se, e1, e2 = engine.aipw_score(w, y, p, m0, m1, want_scores=True)
    t, s, r1, r2 = oracle.aipw(w, y, p, m0, m1, want_scores=True)
    assert (e1 == r1).all() and (e2 == r2).all()          # elementwise scores: bit-exact
    assert abs(tau - t) <= 1e-12 * max(1.0, abs(t))
    assert abs(se - s) <= 1e-10 * s + 1e-14               # bar: 1e-6 relative (n = 1: exact 0 vs rounding)


def test_aipw_hand_table_and_nan_rule(engine):
    w = np.array([1, 1, 0, 0, 1, 0.0])
    y = np.array([1, 0, 0, 1, 1, 0.0])
    p = np.array([0.5, 0.25, 0.5, 0.2, 0.8, 0.4])
    m0 = np.array([0.2, 0.1, 0.3, 0.4, 0.5, 0.1])
    m1 = np.array([0.6, 0.3, 0.5, 0.6, 0.9, 0.2])
    est1 = w * (y - m1) / p + (1 - w) * (y - m0) / (1 - p)
    tau_ref = est1.mean() + (m1 - m0).mean()
    tau, _ = engine.aipw_score(w, y, p, m0, m1)
    assert abs(tau - tau_ref) < 1e-15
    # 0/0 -> NaN is dropped from the first mean only (mean(est1, na.rm = TRUE) + mean(est2))
    p2 = p.copy(); p2[0] = 0.0
    m1b = m1.copy(); m1b[0] = 1.0  # y - m1 = 0 -> 0/0
    tau2, _, e1, _ = engine.aipw_score(w, y, p2, m0, m1b, want_scores=True)
    assert np.isnan(e1[0])
    assert abs(tau2 - (np.nanmean(e1) + (m1b - m0).mean())) < 1e-15
    # Inf is kept, as in R
    p3 = p.copy(); p3[1] = 0.0    # y - m1 != 0 -> -Inf
    tau3, _ = engine.aipw_score(w, y, p3, m0, m1)
    assert np.isinf(tau3)


# ---------------------------------------------------------------- a9 / a10
@pytest.mark.parametrize("n,B", [(1, 3), (50, 7), (10_007, 33)])
def test_bootstrap_index_mode_matches_oracle(engine, oracle, n, B):
    w, y, p, m0, m1 = _nuisances(n, 5)
    rng = np.random.default_rng(9)
    idx = rng.integers(1, n + 1, (B, n)).astype(np.int32)
    sd, boot = engine.bootstrap(w, y, p, m0, m1, idx=idx)
    osd, oboot = oracle.bootstrap_idx(w, y, p, m0, m1, idx)
    assert np.abs(boot - oboot).max() <= 1e-12 * max(1.0, np.abs(oboot).max())
    assert abs(sd - osd) <= 1e-9 * osd or (np.isnan(sd) and np.isnan(osd))
    assert abs(engine.tau_hat_dr_est(w, y, p, m0, m1, idx[0]) - oracle.tau_hat_dr_est(w, y, p, m0, m1, idx[0])) < 1e-12


def test_bootstrap_with_nan_scores_uses_two_tables(engine, oracle):
    n, B = 400, 9
    w, y, p, m0, m1 = _nuisances(n, 6)
    p[3] = 0.0; w[3] = 1.0; y[3] = 1.0; m1[3] = 1.0     # 0/0
    idx = np.random.default_rng(2).integers(1, n + 1, (B, n)).astype(np.int32)
    idx[0, :5] = 4                                      # make sure the NaN row is drawn
    sd, boot = engine.bootstrap(w, y, p, m0, m1, idx=idx)
    osd, oboot = oracle.bootstrap_idx(w, y, p, m0, m1, idx)
    assert np.isfinite(boot).all() and np.abs(boot - oboot).max() < 1e-12


def _philox_indices(n, B, seed, rep0=0):
    """numpy twin of the kernel's index stream: Philox4x32-10, counter (pair, 0, rep), key seed;
    idx = mulhi64(r64, n)."""
    M0, M1, W0, W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
    npairs = (n + 1) // 2
    out = np.empty((B, n), dtype=np.int64)
    for b in range(B):
        rep = rep0 + b
        c0 = np.arange(npairs, dtype=np.uint64) & np.uint64(0xFFFFFFFF)
        c1 = np.arange(npairs, dtype=np.uint64) >> np.uint64(32)
        c2 = np.full(npairs, rep & 0xFFFFFFFF, dtype=np.uint64)
        c3 = np.full(npairs, rep >> 32, dtype=np.uint64)
        k0, k1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
        for _ in range(10):
            p0_ = np.uint64(M0) * c0
            p1_ = np.uint64(M1) * c2
            hi0, lo0 = p0_ >> np.uint64(32), p0_ & np.uint64(0xFFFFFFFF)
            hi1, lo1 = p1_ >> np.uint64(32), p1_ & np.uint64(0xFFFFFFFF)
            c0, c1, c2, c3 = hi1 ^ c1 ^ np.uint64(k0), lo1, hi0 ^ c3 ^ np.uint64(k1), lo0
            k0 = (k0 + W0) & 0xFFFFFFFF
            k1 = (k1 + W1) & 0xFFFFFFFF
        ra = [int(v) for v in ((c1 << np.uint64(32)) | c0)]
        rb = [int(v) for v in ((c3 << np.uint64(32)) | c2)]
        ia = np.array([(r * n) >> 64 for r in ra], dtype=np.int64)
        ib = np.array([(r * n) >> 64 for r in rb], dtype=np.int64)
        row = np.empty(2 * npairs, dtype=np.int64)
        row[0::2], row[1::2] = ia, ib
        out[b] = row[:n]
    return out + 1


def test_bootstrap_philox_stream_is_the_documented_one(engine, oracle):
    n, B, seed = 1001, 6, 0x1234_5678_9ABC
    w, y, p, m0, m1 = _nuisances(n, 8)
    sd, boot = engine.bootstrap(w, y, p, m0, m1, B=B, seed=seed)
    idx = _philox_indices(n, B, seed).astype(np.int32)
    assert idx.min() >= 1 and idx.max() <= n
    _, oboot = oracle.bootstrap_idx(w, y, p, m0, m1, idx)
    assert np.abs(boot - oboot).max() < 1e-12


def test_bootstrap_philox_sharding_invariance(engine):
    n, B = 5000, 40
    w, y, p, m0, m1 = _nuisances(n, 10)
    _, full = engine.bootstrap(w, y, p, m0, m1, B=B, seed=77)
    _, a = engine.bootstrap(w, y, p, m0, m1, B=15, seed=77, rep_offset=0)
    _, b = engine.bootstrap(w, y, p, m0, m1, B=25, seed=77, rep_offset=15)
    assert (np.r_[a, b] == full).all()                   # bit-identical however it is sharded
    _, other = engine.bootstrap(w, y, p, m0, m1, B=B, seed=78)
    assert not (other == full).any()


def test_bootstrap_se_is_statistically_sane(engine):
    n = 200_000
    w, y, p, m0, m1 = _nuisances(n, 12)
    tau, se_sand = engine.aipw_score(w, y, p, m0, m1)
    sd, boot = engine.bootstrap(w, y, p, m0, m1, B=400, seed=1)
    assert abs(boot.mean() - tau) < 4 * sd / np.sqrt(400) + 1e-4
    assert 0.8 < sd / se_sand < 1.25                      # bootstrap and sandwich agree roughly


# ---------------------------------------------------------------- f1: R's sample()
def test_r_rng_bit_exact_with_oracle_and_known_answers(engine, oracle, golden):
    k = golden["rng"]
    engine.r_set_seed(1)
    perm = engine.r_foldid(10, 10, 1)                    # nfolds = n: fold ids are the permutation
    assert perm.tolist() == k["set.seed(1); sample(10) [R >= 3.6, Rejection]"]
    engine.r_set_seed(1)
    assert engine.r_foldid(10, 10, 0).tolist() == k["set.seed(1); sample(10) [R <= 3.5, Rounding]"]
    for kind in (0, 1):
        engine.r_set_seed(1991); oracle.set_seed(1991)
        assert (engine.r_sample(1_000_000, 20_000, kind) == oracle.sample_replace(1_000_000, 20_000, kind)).all()
        assert (engine.r_foldid(12_345, 10, kind) == oracle.foldid(12_345, 10, kind)).all()


# ---------------------------------------------------------------- weighted Gram
@pytest.mark.parametrize("n,p", [(1, 1), (31, 5), (1000, 40), (4097, 256), (3000, 70), (513, 288),
                                 (2500, 417), (3000, 515), (1500, 1027)])   # the last three: panel-pair path
def test_weighted_gram_matches_numpy(engine, n, p):
    rng = np.random.default_rng(n + p)
    x = rng.standard_normal((n, p))
    rows = np.sort(rng.choice(n, max(1, n * 3 // 4), replace=False)).astype(np.int32)
    v = rng.uniform(0.1, 1.0, rows.size)
    G = engine.weighted_gram(x, v, rows)
    ref = (x[rows] * v[:, None]).T @ x[rows]
    assert np.abs(G - ref).max() <= 1e-12 * np.abs(ref).max()
    assert (G == G.T).all()
    G1 = engine.weighted_gram(x)
    assert np.abs(G1 - x.T @ x).max() <= 1e-12 * max(1.0, np.abs(x.T @ x).max())


def test_weighted_gram_is_reproducible_and_additive_over_row_shards(engine):
    rng = np.random.default_rng(0)
    n, p = 20_000, 96
    x = rng.standard_normal((n, p))
    v = rng.uniform(0.0, 1.0, n)
    a = engine.weighted_gram(x, v)
    b = engine.weighted_gram(x, v)
    assert (a == b).all()                                 # fixed fold order: bit-reproducible
    r1 = np.arange(0, n // 3, dtype=np.int32)
    r2 = np.arange(n // 3, n, dtype=np.int32)
    s = engine.weighted_gram(x, v[r1], r1) + engine.weighted_gram(x, v[r2], r2)
    assert np.abs(s - a).max() <= 1e-12 * np.abs(a).max()  # what a row-sharded all-reduce relies on


def test_weighted_gram_full_size_properties(engine):
    """BASELINE size (n = 1M, p = 256) through size-independent properties, on device memory."""
    import torch
    n, p = 1_000_000, 256
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(1)
    X = torch.randn(n, p, dtype=torch.float64, device=dev, generator=g)
    v = torch.rand(n, dtype=torch.float64, device=dev, generator=g)
    G = torch.empty(p, p, dtype=torch.float64, device=dev)
    engine.weighted_gram_device(n, p, X.data_ptr(), p, v.data_ptr(), G.data_ptr())
    assert torch.equal(G, G.T)
    # linearity in v: G(v) + G(1 - v) = G(1)
    v2 = 1.0 - v
    G2 = torch.empty_like(G); G3 = torch.empty_like(G)
    engine.weighted_gram_device(n, p, X.data_ptr(), p, v2.data_ptr(), G2.data_ptr())
    engine.weighted_gram_device(n, p, X.data_ptr(), p, None, G3.data_ptr())
    assert ((G + G2 - G3).abs().max() / G3.abs().max()).item() < 1e-12
    # trace and a random quadratic form against plain reductions
    tr = (X * X * v[:, None]).sum()
    assert abs((torch.trace(G) - tr) / tr).item() < 1e-12
    z = torch.randn(p, dtype=torch.float64, device=dev, generator=g)
    q = ((X @ z) ** 2 * v).sum()
    assert abs((z @ G @ z - q) / q).item() < 1e-11
