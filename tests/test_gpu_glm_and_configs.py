"""GPU parity tests of the unpenalised logistic glm (row f2: doubly_robust_glm exactly as the
reference writes it) and of BASELINE.json's other configurations at sizes the oracle can check
(C0) or through size-independent properties (C2)."""
import numpy as np
import pytest

from ate_replication_causalml_b200 import synth
from conftest import make_folds

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,p", [(50, 1), (500, 3), (5000, 8), (50_000, 22)])
def test_glm_binomial_matches_oracle(engine, oracle, n, p):
    d = synth.make_dataset(n, max(p, 3))
    x = d["x_raw"][:, :p]
    for y in (d["w"], d["y"]):
        g = engine.glm_binomial(x, y)
        o = oracle.glm_binomial(x, y)
        assert g["converged"] and o["converged"] and g["iters"] == o["iters"]
        assert np.abs(g["coef"] - o["coef"]).max() <= 1e-9 * max(1.0, np.abs(o["coef"]).max())
        assert abs(g["deviance"] - o["deviance"]) <= 1e-10 * o["deviance"]
    # the MLE's score equations hold: X'(y - mu) = 0
    xa = np.column_stack([np.ones(n), x])
    mu = 1 / (1 + np.exp(-xa @ g["coef"]))
    assert np.abs(xa.T @ (y - mu)).max() < 1e-6 * n


def test_glm_binomial_intercept_only_and_collinear(engine):
    from ate_replication_causalml_b200._lib import AteError
    y = (np.arange(200) % 3 == 0).astype(float)
    g = engine.glm_binomial(np.zeros((200, 0)), y)
    assert abs(g["coef"][0] - np.log(y.mean() / (1 - y.mean()))) < 1e-10
    x = np.random.default_rng(0).standard_normal((200, 2))
    with pytest.raises(AteError):
        engine.glm_binomial(np.column_stack([x, x[:, 0]]), y)     # exactly collinear -> reported, not NaN


def test_doubly_robust_glm_as_written(engine, oracle):
    """ate_functions.R:211-264 end to end against the oracle's pieces."""
    import pandas as pd
    from ate_replication_causalml_b200 import ate_functions as af
    n, p0 = 8000, 21
    d = synth.make_dataset(n, p0)
    xs, _, _ = oracle.scale(d["x_raw"], cols=np.arange(d["n_continuous"]))
    names = [f"x{j}" for j in range(p0)]
    df = pd.DataFrame(xs, columns=names)
    df["W"] = d["w"]; df["Y"] = d["y"]
    af.set_engine(engine)
    row = af.doubly_robust_glm(df, "W", "Y")
    xo = np.column_stack([xs, d["w"]])
    cm = oracle.glm_binomial(xo, d["y"])["coef"]
    m1 = oracle.predict(np.column_stack([xs, np.ones(n)]), cm[0], cm[1:], True)
    m0 = oracle.predict(np.column_stack([xs, np.zeros(n)]), cm[0], cm[1:], True)
    pc = oracle.glm_binomial(xs, d["w"])["coef"]
    p = oracle.predict(xs, pc[0], pc[1:], True)
    tau, se = oracle.aipw(d["w"], d["y"], p, m0, m1)
    assert row["Method"][0] == "Doubly Robust with logistic regression PS"
    assert abs(row["ATE"][0] - tau) <= 1e-8 * abs(tau)                       # bar: 1e-6
    assert abs(row["upper_ci"][0] - (tau + 1.96 * se)) <= 1e-8 * abs(tau + 1.96 * se)
    # bootstrap branch (B <- 1000 in the reference) with R-drawn indices for a few replicates
    oracle.set_seed(3)
    bidx = np.stack([oracle.sample_replace(n, n, 1) for _ in range(12)])
    rowb = af.doubly_robust_glm(df, "W", "Y", bootstrap_se=True, boot_idx=bidx)
    sd, _ = oracle.bootstrap_idx(d["w"], d["y"], p, m0, m1, bidx)
    assert abs((rowb["upper_ci"][0] - rowb["ATE"][0]) / 1.96 - sd) <= 1e-7 * sd


def test_config_c0_reference_sized_case(engine, oracle):
    """BASELINE.json configs[0]: n = 50k, p = 21 (15 continuous + 6 binary), 5-fold cross-fit AIPW with
    glmnet nuisances, bootstrap SE B = 200 -- the size the reference itself runs on a CPU."""
    n, p0, B = 50_000, 21, 200
    d = synth.make_dataset(n, p0)
    assert d["n_continuous"] == 15
    fo, ips, iout = make_folds(n)
    r = engine.crossfit_aipw(d["x_raw"], d["w"], d["y"], foldid=fo, inner_ps=ips, inner_out=iout,
                             n_continuous=15, expand=0, B=B, boot_mode=1, boot_seed=1991)
    xs, _, _ = oracle.scale(d["x_raw"], cols=np.arange(15))
    o = oracle.crossfit_aipw(xs, d["w"], d["y"], fo, ips, iout)
    assert abs(r["tau"] - o["tau"]) <= 1e-9 * abs(o["tau"]) and abs(r["se_sandwich"] - o["se"]) <= 1e-9 * o["se"]
    assert np.abs(r["phat"] - o["phat"]).max() < 1e-9
    assert 0.7 < r["se_bootstrap"] / r["se_sandwich"] < 1.4 and r["boot"].size == B
    assert abs(r["tau"] - d["true_ate"]) < 5 * r["se_sandwich"]               # and it estimates the right thing


def test_config_c2_bootstrap_many_replicates_properties(engine):
    """BASELINE.json configs[2]: B = 10,000 replicates over n = 1M scores, sharded.  Properties:
    shards reproduce the unsharded replicates bit for bit; the replicate mean/sd are what the
    central limit theorem says."""
    n, B = 1_000_000, 10_000
    rng = np.random.default_rng(4)
    w = (rng.random(n) < 0.4).astype(float); y = (rng.random(n) < 0.3).astype(float)
    p = rng.uniform(0.05, 0.95, n); m0 = rng.uniform(0.1, 0.5, n); m1 = m0 + rng.uniform(-0.1, 0.3, n)
    tau, se_sand = engine.aipw_score(w, y, p, m0, m1)
    sd, boot = engine.bootstrap(w, y, p, m0, m1, B=B, seed=2024)
    parts = [engine.bootstrap(w, y, p, m0, m1, B=B // 8, seed=2024, rep_offset=r * (B // 8))[1] for r in (0, 3, 7)]
    for r, part in zip((0, 3, 7), parts):
        assert (part == boot[r * (B // 8):(r + 1) * (B // 8)]).all()
    assert abs(boot.mean() - tau) < 5 * sd / np.sqrt(B)
    assert abs(sd / se_sand - 1) < 0.05
