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


# ------------------------------------------------------------------ rows f3 / f4: lm-based estimators
def _lm_close(g, o, tol=1e-9):
    assert g["rank"] == o["rank"] and g["df_residual"] == o["df_residual"]
    assert (np.isnan(g["coef"]) == np.isnan(o["coef"])).all()
    k = ~np.isnan(o["coef"])
    assert np.abs(g["coef"][k] - o["coef"][k]).max() <= tol * max(1.0, np.abs(o["coef"][k]).max())
    assert np.abs(g["std_err"][k] - o["std_err"][k]).max() <= tol * np.abs(o["std_err"][k]).max()
    assert abs(g["rss"] - o["rss"]) <= tol * o["rss"] and abs(g["sigma"] - o["sigma"]) <= tol * o["sigma"]


@pytest.mark.parametrize("n,p,weighted", [(40, 1, False), (3000, 7, True), (50_000, 22, False), (20_000, 120, True)])
def test_lm_matches_oracle_qr(engine, oracle, n, p, weighted):
    rng = np.random.default_rng(n + p)
    x = rng.standard_normal((n, p)) * rng.uniform(0.2, 5, p) + rng.uniform(-2, 2, p)
    y = x @ rng.standard_normal(p) + 1.5 + rng.standard_normal(n)
    w = rng.uniform(0.2, 3, n) if weighted else None
    if weighted:
        w[::17] = 0.0                      # lm.wfit drops zero weights from the degrees of freedom
    g = engine.lm(x, y, w, want_resid=True)
    o = oracle.lm(x, y, w, want_resid=True)
    _lm_close(g, o)
    assert np.abs(g["resid"] - o["resid"]).max() <= 1e-9 * np.abs(y).max()


def test_lm_aliased_columns_are_na_like_r(engine, oracle):
    from ate_replication_causalml_b200._lib import AteError
    rng = np.random.default_rng(9)
    n = 2000
    x = rng.standard_normal((n, 5))
    xa = np.column_stack([x[:, :3], x[:, 1], x[:, 3:], np.zeros(n), x[:, 0] - 2 * x[:, 4]])
    y = x @ np.arange(1.0, 6.0) + rng.standard_normal(n)
    g, o = engine.lm(xa, y), oracle.lm(xa, y)
    assert np.isnan(g["coef"][[4, 7, 8]]).all() and g["rank"] == 6
    _lm_close(g, o)
    with pytest.raises(AteError):
        engine.lm(x, y, -np.ones(n))


def _frame(d, xs, p0):
    import pandas as pd
    names = [f"x{j}" for j in range(p0)]
    df = pd.DataFrame(np.asarray(xs), columns=names)
    df["W"] = d["w"]; df["Y"] = d["y"]
    return df, names


def test_regression_estimators_as_written(engine, oracle):
    """ate_condmean_ols (ate_functions.R:24-39), prop_score_weight (:44-64), prop_score_ols (:67-86)
    with a glm propensity as ate_replication.Rmd:165-176 feeds them."""
    from ate_replication_causalml_b200 import ate_functions as af
    n, p0 = 20_000, 21
    d = synth.make_dataset(n, p0)
    xs, _, _ = oracle.scale(d["x_raw"], cols=np.arange(d["n_continuous"]))
    df, names = _frame(d, xs, p0)
    af.set_engine(engine)
    w, y = d["w"], d["y"]

    row = af.ate_condmean_ols(df, "W", "Y")
    o = oracle.lm(np.column_stack([xs, w]), y)
    assert row["Method"][0] == "Direct Method"
    assert abs(row["ATE"][0] - o["coef"][-1]) <= 1e-9 * abs(o["coef"][-1])
    assert abs(row["upper_ci"][0] - (o["coef"][-1] + 1.96 * o["std_err"][-1])) <= 1e-9

    pc = oracle.glm_binomial(xs, w)["coef"]
    p = oracle.predict(xs, pc[0], pc[1:], True)

    row = af.prop_score_weight(df, p, "W", "Y", names)
    tau_i = ((w - p) * y) / (p * (1 - p))
    e = oracle.lm(np.asarray(xs) * (w - p)[:, None], tau_i, want_resid=True)["resid"]
    tau, se = oracle.r_mean(tau_i), np.sqrt(oracle.r_mean(e * e)) / np.sqrt(n)
    assert row["Method"][0] == "Propensity_Weighting"
    assert abs(row["ATE"][0] - tau) <= 1e-12 * abs(tau)
    assert abs(row["lower_ci"][0] - (tau - 1.96 * se)) <= 1e-9 * abs(tau)

    row = af.prop_score_ols(df, p, "W", "Y")
    o = oracle.lm(w.reshape(-1, 1), y, w / p + (1 - w) / (1 - p))
    assert row["Method"][0] == "Propensity_Regression"
    assert abs(row["ATE"][0] - o["coef"][1]) <= 1e-9 * abs(o["coef"][1])
    assert abs(row["upper_ci"][0] - (o["coef"][1] + 1.96 * o["std_err"][1])) <= 1e-9
    # the Horvitz-Thompson-weighted slope is the difference of the two weighted means
    wt = w / p + (1 - w) / (1 - p)
    diff = (wt * w * y).sum() / (wt * w).sum() - (wt * (1 - w) * y).sum() / (wt * (1 - w)).sum()
    assert abs(row["ATE"][0] - diff) <= 1e-9


def test_belloni_end_to_end(engine, oracle):
    """belloni() (ate_functions.R:286-328): expansion, two cv.glmnet fits, support at model_xw's
    lambda.min with the `> 0` / `- 1` quirks, post-lasso OLS; each stage against the oracle."""
    from ate_replication_causalml_b200 import ate_functions as af
    n, p0 = 6000, 6
    d = synth.make_dataset(n, p0)
    xs, _, _ = oracle.scale(d["x_raw"], cols=np.arange(d["n_continuous"]))
    df, names = _frame(d, xs, p0)
    af.set_engine(engine); af.set_covariates(names)
    oracle.set_seed(7); f1 = oracle.foldid(n, 10)
    f2 = oracle.foldid(n, 10)
    row = af.belloni(df, "W", "Y", foldid_xw=f1, foldid_xy=f2)
    x = oracle.expand_deg2(xs, include_main=True)
    mw = oracle.cv_glmnet(x, d["w"], f1)
    my = oracle.cv_glmnet(x, d["y"], f2)
    s = mw["lambda_min"]
    c_xw = mw["beta"][:, mw["idx_min"]]
    c_xy = af._coef_at(dict(lambda_=my["lambda_"], beta=my["beta"]), s)
    import pandas as pd
    cu = pd.unique(np.r_[np.nonzero(c_xw > 0)[0] + 1, np.nonzero(c_xy > 0)[0] + 1]) - 1
    cu = cu[cu > 0]
    # x_a*x_b and x_b*x_a are identical columns, so WHICH twin carries the weight is not determined
    #  and coordinate descent stops at thresh = 1e-7: the supports agree up to
    # the twin symmetry except for columns that are only just entering at this lambda
    def canon(j):
        a, b = divmod(int(j) - p0, p0)
        return ("main", int(j)) if j < p0 else ("prod", min(a, b), max(a, b))
    got = row.attrs["c_union"]
    big = lambda c: {canon(j) for j in np.nonzero(c > 1e-4)[0]}          # noqa: E731
    assert big(row.attrs["c_xw"]) == big(c_xw) and big(row.attrs["c_xy"]) == big(c_xy)
    for j in set(map(canon, got - 1)) ^ set(map(canon, cu - 1)):
        twins = [k for k in range(x.shape[1]) if canon(k) == j]
        assert max(np.abs(c[twins]).max() for c in (row.attrs["c_xw"], row.attrs["c_xy"], c_xw, c_xy)) < 1e-4
    assert cu.size > 2 and row["Method"][0] == "Belloni et.al"
    # post-lasso OLS: exact on the support the GPU selected, and close on the oracle's own
    o = oracle.lm(np.column_stack([x[:, got - 1], d["w"]]), d["y"])
    assert abs(row["ATE"][0] - o["coef"][-1]) <= 1e-9 * abs(o["coef"][-1])
    assert abs(row["upper_ci"][0] - row["ATE"][0] - 1.96 * o["std_err"][-1]) <= 1e-9 * o["std_err"][-1]
    assert row.attrs["fit"]["rank"] == o["rank"]
    o2 = oracle.lm(np.column_stack([x[:, cu - 1], d["w"]]), d["y"])
    assert abs(row["ATE"][0] - o2["coef"][-1]) <= 0.1 * o2["std_err"][-1]   # one marginal regressor more or less


def test_wide_designs_beyond_one_staged_row(engine, oracle):
    """BASELINE.json configs[3]/[4] are wider (p = 1024 / 512) than a row the staging ring holds; the
    Gram then runs panel by panel.  Gaussian cv.glmnet (C3 is a 10-fold CV path) and lm at such widths
    against the oracle, at an n the oracle finishes quickly."""
    rng = np.random.default_rng(5)
    n, p = 3000, 530
    x = rng.standard_normal((n, p)) * rng.uniform(0.5, 2.0, p)
    beta = np.zeros(p); beta[rng.choice(p, 12, replace=False)] = rng.uniform(-1, 1, 12)
    y = x @ beta + rng.standard_normal(n)
    fid = synth.fold_ids(n, 10, 7)
    g = engine.cv_glmnet(x, y, "gaussian", foldid=fid, nlambda=40)
    o = oracle.cv_glmnet(x, y, fid, "gaussian", nlambda=40)
    assert g["lmu"] == o["lmu"] and g["idx_min"] == o["idx_min"] and g["idx_1se"] == o["idx_1se"]
    assert np.abs(g["lambda_"] - o["lambda_"]).max() <= 1e-12 * o["lambda_"][0]
    assert np.abs(g["beta"] - o["beta"]).max() <= 1e-8 and np.abs(g["cvm"] - o["cvm"]).max() <= 1e-9 * o["cvm"].max()
    gl, ol = engine.lm(x[:, :450], y), oracle.lm(x[:, :450], y)
    _lm_close(gl, ol, tol=1e-8)
    # the binomial step stages such rows in 16-row slots (round 1 refused them)
    yb = (y > 0).astype(float)
    gb = engine.cv_glmnet(x, yb, "binomial", foldid=fid, nlambda=12)
    ob = oracle.cv_glmnet(x, yb, fid, "binomial", nlambda=12)
    assert gb["lmu"] == ob["lmu"] and gb["idx_min"] == ob["idx_min"] and gb["idx_1se"] == ob["idx_1se"]
    assert np.abs(gb["beta"] - ob["beta"]).max() <= 1e-8 and np.allclose(gb["cvm"], ob["cvm"], rtol=1e-9)
