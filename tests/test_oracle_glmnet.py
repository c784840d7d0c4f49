"""Pin the oracle's glmnet restatement: closed forms, scikit-learn's independent solvers (golden
fixtures), KKT conditions, and its own regression fixtures."""
import numpy as np
import pytest

from ate_replication_causalml_b200 import synth


def test_gaussian_path_matches_sklearn_golden(oracle, golden):
    sk = golden["sk"]
    r = oracle.glmnet(sk["X"], sk["y"], "gaussian", thresh=1e-15)
    assert np.allclose(r["lambda_"], sk["lambda_gauss"], rtol=1e-12)
    assert np.abs(r["beta"][:, 1:] - sk["sk_gauss"]).max() < 1e-8


def test_binomial_path_matches_sklearn_golden(oracle, golden):
    sk = golden["sk"]
    r = oracle.glmnet(sk["X"], sk["w"], "binomial", thresh=1e-15)
    assert np.allclose(r["lambda_"], sk["lambda_binom"], rtol=1e-12)
    for i, l in enumerate(sk["sel_binom"]):
        assert abs(r["a0"][l] - sk["sk_binom"][i][0]) < 1e-7
        assert np.abs(r["beta"][:, l] - sk["sk_binom"][i][1:]).max() < 1e-7


def test_default_thresh_is_within_glmnet_tolerance_of_optimum(oracle, golden):
    sk = golden["sk"]
    r = oracle.glmnet(sk["X"], sk["y"], "gaussian")  # thresh 1e-7
    L = min(r["lmu"], sk["sk_gauss"].shape[1] + 1)
    assert np.abs(r["beta"][:, 1:L] - sk["sk_gauss"][:, :L - 1]).max() < 5e-4


def test_orthonormal_design_soft_threshold(oracle):
    rng = np.random.default_rng(0)
    n, p = 64, 8
    q, _ = np.linalg.qr(rng.standard_normal((n, p)) - 0.0)
    q = q - q.mean(0)
    q, _ = np.linalg.qr(q)  # centred, orthonormal columns: sd = 1/sqrt(n) each
    y = q @ np.array([3, -2, 1, 0.5, 0, 0, 0, 0.0]) + 0.1 * rng.standard_normal(n)
    r = oracle.glmnet(q, y, "gaussian", thresh=1e-16)
    xs = q.std(0)
    # on standardised columns beta_std = S(<x_std, y>/n, lambda); beta = beta_std / xs
    z = ((q / xs).T @ (y - y.mean())) / n
    for l in (3, 10, 20):
        lam = r["lambda_"][l]
        expect = np.sign(z) * np.maximum(np.abs(z) - lam, 0) / xs
        assert np.abs(r["beta"][:, l] - expect).max() < 1e-9


def test_lambda_max_gives_null_model_and_path_is_decreasing(oracle):
    d = synth.make_dataset(800, 6)
    r = oracle.glmnet(d["x_raw"], d["y"], "gaussian")
    assert (r["beta"][:, 0] == 0).all()
    assert (np.diff(r["lambda_"]) < 0).all()
    rb = oracle.glmnet(d["x_raw"], d["w"], "binomial")
    assert (rb["beta"][:, 0] == 0).all()
    assert abs(rb["a0"][0] - np.log(d["w"].mean() / (1 - d["w"].mean()))) < 1e-10


def test_binomial_kkt_along_path(oracle):
    d = synth.make_dataset(3000, 8)
    X, w = d["x_raw"], d["w"]
    r = oracle.glmnet(X, w, "binomial", thresh=1e-13)
    Xs = (X - X.mean(0)) / X.std(0)
    for l in range(1, r["lmu"], 7):
        pr = 1 / (1 + np.exp(-(r["a0"][l] + X @ r["beta"][:, l])))
        g = (Xs * (w - pr)[:, None]).mean(0)
        lam = r["lambda_"][l]
        act = r["beta"][:, l] != 0
        assert (np.abs(g[~act]) <= lam * (1 + 1e-6) + 1e-9).all()
        if act.any():
            assert np.abs(np.abs(g[act]) - lam).max() < 1e-6
        assert abs((w - pr).mean()) < 1e-7


def test_penalty_factor_zero_keeps_treatment_in_model(oracle):
    d = synth.make_dataset(1500, 6)
    xw = np.column_stack([d["x_raw"], d["w"]])
    vp = np.r_[np.ones(6), 0.0]
    r = oracle.glmnet(xw, d["y"], "gaussian", penalty_factor=vp)
    assert (r["beta"][6, :] != 0).all()         # the unpenalised coefficient is in from lambda_1 on
    assert (r["beta"][:6, 0] == 0).all()


def test_user_lambda_sequence_reproduces_path(oracle):
    d = synth.make_dataset(1200, 6)
    r = oracle.glmnet(d["x_raw"], d["y"], "gaussian", thresh=1e-14)
    r2 = oracle.glmnet(d["x_raw"], d["y"], "gaussian", lambda_=r["lambda_"], thresh=1e-14)
    assert r2["lmu"] == r["lmu"]
    assert np.abs(r2["beta"] - r["beta"]).max() < 1e-8


def test_cv_summary_and_getmin(oracle):
    d = synth.make_dataset(1000, 5)
    fid = synth.fold_ids(1000, 10, 3)
    c = oracle.cv_glmnet(d["x_raw"], d["y"], fid, "gaussian")
    i = c["idx_min"]
    assert c["cvm"][i] == c["cvm"].min()
    assert c["idx_1se"] <= i                                   # larger lambda = earlier index
    assert c["cvm"][c["idx_1se"]] <= c["cvm"][i] + c["cvsd"][i]
    if c["idx_1se"] > 0:
        assert c["cvm"][c["idx_1se"] - 1] > c["cvm"][i] + c["cvsd"][i]


@pytest.mark.parametrize("tag,n,p0", [("a", 200, 5), ("b", 5000, 21)])
def test_oracle_regression_fixtures(oracle, golden, tag, n, p0):
    reg = golden["reg"]
    d = synth.make_dataset(n, p0)
    xs, cen, scl = oracle.scale(d["x_raw"], cols=np.arange(d["n_continuous"]))
    assert np.allclose(cen, reg[f"{tag}_center"], rtol=1e-13) and np.allclose(scl, reg[f"{tag}_scale"], rtol=1e-13)
    xw = np.column_stack([xs, d["w"]])
    g = oracle.glmnet(xw, d["y"], "gaussian", penalty_factor=np.r_[np.ones(p0), 0.0])
    assert g["npasses"] == int(reg[f"{tag}_g_npasses"])
    assert np.abs(g["beta"] - reg[f"{tag}_g_beta"]).max() < 1e-12
    b = oracle.glmnet(xs, d["w"], "binomial")
    assert b["npasses"] == int(reg[f"{tag}_b_npasses"])
    assert np.abs(b["beta"] - reg[f"{tag}_b_beta"]).max() < 1e-11


def test_aipw_hand_computed_table(oracle):
    # 6-row table worked by hand from ate_functions.R:241-243,255-256
    w = np.array([1, 1, 0, 0, 1, 0.0])
    y = np.array([1, 0, 0, 1, 1, 0.0])
    p = np.array([0.5, 0.25, 0.5, 0.2, 0.8, 0.4])
    m0 = np.array([0.2, 0.1, 0.3, 0.4, 0.5, 0.1])
    m1 = np.array([0.6, 0.3, 0.5, 0.6, 0.9, 0.2])
    est1 = w * (y - m1) / p + (1 - w) * (y - m0) / (1 - p)
    est2 = m1 - m0
    tau = est1.mean() + est2.mean()
    Ii = (w * y) / p - m1 * (w - p) / p - (((1 - w) * y / (1 - p)) + (m0 * (w - p) / (1 - p))) - tau
    se = np.sqrt((Ii ** 2).sum()) / 6
    t, s = oracle.aipw(w, y, p, m0, m1)
    assert abs(t - tau) < 1e-15 and abs(s - se) < 1e-15


def test_aipw_nan_only_shrinks_first_mean(oracle):
    w = np.array([1, 0, 1, 0.0])
    y = np.array([1, 0, 0, 1.0])
    p = np.array([0.5, 0.5, 0.0, 0.5])   # p = 0 with w = 1, y - m1 = 0 -> 0/0 = NaN
    m0 = np.array([0.2, 0.2, 0.2, 0.2])
    m1 = np.array([0.4, 0.4, 0.0, 0.4])
    t, _ = oracle.aipw(w, y, p, m0, m1)
    est1 = w * (y - m1) / np.where(p == 0, np.nan, p) + (1 - w) * (y - m0) / (1 - p)
    assert np.isnan(est1[2])
    expect = np.nanmean(est1) + (m1 - m0).mean()
    assert abs(t - expect) < 1e-15


def test_bootstrap_fixed_index_matrix(oracle):
    rng = np.random.default_rng(5)
    n, B = 50, 7
    w = (rng.random(n) < 0.5).astype(float)
    y = (rng.random(n) < 0.5).astype(float)
    p = rng.uniform(0.2, 0.8, n)
    m0 = rng.uniform(0.1, 0.5, n)
    m1 = rng.uniform(0.3, 0.8, n)
    idx = rng.integers(1, n + 1, (B, n)).astype(np.int32)
    sd, boot = oracle.bootstrap_idx(w, y, p, m0, m1, idx)
    est1 = w * (y - m1) / p + (1 - w) * (y - m0) / (1 - p)
    est2 = m1 - m0
    ref = np.array([est1[i - 1].mean() + est2[i - 1].mean() for i in idx])
    assert np.abs(boot - ref).max() < 1e-14
    assert abs(sd - ref.std(ddof=1)) < 1e-14
    assert abs(oracle.tau_hat_dr_est(w, y, p, m0, m1, idx[0]) - ref[0]) < 1e-14


def test_expand_order_matches_belloni_loop(oracle):
    x = np.arange(1, 13, dtype=float).reshape(4, 3)
    out = oracle.expand_deg2(x, True)
    cols = [x[:, j] for j in range(3)]
    for c1 in range(3):
        for c2 in range(3):
            cols.append(x[:, c1] * x[:, c2])
    assert (out == np.column_stack(cols)).all()
    assert (oracle.expand_deg2(x, False) == np.column_stack(cols[3:])).all()


def test_glm_binomial_is_the_mle_and_matches_glm_fit_conventions(oracle):
    """stats::glm.fit (ate_functions.R:218-220, 231-234): the fit is the unpenalised MLE (score = 0),
    stops by |dev - devold| / (|dev| + 0.1) < 1e-8, and agrees with the L1 path's lambda -> 0 end."""
    rng = np.random.default_rng(11)
    n, p = 4000, 6
    x = rng.standard_normal((n, p))
    y = (rng.random(n) < 1 / (1 + np.exp(-(0.3 + x @ np.array([0.8, -0.5, 0, 0.2, 0, 1.0]))))).astype(float)
    g = oracle.glm_binomial(x, y)
    assert g["converged"] and 3 <= g["iters"] <= 8
    xa = np.column_stack([np.ones(n), x])
    mu = 1 / (1 + np.exp(-xa @ g["coef"]))
    assert np.abs(xa.T @ (y - mu)).max() < 1e-6
    dev = -2 * np.sum(y * np.log(mu) + (1 - y) * np.log1p(-mu))
    assert abs(dev - g["deviance"]) < 1e-9 * dev
    path = oracle.glmnet(x, y, "binomial", lambda_=np.array([1e-1, 1e-2, 1e-3, 1e-4, 1e-6, 1e-8]), thresh=1e-12)
    assert np.abs(path["beta"][:, -1] - g["coef"][1:]).max() < 1e-4
    # maxit = 1 reports non-convergence, like R's warning
    assert not oracle.glm_binomial(x, y, maxit=1)["converged"]


def test_lm_qr_matches_normal_equations_and_summary_formulas(oracle):
    """stats::lm + summary.lm as the regression estimators use them (ate_functions.R:28, 53, 74, 320):
    Estimate, Std. Error = sqrt(rss/rdf * diag((X'WX)^-1)), zero weights leave the df, aliased columns
    come back NA in dqrdc2's order (later duplicates are the ones dropped)."""
    rng = np.random.default_rng(0)
    n, p = 500, 6
    x = rng.standard_normal((n, p))
    y = x @ rng.standard_normal(p) + 0.5 + rng.standard_normal(n)
    w = rng.uniform(0.5, 2, n); w[:7] = 0
    for ww in (None, w):
        r = oracle.lm(x, y, ww, want_resid=True)
        xa = np.column_stack([np.ones(n), x]); W = np.ones(n) if ww is None else ww
        xtx = xa.T @ (W[:, None] * xa)
        b = np.linalg.solve(xtx, xa.T @ (W * y))
        e = y - xa @ b
        rss, df = (W * e * e).sum(), int((W != 0).sum()) - p - 1
        assert np.abs(r["coef"] - b).max() < 1e-12 and r["df_residual"] == df and r["rank"] == p + 1
        assert np.abs(r["std_err"] - np.sqrt(rss / df * np.diag(np.linalg.inv(xtx)))).max() < 1e-12
        assert np.abs(r["resid"] - e).max() < 1e-11
    x2 = np.column_stack([x[:, :3], x[:, 0], x[:, 3:], np.zeros(n)])
    r2, r0 = oracle.lm(x2, y), oracle.lm(x, y)
    assert np.isnan(r2["coef"][[4, 8]]).all() and r2["rank"] == p + 1
    assert np.abs(r2["coef"][[0, 1, 2, 3, 5, 6, 7]] - r0["coef"]).max() < 1e-12


# ---------------------------------------------------------------- properties
@pytest.mark.parametrize("family", ["gaussian", "binomial"])
def test_row_permutation_leaves_the_path_unchanged(oracle, family):
    rng = np.random.default_rng(3)
    n, p = 800, 9
    x = rng.standard_normal((n, p)); x[:, 2] = (x[:, 2] > 0.3)
    eta = x @ rng.uniform(-1, 1, p)
    y = eta + rng.standard_normal(n) if family == "gaussian" else (rng.random(n) < 1 / (1 + np.exp(-eta))).astype(float)
    perm = rng.permutation(n)
    a, b = oracle.glmnet(x, y, family), oracle.glmnet(x[perm], y[perm], family)
    assert a["lmu"] == b["lmu"]
    assert np.abs(a["lambda_"] - b["lambda_"]).max() <= 1e-12 * a["lambda_"][0]
    assert np.abs(a["beta"] - b["beta"]).max() <= 1e-9          # sums in another order: rounding only
    fid = (np.arange(n) % 10 + 1).astype(np.int32)
    ca, cb = oracle.cv_glmnet(x, y, fid, family), oracle.cv_glmnet(x[perm], y[perm], fid[perm], family)
    assert (ca["idx_min"], ca["idx_1se"]) == (cb["idx_min"], cb["idx_1se"])
    assert np.abs(ca["cvm"] - cb["cvm"]).max() <= 1e-9 * ca["cvm"].max()
    relabel = (10 - fid + 1).astype(np.int32)                     # fold labels are names, not an order
    cc = oracle.cv_glmnet(x, y, relabel, family)
    assert np.abs(ca["cvm"] - cc["cvm"]).max() <= 1e-12 * ca["cvm"].max()
    assert np.abs(ca["cvsd"] - cc["cvsd"]).max() <= 1e-10 * ca["cvsd"].max()


def test_aipw_and_bootstrap_identities_hypothesis(oracle):
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=40, deadline=None)
    @given(st.integers(5, 200), st.integers(0, 2**31 - 1))
    def prop(n, seed):
        rng = np.random.default_rng(seed)
        w = (rng.random(n) < 0.5).astype(float); y = (rng.random(n) < 0.4).astype(float)
        p = rng.uniform(0.05, 0.95, n); m0 = rng.uniform(0, 1, n); m1 = rng.uniform(0, 1, n)
        tau, se, e1, e2 = oracle.aipw(w, y, p, m0, m1, want_scores=True)
        # the definition, ate_functions.R:241-243 and :255-256
        est1 = w * (y - m1) / p + (1 - w) * (y - m0) / (1 - p)
        assert abs(tau - (est1.mean() + (m1 - m0).mean())) <= 1e-12 * max(1.0, abs(tau))
        ii = w * y / p - m1 * (w - p) / p - ((1 - w) * y / (1 - p) + m0 * (w - p) / (1 - p)) - tau
        assert abs(se - np.sqrt((ii ** 2).sum()) / n) <= 1e-12 * se
        # a replicate over the identity index is the point estimate; over a permutation too
        ident = np.arange(1, n + 1, dtype=np.int32)
        assert abs(oracle.tau_hat_dr_est(w, y, p, m0, m1, ident) - tau) <= 1e-12 * max(1.0, abs(tau))
        perm = rng.permutation(n)
        assert abs(oracle.tau_hat_dr_est(w, y, p, m0, m1, (perm + 1).astype(np.int32)) - tau) <= 1e-12 * max(1.0, abs(tau))
        # a replicate is a function of the scores alone (nuisances are not refit, :267-283)
        idx = rng.integers(1, n + 1, n).astype(np.int32)
        assert abs(oracle.tau_hat_dr_est(w, y, p, m0, m1, idx) - (e1[idx - 1].mean() + e2[idx - 1].mean())) <= 1e-12

    prop()


@pytest.mark.parametrize("tag,n,p0", [("a", 200, 5), ("b", 5000, 21)])
def test_glm_and_lm_regression_fixtures(oracle, golden, tag, n, p0):
    """Pins of the oracle's glm.fit / lm restatements on the fixture inputs (tools/make_golden.py)."""
    reg = golden["reg"]
    d = synth.make_dataset(n, p0)
    xs, _, _ = oracle.scale(d["x_raw"], cols=np.arange(d["n_continuous"]))
    xw = np.column_stack([xs, d["w"]])
    gm, gp = oracle.glm_binomial(xw, d["y"]), oracle.glm_binomial(xs, d["w"])
    assert [gm["iters"], gp["iters"]] == reg[f"{tag}_glm_iters"].tolist()
    assert np.abs(gm["coef"] - reg[f"{tag}_glm_out"]).max() < 1e-11 and np.abs(gp["coef"] - reg[f"{tag}_glm_ps"]).max() < 1e-11
    assert np.allclose([gm["deviance"], gp["deviance"]], reg[f"{tag}_glm_dev"], rtol=1e-12)
    lm = oracle.lm(xw, d["y"])
    assert np.abs(lm["coef"] - reg[f"{tag}_lm_coef"]).max() < 1e-12 and np.abs(lm["std_err"] - reg[f"{tag}_lm_se"]).max() < 1e-12
    assert abs(lm["sigma"] - float(reg[f"{tag}_lm_sigma"])) < 1e-13


def test_glm_binomial_against_sklearn_unpenalised_logistic(oracle):
    """Independent check of the glm.fit restatement: scikit-learn's unpenalised logistic regression
    (a different algorithm, L-BFGS) lands on the same MLE."""
    from sklearn.linear_model import LogisticRegression
    rng = np.random.default_rng(21)
    n, p = 3000, 5
    x = rng.standard_normal((n, p))
    y = (rng.random(n) < 1 / (1 + np.exp(-(x @ np.array([0.5, -1.0, 0.0, 0.3, 0.8]) - 0.2)))).astype(float)
    g = oracle.glm_binomial(x, y)
    sk = LogisticRegression(penalty=None, tol=1e-12, max_iter=5000).fit(x, y)
    assert np.abs(np.r_[sk.intercept_, sk.coef_[0]] - g["coef"]).max() < 1e-5
