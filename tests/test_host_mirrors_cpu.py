"""CPU tests of the host-side mirrors of the R closures (ate_replication_causalml_b200/ate_functions.py)
with a stand-in engine whose numerics come from the oracle.  What is checked here is the HOST logic
the mirrors add on top of the C ABI -- argument plumbing, the reference's quirks, the return frame --
not the kernels (those are the -m gpu tests)."""
import numpy as np
import pandas as pd
import pytest

from ate_replication_causalml_b200 import ate_functions as af
from ate_replication_causalml_b200 import synth


class OracleEngine:
    """Same method names and return shapes as engine.Engine, computed by the CPU oracle."""

    def __init__(self, O):
        self.O = O
        self.calls = []

    def cv_glmnet(self, x, y, family="gaussian", penalty_factor=None, foldid=None, seed=0, **kw):
        self.calls.append(("cv_glmnet", family, None if penalty_factor is None else tuple(penalty_factor)))
        if foldid is None:
            self.O.set_seed(seed)
            foldid = self.O.foldid(len(y), 10)
        return self.O.cv_glmnet(x, y, foldid, family, penalty_factor=penalty_factor)

    def predict(self, x, a0, beta, response_binomial=False):
        return self.O.predict(x, a0, beta, response_binomial)

    def expand_deg2(self, x, include_main=True):
        return self.O.expand_deg2(x, include_main)

    def glm_binomial(self, x, y, **kw):
        return self.O.glm_binomial(x, y)

    def lm(self, x, y, weights=None, want_resid=False):
        return self.O.lm(x, y, weights, want_resid)

    def aipw_score(self, w, y, p, m0, m1, want_scores=False):
        return self.O.aipw(w, y, p, m0, m1)

    def bootstrap(self, w, y, p, m0, m1, B=1000, idx=None, seed=0, **kw):
        assert idx is not None, "the stand-in only does validation mode"
        return self.O.bootstrap_idx(w, y, p, m0, m1, idx)

    def tau_hat_dr_est(self, w, y, p, m0, m1, idx):
        return self.O.tau_hat_dr_est(w, y, p, m0, m1, idx)

    def r_sample(self, n, k, kind):
        return self.O.sample_replace(n, k, kind)

    def crossfit_aipw(self, x, w, y, foldid=None, inner_ps=None, inner_out=None, boot_idx=None, B=0, **kw):
        self.calls.append(("crossfit_aipw", dict(kw, B=B)))
        o = self.O.crossfit_aipw(x, w, y, foldid, inner_ps, inner_out, use_1se=bool(kw.get("use_1se", 1)))
        r = dict(tau=o["tau"], se_sandwich=o["se"], se_bootstrap=float("nan"), phat=o["phat"], m0=o["m0"], m1=o["m1"])
        if B > 0:
            r["se_bootstrap"], r["boot"] = self.O.bootstrap_idx(w, y, o["phat"], o["m0"], o["m1"], boot_idx)
        return r


@pytest.fixture()
def setup(oracle):
    n, p0 = 1200, 6
    d = synth.make_dataset(n, p0)
    xs, _, _ = oracle.scale(d["x_raw"], cols=np.arange(d["n_continuous"]))
    names = [f"x{j}" for j in range(p0)]
    df = pd.DataFrame(np.asarray(xs), columns=names)
    df["W"] = d["w"]; df["Y"] = d["y"]
    eng = OracleEngine(oracle)
    af.set_engine(eng); af.set_covariates(names)
    yield df, names, d, eng
    af.set_engine(None)


def test_lasso_mirrors_use_the_references_penalty_factors_and_lambda_1se(setup, oracle):
    df, names, d, eng = setup
    fid = synth.fold_ids(len(df), 10, 3)
    row = af.ate_condmean_lasso(df, "W", "Y", foldid=fid)
    assert eng.calls[-1] == ("cv_glmnet", "gaussian", tuple([1.0] * len(names) + [0.0]))        # :98
    m = oracle.cv_glmnet(df[names + ["W"]].to_numpy(), df["Y"].to_numpy(), fid, penalty_factor=np.r_[np.ones(6), 0.0])
    assert row["ATE"][0] == m["beta"][6, m["idx_1se"]] and row["lower_ci"][0] == row["upper_ci"][0] == row["ATE"][0]  # :106-107
    assert list(row.columns) == ["Method", "ATE", "lower_ci", "upper_ci"] and row["Method"][0] == "Single-equation LASSO"
    row = af.ate_lasso(df, "W", "Y", foldid=fid)
    assert eng.calls[-1][2] == tuple([1.0] * 7) and row["Method"][0] == "Usual LASSO"          # :120
    p = af.prop_score_lasso(df, "W", foldid=fid)
    assert p.shape == (len(df), 1) and 0 < p.min() and p.max() < 1                               # n x 1 matrix, :144-145


def test_doubly_robust_glm_mirror_and_tail(setup, oracle):
    df, names, d, eng = setup
    row = af.doubly_robust_glm(df, "W", "Y")
    x = df[names].to_numpy(); w = df["W"].to_numpy(); y = df["Y"].to_numpy()
    cm = oracle.glm_binomial(np.column_stack([x, w]), y)["coef"]
    pc = oracle.glm_binomial(x, w)["coef"]
    m1 = oracle.predict(np.column_stack([x, np.ones(len(w))]), cm[0], cm[1:], True)
    m0 = oracle.predict(np.column_stack([x, np.zeros(len(w))]), cm[0], cm[1:], True)
    p = oracle.predict(x, pc[0], pc[1:], True)
    tau, se = oracle.aipw(w, y, p, m0, m1)
    assert abs(row["ATE"][0] - tau) < 1e-14 and abs(row["lower_ci"][0] - (tau - 1.96 * se)) < 1e-14
    oracle.set_seed(5)
    bidx = np.stack([oracle.sample_replace(len(w), len(w), 1) for _ in range(7)])
    rowb = af.doubly_robust_glm(df, "W", "Y", bootstrap_se=True, boot_idx=bidx)
    sd, _ = oracle.bootstrap_idx(w, y, p, m0, m1, bidx)
    assert abs(rowb["upper_ci"][0] - (tau + 1.96 * sd)) < 1e-14
    # tau_hat_dr_est draws its own indices from the engine's R stream when none are passed (:269)
    oracle.set_seed(11); expect_idx = oracle.sample_replace(len(w), len(w), 1)
    oracle.set_seed(11); got = af.tau_hat_dr_est(w, y, p, m0, m1)
    assert got == oracle.tau_hat_dr_est(w, y, p, m0, m1, expect_idx)


def test_ipw_and_ols_mirrors(setup, oracle):
    df, names, d, eng = setup
    x = df[names].to_numpy(); w = df["W"].to_numpy(); y = df["Y"].to_numpy(); n = len(w)
    pc = oracle.glm_binomial(x, w)["coef"]; p = oracle.predict(x, pc[0], pc[1:], True)
    row = af.prop_score_weight(df, p, "W", "Y", names)
    ti = (w - p) * y / (p * (1 - p))
    e = oracle.lm(x * (w - p)[:, None], ti, want_resid=True)["resid"]
    assert abs(row["ATE"][0] - ti.mean()) < 1e-13
    assert abs((row["upper_ci"][0] - row["ATE"][0]) / 1.96 - np.sqrt((e ** 2).mean()) / np.sqrt(n)) < 1e-13   # :57
    row = af.prop_score_ols(df, p.reshape(-1, 1), "W", "Y")                      # an n x 1 matrix works as in R
    o = oracle.lm(w.reshape(-1, 1), y, w / p + (1 - w) / (1 - p))
    assert abs(row["ATE"][0] - o["coef"][1]) < 1e-13
    row = af.ate_condmean_ols(df, "W", "Y")
    o = oracle.lm(np.column_stack([x, w]), y)
    assert abs(row["ATE"][0] - o["coef"][-1]) < 1e-13 and abs(row["upper_ci"][0] - o["coef"][-1] - 1.96 * o["std_err"][-1]) < 1e-13
    # an aliased treatment column has no row in summary(lm)'s table: R stops with subscript out of bounds
    df2 = df.copy(); df2["W2"] = df2["W"]
    with pytest.raises(IndexError):
        af.ate_condmean_ols(df2, "W2", "Y")


def test_belloni_quirks(setup, oracle):
    """:308-314: BOTH models are read at model_xw's lambda.min, only POSITIVE coefficients count, and
    the `- 1` shifts the union down by one column (index 0 silently vanishes)."""
    df, names, d, eng = setup
    f1 = synth.fold_ids(len(df), 10, 1); f2 = synth.fold_ids(len(df), 10, 2)
    c_union, x, c_xw, c_xy = af.belloni_selection(df, "W", "Y", foldid_xw=f1, foldid_xy=f2, return_coefs=True)
    assert x.shape == (len(df), 6 + 36)                                       # mains, then c1-major products
    assert np.array_equal(x[:, 6 + 1], x[:, 0] * x[:, 1]) and np.array_equal(x[:, 6 + 6], x[:, 1] * x[:, 0])
    mw = oracle.cv_glmnet(x, df["W"].to_numpy(), f1)
    assert np.array_equal(c_xw, mw["beta"][:, mw["idx_min"]])
    pos = np.unique(np.r_[np.nonzero(c_xw > 0)[0], np.nonzero(c_xy > 0)[0]])   # 0-based positions of positive coefs
    expect = pos[pos > 0]                                                      # which() is 1-based; minus one; x[, 0] drops
    assert sorted(c_union) == sorted(expect)
    row = af.belloni(df, "W", "Y", foldid_xw=f1, foldid_xy=f2)
    o = oracle.lm(np.column_stack([x[:, c_union - 1], df["W"].to_numpy()]), df["Y"].to_numpy())
    assert abs(row["ATE"][0] - o["coef"][-1]) < 1e-12 and row["Method"][0] == "Belloni et.al"


def test_crossfit_mirror_frame_and_options(setup, oracle):
    from conftest import make_folds
    df, names, d, eng = setup
    n = len(df)
    fo, ips, iout = make_folds(n)
    oracle.set_seed(2)
    bidx = np.stack([oracle.sample_replace(n, n, 1) for _ in range(9)])
    row = af.crossfit_aipw(df, "W", "Y", B=9, foldid=fo, inner_ps=ips, inner_out=iout, boot_idx=bidx)
    det = row.attrs["detail"]
    name, opts = eng.calls[-1]
    assert name == "crossfit_aipw" and opts["B"] == 9 and opts["boot_mode"] == 0 and opts["K"] == 5 and opts["use_1se"] == 1
    assert list(row.columns) == ["Method", "ATE", "lower_ci", "upper_ci"]
    assert row["ATE"][0] == det["tau"] and abs(row["upper_ci"][0] - (det["tau"] + 1.96 * det["se_bootstrap"])) < 1e-15
    row2 = af.crossfit_aipw(df, "W", "Y", bootstrap_se=False, foldid=fo, inner_ps=ips, inner_out=iout)
    assert eng.calls[-1][1]["B"] == 0 and abs(row2["lower_ci"][0] - (det["tau"] - 1.96 * det["se_sandwich"])) < 1e-15
