"""Host-side mirror of the reference's R entry points (`/root/reference/ate_functions.R`).

R is not installed in this image, so the closures a maintainer would keep in `ate_functions.R`
(see INTEGRATION.md for the `.Call` versions) are mirrored here in Python with the same names,
argument meaning, return shape (`Method, ATE, lower_ci, upper_ci`) and quirks:

* like the R code, the lasso estimators read the module-level global ``covariates``
  (ate_functions.R:91,113,135,289) -- set it with ``set_covariates([...])``;
* ``coef(model)`` / ``predict(model)`` on a cv.glmnet object use ``s = "lambda.1se"``
  (ate_functions.R:106,128,144); ``belloni`` evaluates BOTH models at ``model_xw$lambda.min``
  (:308-309), keeps only positive coefficients (:312-313) and shifts the union by one (:314);
* ``doubly_robust_glm`` hard-codes ``B <- 1000`` (:247) -- kept as the default of a new
  trailing argument.

All numerical work goes through the C ABI (``Engine``); nothing here falls back to the CPU.
``dataset`` is a pandas DataFrame (the stand-in for R's data.frame).
"""
from __future__ import annotations

import numpy as np
import pandas as pd

from .engine import Engine, RNG_REJECTION

covariates: list[str] = []
_engine: Engine | None = None


def set_covariates(names):
    global covariates
    covariates = list(names)


def engine() -> Engine:
    global _engine
    if _engine is None:
        _engine = Engine(0)
    return _engine


def set_engine(e: Engine):
    global _engine
    _engine = e


def _row(method, tau_hat, lower_ci, upper_ci):
    return pd.DataFrame({"Method": [method], "ATE": [tau_hat], "lower_ci": [lower_ci],
                         "upper_ci": [upper_ci]})


def _r_mean(x):
    """base::mean on a double vector: long-double sum / n, then one long-double correction pass
    (summary.c: real_mean).  The R closure keeps calling mean() itself; this is its host stand-in."""
    x = np.asarray(x, dtype=np.float64)
    s = np.sum(x, dtype=np.longdouble) / x.size
    s += np.sum(x.astype(np.longdouble) - s) / x.size
    return float(s)


def _table_row(method, fit, j):
    """tau_hat / se_hat from row j of summary(lm)$coefficients, then the +-1.96 interval."""
    tau_hat, se_hat = float(fit["coef"][j]), float(fit["std_err"][j])
    if tau_hat != tau_hat:
        # R: reg_coef["W", "Estimate"] on a table that dropped the aliased row -> subscript out of bounds
        raise IndexError("subscript out of bounds: the treatment coefficient is aliased (NA)")
    return _row(method, tau_hat, tau_hat - se_hat * 1.96, tau_hat + se_hat * 1.96)


# ---------------------------------------------------------------- ate_functions.R:24-39
def ate_condmean_ols(dataset, treatment_var, outcome_var, method="Direct Method"):
    """lm(Y ~ .) on every other column (:25-29); Estimate and Std. Error of the treatment row (:33-34)."""
    cols = [c for c in dataset.columns if c != outcome_var]
    fit = engine().lm(dataset[cols].to_numpy(dtype=np.float64), dataset[outcome_var].to_numpy(dtype=np.float64))
    return _table_row(method, fit, 1 + cols.index(treatment_var))


# ---------------------------------------------------------------- ate_functions.R:44-64
def prop_score_weight(dataset, p, treatment_var, outcome_var, covariates, method="Propensity_Weighting"):
    """tau_hat_i = (W - p) Y / (p (1 - p)) (:47); its mean is the estimate (:59); the SE comes from the
    residuals of tau_hat_i regressed on covariates * (W - p) (:50-57).  `covariates` is an ARGUMENT here,
    as in the reference."""
    w = dataset[treatment_var].to_numpy(dtype=np.float64)
    y = dataset[outcome_var].to_numpy(dtype=np.float64)
    p = np.asarray(p, dtype=np.float64).ravel()
    tau_i = ((w - p) * y) / (p * (1 - p))
    ps_er = w - p
    d = dataset[list(covariates)].to_numpy(dtype=np.float64) * ps_er[:, None]
    fit = engine().lm(d, tau_i)
    n = w.size
    se_hat = np.sqrt(fit["rss"] / n) / np.sqrt(n)            # sqrt(mean(e^2)) / sqrt(NROW)
    tau_hat = _r_mean(tau_i)
    return _row(method, tau_hat, tau_hat - se_hat * 1.96, tau_hat + se_hat * 1.96)


# ---------------------------------------------------------------- ate_functions.R:67-86
def prop_score_ols(dataset, p, treatment_var, outcome_var, method="Propensity_Regression"):
    """lm(Y ~ W, weights = W/p + (1 - W)/(1 - p)) (:72-76); treatment row of the table (:80-81)."""
    w = dataset[treatment_var].to_numpy(dtype=np.float64)
    y = dataset[outcome_var].to_numpy(dtype=np.float64)
    p = np.asarray(p, dtype=np.float64).ravel()
    weights = (w / p) + ((1 - w) / (1 - p))
    fit = engine().lm(w.reshape(-1, 1), y, weights=weights)
    return _table_row(method, fit, 1)


# ---------------------------------------------------------------- ate_functions.R:89-108
def ate_condmean_lasso(dataset, treatment_var, outcome_var, foldid=None, seed=None):
    regs = list(covariates) + [treatment_var]
    x = dataset[regs].to_numpy(dtype=np.float64)
    y = dataset[outcome_var].to_numpy(dtype=np.float64)
    pfac = np.r_[np.ones(len(covariates)), 0.0]          # :98
    model = engine().cv_glmnet(x, y, "gaussian", penalty_factor=pfac, foldid=foldid,
                               seed=0 if seed is None else seed)
    betaw = float(model["beta"][len(covariates), model["idx_1se"]])  # coef(model)[treatment_var,]
    return _row("Single-equation LASSO", betaw, betaw, betaw)


# ---------------------------------------------------------------- ate_functions.R:111-130
def ate_lasso(dataset, treatment_var, outcome_var, foldid=None, seed=None):
    regs = list(covariates) + [treatment_var]
    x = dataset[regs].to_numpy(dtype=np.float64)
    y = dataset[outcome_var].to_numpy(dtype=np.float64)
    pfac = np.r_[np.ones(len(covariates)), 1.0]          # :120
    model = engine().cv_glmnet(x, y, "gaussian", penalty_factor=pfac, foldid=foldid,
                               seed=0 if seed is None else seed)
    betaw = float(model["beta"][len(covariates), model["idx_1se"]])
    return _row("Usual LASSO", betaw, betaw, betaw)


# ---------------------------------------------------------------- ate_functions.R:133-146
def prop_score_lasso(dataset, treatment_var, foldid=None, seed=None):
    """Returns an n x 1 matrix, as predict(cv.glmnet, type = "response") does."""
    x = dataset[list(covariates)].to_numpy(dtype=np.float64)
    w = dataset[treatment_var].to_numpy(dtype=np.float64)
    model = engine().cv_glmnet(x, w, "binomial", foldid=foldid, seed=0 if seed is None else seed)
    s = model["idx_1se"]
    p = engine().predict(x, model["a0"][s], model["beta"][:, s], response_binomial=True)
    return p.reshape(-1, 1)


# ---------------------------------------------------------------- ate_functions.R:267-283
def tau_hat_dr_est(w, y, p, tauhat0x, tauhat1x, bootstrap_sample_idx=None):
    """One bootstrap replicate.  The R function draws sample(length(w), length(w), replace = T)
    from the global RNG; pass that vector (1-based), or None to draw it with the engine's
    R-compatible Mersenne-Twister stream (Engine.r_set_seed)."""
    w = np.asarray(w, dtype=np.float64)
    if bootstrap_sample_idx is None:
        bootstrap_sample_idx = engine().r_sample(w.size, w.size, RNG_REJECTION)
    return engine().tau_hat_dr_est(w, y, p, tauhat0x, tauhat1x, bootstrap_sample_idx)


# ---------------------------------------------------------------- ate_functions.R:241-263
def doubly_robust_scores(w, y, p, tauhat0x, tauhat1x, bootstrap_se=False, B=1000, boot_idx=None,
                         boot_seed=1991, method="Doubly Robust"):
    """Lines :241-263 of doubly_robust_glm for given nuisance vectors: AIPW point estimate,
    sandwich or bootstrap SE, +-1.96 se interval."""
    tau_hat, se_sand = engine().aipw_score(w, y, p, tauhat0x, tauhat1x)
    if bootstrap_se:
        se_hat, _ = engine().bootstrap(w, y, p, tauhat0x, tauhat1x, B=B, idx=boot_idx, seed=boot_seed)
    else:
        se_hat = se_sand
    return _row(method, tau_hat, tau_hat - se_hat * 1.96, tau_hat + se_hat * 1.96)


# ---------------------------------------------------------------- ate_functions.R:211-264
def doubly_robust_glm(dataset, treatment_var, outcome_var, bootstrap_se=False, B=1000, boot_idx=None,
                      boot_seed=1991):
    """The reference function as written: logistic glm for the outcome on all other columns
    (:218-220), predicted at W := 1 and W := 0 (:221-228), logistic glm propensity on everything but
    the outcome (:231-234), AIPW (:241-243), sandwich or bootstrap SE (:245-256), interval (:260-263).
    Every column of `dataset` other than the outcome is a regressor, as `Y ~ .` makes it."""
    cols = [c for c in dataset.columns if c != outcome_var]
    xo = dataset[cols].to_numpy(dtype=np.float64)
    y = dataset[outcome_var].to_numpy(dtype=np.float64)
    w = dataset[treatment_var].to_numpy(dtype=np.float64)
    condmean = engine().glm_binomial(xo, y)["coef"]
    jw = cols.index(treatment_var)
    x1 = xo.copy(); x1[:, jw] = 1.0
    x0 = xo.copy(); x0[:, jw] = 0.0
    tauhat1x = engine().predict(x1, condmean[0], condmean[1:], response_binomial=True)
    tauhat0x = engine().predict(x0, condmean[0], condmean[1:], response_binomial=True)
    pcols = [c for c in cols if c != treatment_var]
    xp = dataset[pcols].to_numpy(dtype=np.float64)
    pcoef = engine().glm_binomial(xp, w)["coef"]
    p = engine().predict(xp, pcoef[0], pcoef[1:], response_binomial=True)
    return doubly_robust_scores(w, y, p, tauhat0x, tauhat1x, bootstrap_se=bootstrap_se, B=B, boot_idx=boot_idx,
                                boot_seed=boot_seed, method="Doubly Robust with logistic regression PS")


def crossfit_aipw(dataset, treatment_var, outcome_var, bootstrap_se=True, B=1000, K=5, seed=1991,
                  foldid=None, inner_ps=None, inner_out=None, boot_idx=None, expand=0,
                  n_continuous=0, use_1se=True, method="Cross-fitted AIPW with glmnet nuisances"):
    """The composite of SURVEY.md section 8: K-fold cross-fitted doubly-robust ATE with
    L1-logistic propensity (:139-141) and lasso outcome (:98-103) nuisances, AIPW score
    (:241-243), sandwich (:255-256) or bootstrap (:245-252) SE, interval (:260-263)."""
    x = dataset[list(covariates)].to_numpy(dtype=np.float64)
    w = dataset[treatment_var].to_numpy(dtype=np.float64)
    y = dataset[outcome_var].to_numpy(dtype=np.float64)
    r = engine().crossfit_aipw(x, w, y, foldid=foldid, inner_ps=inner_ps, inner_out=inner_out,
                               boot_idx=boot_idx, K=K, B=B if bootstrap_se else 0,
                               boot_mode=0 if boot_idx is not None else 1, boot_seed=seed, r_seed=seed,
                               expand=expand, n_continuous=n_continuous,
                               standardize_base=int(n_continuous > 0), use_1se=int(use_1se))
    se = r["se_bootstrap"] if bootstrap_se else r["se_sandwich"]
    out = _row(method, r["tau"], r["tau"] - 1.96 * se, r["tau"] + 1.96 * se)
    out.attrs["detail"] = r
    return out


# ---------------------------------------------------------------- ate_functions.R:286-328
def belloni_selection(dataset, treatment_var, outcome_var, foldid_xw=None, foldid_xy=None, seed=None,
                      return_coefs=False):
    """belloni() up to the post-lasso OLS (which SURVEY.md ranks as the next row, f4): the
    expansion (:288-296), the two gaussian cv.glmnet fits (:304-305), coefficients at
    model_xw$lambda.min for BOTH models (:308-309) and the selected support with the
    reference's `> 0` and `- 1` quirks (:312-314).  Returns (c_union, x_expanded)."""
    x0 = dataset[list(covariates)].to_numpy(dtype=np.float64)
    x = engine().expand_deg2(x0, include_main=True)
    w = dataset[treatment_var].to_numpy(dtype=np.float64)
    y = dataset[outcome_var].to_numpy(dtype=np.float64)
    model_xw = engine().cv_glmnet(x, w, "gaussian", foldid=foldid_xw, seed=0 if seed is None else seed)
    model_xy = engine().cv_glmnet(x, y, "gaussian", foldid=foldid_xy, seed=0 if seed is None else seed + 1)
    s = model_xw["lambda_min"]
    c_xw = model_xw["beta"][:, model_xw["idx_min"]]
    c_xy = _coef_at(model_xy, s)
    nz_xw = np.nonzero(c_xw > 0)[0] + 1       # which(), 1-based
    nz_xy = np.nonzero(c_xy > 0)[0] + 1
    c_union = pd.unique(np.r_[nz_xw, nz_xy]) - 1
    c_union = c_union[c_union > 0]            # x[, 0] is silently dropped by R
    if return_coefs:
        return c_union, x, c_xw, c_xy
    return c_union, x


def belloni(dataset, treatment_var, outcome_var, method="Belloni et.al", foldid_xw=None, foldid_xy=None, seed=None):
    """belloni() end to end (:286-328): the double selection above, then lm(y ~ cbind(x[, c_union], w))
    (:317-320) and the `x_restrictedW` row of its coefficient table (:322-323)."""
    c_union, x, c_xw, c_xy = belloni_selection(dataset, treatment_var, outcome_var, foldid_xw, foldid_xy, seed,
                                               return_coefs=True)
    w = dataset[treatment_var].to_numpy(dtype=np.float64)
    y = dataset[outcome_var].to_numpy(dtype=np.float64)
    x_restricted = np.column_stack([x[:, c_union - 1], w])     # R's 1-based x[, c_union]
    fit = engine().lm(x_restricted, y)
    out = _table_row(method, fit, x_restricted.shape[1])
    out.attrs.update(c_union=c_union, c_xw=c_xw, c_xy=c_xy, fit=fit)
    return out


def _coef_at(model, s):
    """coef(glmnet.fit, s = s): linear interpolation in lambda (lambda.interp)."""
    lam = model["lambda_"]
    if s >= lam[0]:
        return model["beta"][:, 0]
    if s <= lam[-1]:
        return model["beta"][:, -1]
    sfrac = (lam[0] - s) / (lam[0] - lam[-1])
    lamf = (lam[0] - lam) / (lam[0] - lam[-1])
    coord = np.interp(sfrac, lamf, np.arange(lam.size))
    left, right = int(np.floor(coord)), int(np.ceil(coord))
    if left == right:
        return model["beta"][:, left]
    f = (sfrac - lamf[right]) / (lamf[left] - lamf[right])
    return model["beta"][:, left] * f + model["beta"][:, right] * (1 - f)
