"""ctypes front-end of the CPU oracle (oracle/ate_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl reference`` legs
may import this module.  PARITY UNPINNED: see the header of ate_oracle.c and DESIGN.md.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libate_oracle.so")

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "ate_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"], stdout=subprocess.DEVNULL,
                              stderr=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.orc_r_mean.restype = C.c_double
        _lib.orc_r_sd.restype = C.c_double
        _lib.orc_r_unif_rand.restype = C.c_double
        _lib.orc_tau_hat_dr_est.restype = C.c_double
        _lib.orc_bootstrap_idx.restype = C.c_double
        _lib.orc_bootstrap_r.restype = C.c_double
    return _lib


def _d(a):
    return a.ctypes.data_as(_dp) if a is not None else None


def _i(a):
    return a.ctypes.data_as(_ip) if a is not None else None


def _f(a):
    """column-major float64 copy"""
    return np.asfortranarray(np.asarray(a, dtype=np.float64))


def _v(a):
    return None if a is None else np.ascontiguousarray(np.asarray(a, dtype=np.float64).ravel())


def set_threads(t: int) -> int:
    """Deal `t` OpenMP threads over the composite's two levels of independent path fits (1 = what
    R does).  Returns the thread count in effect.  Only bench.py's timed CPU legs call this."""
    return int(lib().orc_set_threads(C.c_int(int(t))))


def get_threads() -> int:
    return int(lib().orc_get_threads())


# ---------------------------------------------------------------- base R numerics
def r_mean(x):
    x = _v(x)
    return lib().orc_r_mean(_d(x), C.c_long(x.size))


def r_sd(x):
    x = _v(x)
    return lib().orc_r_sd(_d(x), C.c_long(x.size))


def set_seed(seed: int):
    lib().orc_r_set_seed(C.c_uint32(seed & 0xFFFFFFFF))


def unif_rand():
    return lib().orc_r_unif_rand()


def sample_replace(n: int, k: int, kind: int = 1):
    out = np.empty(k, dtype=np.int32)
    lib().orc_r_sample_replace(C.c_int(n), C.c_long(k), C.c_int(kind), _i(out))
    return out


def sample_perm(n: int, kind: int = 1):
    out = np.empty(n, dtype=np.int32)
    lib().orc_r_sample_perm(C.c_int(n), C.c_int(kind), _i(out))
    return out


def foldid(n: int, nfolds: int, kind: int = 1):
    out = np.empty(n, dtype=np.int32)
    lib().orc_r_foldid(C.c_int(n), C.c_int(nfolds), C.c_int(kind), _i(out))
    return out


def scale(x, cols=None):
    x = _f(x).copy(order="F")
    n, p = x.shape
    if cols is None:
        m = p
        cp = None
    else:
        cols = np.ascontiguousarray(cols, dtype=np.int32)
        m = cols.size
        cp = _i(cols)
    center = np.empty(m)
    scl = np.empty(m)
    lib().orc_scale(_d(x), C.c_long(n), C.c_int(p), cp, C.c_int(m), _d(center), _d(scl))
    return x, center, scl


def expand_deg2(x, include_main: bool = True):
    x = _f(x)
    n, p0 = x.shape
    pout = p0 * p0 + (p0 if include_main else 0)
    out = np.empty((n, pout), order="F")
    lib().orc_expand_deg2(_d(x), C.c_long(n), C.c_int(p0), C.c_int(int(include_main)), _d(out))
    return out


# ---------------------------------------------------------------- glmnet
def glmnet(x, y, family="gaussian", weights=None, penalty_factor=None, alpha=1.0, nlambda=100,
           lambda_min_ratio=0.0, lambda_=None, thresh=1e-7, maxit=100000, standardize=True,
           intercept=True):
    x = _f(x)
    y = _v(y)
    n, p = x.shape
    w = _v(weights)
    vp = _v(penalty_factor)
    lam_in = _v(lambda_)
    nl = nlambda if lam_in is None else lam_in.size
    lmu = C.c_int(0)
    nlp = C.c_int(0)
    nulldev = C.c_double(0)
    lam = np.zeros(nl)
    a0 = np.zeros(nl)
    beta = np.zeros((p, nl), order="F")
    dev = np.zeros(nl)
    jerr = lib().orc_glmnet(C.c_int(0 if family == "gaussian" else 1), C.c_long(n), C.c_int(p),
                            _d(x), _d(y), _d(w), _d(vp), C.c_double(alpha), C.c_int(nlambda),
                            C.c_double(lambda_min_ratio), _d(lam_in), C.c_int(nl),
                            C.c_double(thresh), C.c_int(maxit), C.c_int(int(standardize)),
                            C.c_int(int(intercept)), C.byref(lmu), _d(lam), _d(a0), _d(beta),
                            _d(dev), C.byref(nlp), C.byref(nulldev))
    L = lmu.value
    return dict(jerr=jerr, lmu=L, lambda_=lam[:L].copy(), a0=a0[:L].copy(),
                beta=np.ascontiguousarray(beta[:, :L]), dev_ratio=dev[:L].copy(), npasses=nlp.value,
                nulldev=nulldev.value)


def cv_glmnet(x, y, foldid_, family="gaussian", penalty_factor=None, alpha=1.0, nlambda=100,
              lambda_min_ratio=0.0, thresh=1e-7, maxit=100000, standardize=True, intercept=True):
    x = _f(x)
    y = _v(y)
    n, p = x.shape
    vp = _v(penalty_factor)
    fid = np.ascontiguousarray(foldid_, dtype=np.int32)
    nfolds = int(fid.max())
    lmu = C.c_int(0)
    imin = C.c_int(0)
    i1se = C.c_int(0)
    lam = np.zeros(nlambda)
    a0 = np.zeros(nlambda)
    beta = np.zeros((p, nlambda), order="F")
    cvm = np.zeros(nlambda)
    cvsd = np.zeros(nlambda)
    dev = np.zeros(nlambda)
    jerr = lib().orc_cv_glmnet(C.c_int(0 if family == "gaussian" else 1), C.c_long(n), C.c_int(p),
                               _d(x), _d(y), _d(vp), C.c_double(alpha), C.c_int(nfolds), _i(fid),
                               C.c_int(nlambda), C.c_double(lambda_min_ratio), C.c_double(thresh),
                               C.c_int(maxit), C.c_int(int(standardize)), C.c_int(int(intercept)),
                               C.byref(lmu), _d(lam), _d(a0), _d(beta), _d(cvm), _d(cvsd),
                               C.byref(imin), C.byref(i1se), _d(dev))
    L = lmu.value
    return dict(jerr=jerr, lmu=L, lambda_=lam[:L].copy(), a0=a0[:L].copy(),
                beta=np.ascontiguousarray(beta[:, :L]), cvm=cvm[:L].copy(), cvsd=cvsd[:L].copy(),
                idx_min=imin.value, idx_1se=i1se.value, dev_ratio=dev[:L].copy(),
                lambda_min=lam[imin.value], lambda_1se=lam[i1se.value])


def predict(x, a0, beta, response_binomial=False):
    x = _f(x)
    n, p = x.shape
    beta = _v(beta)
    out = np.empty(n)
    lib().orc_predict(C.c_long(n), C.c_int(p), _d(x), C.c_double(a0), _d(beta),
                      C.c_int(int(response_binomial)), _d(out))
    return out


# ---------------------------------------------------------------- AIPW + bootstrap
def aipw(w, y, p, m0, m1, want_scores=False):
    w, y, p, m0, m1 = map(_v, (w, y, p, m0, m1))
    n = w.size
    tau = C.c_double(0)
    se = C.c_double(0)
    e1 = np.empty(n) if want_scores else None
    e2 = np.empty(n) if want_scores else None
    lib().orc_aipw(C.c_long(n), _d(w), _d(y), _d(p), _d(m0), _d(m1), C.byref(tau), C.byref(se),
                   _d(e1), _d(e2))
    if want_scores:
        return tau.value, se.value, e1, e2
    return tau.value, se.value


def tau_hat_dr_est(w, y, p, m0, m1, idx):
    w, y, p, m0, m1 = map(_v, (w, y, p, m0, m1))
    idx = np.ascontiguousarray(idx, dtype=np.int32)
    return lib().orc_tau_hat_dr_est(C.c_long(w.size), _d(w), _d(y), _d(p), _d(m0), _d(m1), _i(idx))


def bootstrap_idx(w, y, p, m0, m1, idx):
    """idx: (B, n) 1-based int32.  Returns (sd, replicates[B])."""
    w, y, p, m0, m1 = map(_v, (w, y, p, m0, m1))
    idx = np.ascontiguousarray(idx, dtype=np.int32)
    B, n = idx.shape
    boot = np.empty(B)
    sd = lib().orc_bootstrap_idx(C.c_long(n), _d(w), _d(y), _d(p), _d(m0), _d(m1), _i(idx),
                                 C.c_int(B), _d(boot))
    return sd, boot


def bootstrap_r(w, y, p, m0, m1, B, kind=1):
    w, y, p, m0, m1 = map(_v, (w, y, p, m0, m1))
    boot = np.empty(B)
    sd = lib().orc_bootstrap_r(C.c_long(w.size), _d(w), _d(y), _d(p), _d(m0), _d(m1), C.c_int(B),
                               C.c_int(kind), _d(boot))
    return sd, boot


def crossfit_aipw(x, w, y, foldid_, inner_ps, inner_out, nfolds_inner=10, nlambda=100,
                  thresh=1e-7, maxit=100000, use_1se=True):
    x = _f(x)
    w = _v(w)
    y = _v(y)
    n, p = x.shape
    fid = np.ascontiguousarray(foldid_, dtype=np.int32)
    K = int(fid.max())
    ips = np.ascontiguousarray(inner_ps, dtype=np.int32).reshape(K, n)
    iout = np.ascontiguousarray(inner_out, dtype=np.int32).reshape(K, n)
    phat = np.full(n, np.nan)
    m0 = np.full(n, np.nan)
    m1 = np.full(n, np.nan)
    tau = C.c_double(0)
    se = C.c_double(0)
    lps = np.zeros(K)
    lout = np.zeros(K)
    rc = lib().orc_crossfit_aipw(C.c_long(n), C.c_int(p), _d(x), _d(w), _d(y), C.c_int(K), _i(fid),
                                 C.c_int(nfolds_inner), _i(ips), _i(iout), C.c_int(nlambda),
                                 C.c_double(thresh), C.c_int(maxit), C.c_int(int(use_1se)), _d(phat),
                                 _d(m0), _d(m1), C.byref(tau), C.byref(se), _d(lps), _d(lout))
    return dict(rc=rc, phat=phat, m0=m0, m1=m1, tau=tau.value, se=se.value, lambda_ps=lps,
                lambda_out=lout)


def glm_binomial(x, y, maxit=25, epsilon=1e-8):
    """stats::glm(y ~ x, family = binomial): returns dict(coef[(Intercept), x...], deviance, iters, converged)."""
    x = _f(x)
    y = _v(y)
    n, p = x.shape
    coef = np.zeros(p + 1)
    dev = C.c_double(0)
    it = C.c_int(0)
    rc = lib().orc_glm_binomial(C.c_long(n), C.c_int(p), _d(x), _d(y), C.c_int(maxit), C.c_double(epsilon),
                                _d(coef), C.byref(dev), C.byref(it))
    return dict(coef=coef, deviance=dev.value, iters=it.value, converged=(rc == 0))


def lm(x, y, weights=None, want_resid=False):
    """stats::lm(y ~ x[, weights]) + summary(): coef/std_err [(Intercept), x...] (NaN = NA), sigma, rss,
    rank, df_residual, resid."""
    x = _f(x)
    y = _v(y)
    n, p = x.shape
    w = _v(weights)
    coef = np.zeros(p + 1)
    se = np.zeros(p + 1)
    sigma = C.c_double(0)
    rss = C.c_double(0)
    rdf = C.c_long(0)
    resid = np.empty(n) if want_resid else None
    rank = lib().orc_lm(C.c_long(n), C.c_int(p), _d(x), _d(y), _d(w), _d(coef), _d(se), C.byref(sigma),
                        C.byref(rss), C.byref(rdf), _d(resid))
    return dict(coef=coef, std_err=se, sigma=sigma.value, rss=rss.value, rank=rank, df_residual=rdf.value,
                resid=resid)
