"""Engine: thin numpy-facing wrapper over the C ABI (include/ate_b200.h).

Every method is one C-ABI call with HOST buffers (column-major doubles, 1-based int32
indices: R's conventions); device memory lives inside the library.  Methods that accept
``on_device=True`` take raw CUDA device pointers (ints, e.g. ``tensor.data_ptr()``).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib as L
from ._lib import AteError, CrossfitOpts, CrossfitResult

GAUSSIAN, BINOMIAL = 0, 1
BOOT_INDEX, BOOT_PHILOX = 0, 1
RNG_ROUNDING, RNG_REJECTION = 0, 1


class Engine:
    def __init__(self, device: int = 0):
        self._lib = L.load()
        h = C.c_void_p()
        rc = self._lib.ate_ctx_create(int(device), C.byref(h))
        if rc != 0 or not h:
            msg = self._lib.ate_last_error(None)
            raise AteError(f"ate_ctx_create(device={device}) failed (rc={rc}): "
                           f"{msg.decode() if msg else ''} -- this library has no CPU fallback")
        self._h = h
        self.device = device

    def close(self):
        if getattr(self, "_h", None):
            self._lib.ate_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ------------------------------------------------------------------ helpers
    def _chk(self, rc, what):
        if rc < 0:
            msg = self._lib.ate_last_error(self._h)
            raise AteError(f"{what} failed (rc={rc}): {msg.decode() if msg else ''}")
        return rc

    @property
    def stream_ptr(self) -> int:
        """cudaStream_t of the ctx (wrap with torch.cuda.ExternalStream to record events)."""
        return int(self._lib.ate_ctx_stream(self._h) or 0)

    @property
    def launches(self) -> int:
        return int(self._lib.ate_launch_count(self._h))

    def profile_enable(self, on=1):
        """0 off, 1 the IRLS step kernel, 2 the full-width Gram kernel, 3 both."""
        self._chk(self._lib.ate_profile_enable(self._h, int(on)), "ate_profile_enable")

    def profile_read(self):
        ms = C.c_double()
        nl = C.c_longlong()
        fl = C.c_double()
        by = C.c_double()
        self._chk(self._lib.ate_profile_read(self._h, C.byref(ms), C.byref(nl), C.byref(fl), C.byref(by)),
                  "ate_profile_read")
        return dict(gram_ms=ms.value, gram_launches=nl.value, gram_flops=fl.value, gram_bytes=by.value)

    # ------------------------------------------------------------------ a1 / a2
    def standardize(self, x, cols=None):
        """scale() on the selected columns (ate_replication.Rmd:72-74).  Returns (x_scaled,
        center, scale)."""
        x = L.fmat(x).copy(order="F")
        n, p = x.shape
        if cols is None:
            cp, m = None, p
        else:
            cols = L.i32(cols)
            cp, m = cols, cols.size
        center = np.empty(m)
        scl = np.empty(m)
        self._chk(self._lib.ate_standardize(self._h, L.ptr(x), n, p, L.ptr(cp), m, L.ptr(center), L.ptr(scl)),
                  "ate_standardize")
        return x, center, scl

    def expand_deg2(self, x, include_main=True):
        """belloni()'s interaction expansion (ate_functions.R:288-296)."""
        x = L.fmat(x)
        n, p0 = x.shape
        out = np.empty((n, p0 * p0 + (p0 if include_main else 0)), order="F")
        self._chk(self._lib.ate_expand_deg2(self._h, L.ptr(x), n, p0, int(include_main), L.ptr(out)),
                  "ate_expand_deg2")
        return out

    # ------------------------------------------------------------------ a3 / a4 / a5
    def glmnet(self, x, y, family="gaussian", penalty_factor=None, nlambda=100, lambda_min_ratio=0.0,
               lambda_=None, thresh=1e-7, maxit=100000):
        x = L.fmat(x)
        y = L.f64(y)
        n, p = x.shape
        if y.size != n:
            raise ValueError("number of observations in y not equal to the number of rows of x")
        vp = None if penalty_factor is None else L.f64(penalty_factor)
        lam_in = None if lambda_ is None else L.f64(lambda_)
        nl = nlambda if lam_in is None else lam_in.size
        lmu = C.c_int()
        npass = C.c_int()
        jerr = C.c_int()
        nulldev = C.c_double()
        lam = np.zeros(nl)
        a0 = np.zeros(nl)
        beta = np.zeros((p, nl), order="F")
        dev = np.zeros(nl)
        fam = GAUSSIAN if family == "gaussian" else BINOMIAL
        self._chk(self._lib.ate_glmnet(self._h, fam, n, p, L.ptr(x), L.ptr(y), L.ptr(vp), nlambda,
                                       float(lambda_min_ratio), L.ptr(lam_in), nl, float(thresh), int(maxit),
                                       C.byref(lmu), L.ptr(lam), L.ptr(a0), L.ptr(beta), L.ptr(dev),
                                       C.byref(npass), C.byref(nulldev), C.byref(jerr)), "ate_glmnet")
        Lm = lmu.value
        return dict(jerr=jerr.value, lmu=Lm, lambda_=lam[:Lm].copy(), a0=a0[:Lm].copy(),
                    beta=np.ascontiguousarray(beta[:, :Lm]), dev_ratio=dev[:Lm].copy(),
                    npasses=npass.value, nulldev=nulldev.value)

    def cv_glmnet(self, x, y, family="gaussian", penalty_factor=None, nfolds=10, foldid=None, seed=0,
                  rng_kind=RNG_REJECTION, nlambda=100, thresh=1e-7, maxit=100000):
        x = L.fmat(x)
        y = L.f64(y)
        n, p = x.shape
        vp = None if penalty_factor is None else L.f64(penalty_factor)
        fid = None
        if foldid is not None:
            fid = L.i32(foldid)
            nfolds = int(fid.max())
        lmu = C.c_int()
        imin = C.c_int()
        i1se = C.c_int()
        jerr = C.c_int()
        lam = np.zeros(nlambda)
        a0 = np.zeros(nlambda)
        beta = np.zeros((p, nlambda), order="F")
        cvm = np.zeros(nlambda)
        cvsd = np.zeros(nlambda)
        dev = np.zeros(nlambda)
        fam = GAUSSIAN if family == "gaussian" else BINOMIAL
        self._chk(self._lib.ate_cv_glmnet(self._h, fam, n, p, L.ptr(x), L.ptr(y), L.ptr(vp), int(nfolds),
                                          L.ptr(fid), int(seed), int(rng_kind), int(nlambda), float(thresh),
                                          int(maxit), C.byref(lmu), L.ptr(lam), L.ptr(a0), L.ptr(beta),
                                          L.ptr(cvm), L.ptr(cvsd), C.byref(imin), C.byref(i1se), L.ptr(dev),
                                          C.byref(jerr)), "ate_cv_glmnet")
        Lm = lmu.value
        return dict(jerr=jerr.value, lmu=Lm, lambda_=lam[:Lm].copy(), a0=a0[:Lm].copy(),
                    beta=np.ascontiguousarray(beta[:, :Lm]), cvm=cvm[:Lm].copy(), cvsd=cvsd[:Lm].copy(),
                    idx_min=imin.value, idx_1se=i1se.value, dev_ratio=dev[:Lm].copy(),
                    lambda_min=lam[imin.value], lambda_1se=lam[i1se.value])

    def cv_glmnet_device(self, X_ptr, ld, n, p, foldid_ptr, nfolds=10, family="gaussian", penalty_factor=None,
                         nlambda=100, thresh=1e-7, maxit=100000, shard_rank=0, shard_count=0):
        """cv.glmnet on a device-resident row-major design [x | y | 1] (this rank's rows); see the header."""
        vp = None if penalty_factor is None else L.f64(penalty_factor)
        lmu = C.c_int()
        imin = C.c_int()
        i1se = C.c_int()
        jerr = C.c_int()
        lam = np.zeros(nlambda)
        a0 = np.zeros(nlambda)
        beta = np.zeros((p, nlambda), order="F")
        cvm = np.zeros(nlambda)
        cvsd = np.zeros(nlambda)
        dev = np.zeros(nlambda)
        fam = GAUSSIAN if family == "gaussian" else BINOMIAL
        self._chk(self._lib.ate_cv_glmnet_device(self._h, fam, int(n), int(p), L.ptr(int(X_ptr)), int(ld),
                                                 L.ptr(int(foldid_ptr)), int(nfolds), L.ptr(vp), int(nlambda),
                                                 float(thresh), int(maxit), int(shard_rank), int(shard_count),
                                                 C.byref(lmu), L.ptr(lam), L.ptr(a0), L.ptr(beta), L.ptr(cvm),
                                                 L.ptr(cvsd), C.byref(imin), C.byref(i1se), L.ptr(dev), C.byref(jerr)),
                  "ate_cv_glmnet_device")
        Lm = lmu.value
        return dict(jerr=jerr.value, lmu=Lm, lambda_=lam[:Lm].copy(), a0=a0[:Lm].copy(),
                    beta=np.ascontiguousarray(beta[:, :Lm]), cvm=cvm[:Lm].copy(), cvsd=cvsd[:Lm].copy(),
                    idx_min=imin.value, idx_1se=i1se.value, dev_ratio=dev[:Lm].copy(),
                    lambda_min=lam[imin.value], lambda_1se=lam[i1se.value])

    def glm_binomial(self, x, y, maxit=25, epsilon=1e-8):
        """stats::glm(y ~ x, family = binomial): coef = [(Intercept), x...] (ate_functions.R:218-234)."""
        x = L.fmat(x)
        y = L.f64(y)
        n, p = x.shape
        coef = np.zeros(p + 1)
        dev = C.c_double()
        it = C.c_int()
        conv = C.c_int()
        self._chk(self._lib.ate_glm_binomial(self._h, n, p, L.ptr(x), L.ptr(y), int(maxit), float(epsilon),
                                             L.ptr(coef), C.byref(dev), C.byref(it), C.byref(conv)),
                  "ate_glm_binomial")
        return dict(coef=coef, deviance=dev.value, iters=it.value, converged=bool(conv.value))

    def lm(self, x, y, weights=None, want_resid=False):
        """stats::lm(y ~ x[, weights]) + summary(): coef / std_err = [(Intercept), x...], NaN where R says NA
        (ate_functions.R:25-34, :52-56, :72-81, :320-323)."""
        x = L.fmat(x)
        y = L.f64(y)
        n, p = x.shape
        w = None if weights is None else L.f64(weights)
        coef = np.zeros(p + 1)
        se = np.zeros(p + 1)
        sigma = C.c_double()
        rss = C.c_double()
        rank = C.c_int()
        df = C.c_longlong()
        resid = np.empty(n) if want_resid else None
        self._chk(self._lib.ate_lm(self._h, n, p, L.ptr(x), L.ptr(y), L.ptr(w) if w is not None else None,
                                   L.ptr(coef), L.ptr(se), C.byref(sigma), C.byref(rss), C.byref(rank),
                                   C.byref(df), L.ptr(resid) if want_resid else None), "ate_lm")
        return dict(coef=coef, std_err=se, sigma=sigma.value, rss=rss.value, rank=rank.value,
                    df_residual=df.value, resid=resid)

    def predict(self, x, a0, beta, response_binomial=False):
        x = L.fmat(x)
        n, p = x.shape
        beta = L.f64(beta)
        out = np.empty(n)
        self._chk(self._lib.ate_predict(self._h, n, p, L.ptr(x), float(a0), L.ptr(beta),
                                        int(response_binomial), L.ptr(out)), "ate_predict")
        return out

    # ------------------------------------------------------------------ a7 - a10
    def aipw_score(self, w, y, p, m0, m1, want_scores=False):
        w, y, p, m0, m1 = map(L.f64, (w, y, p, m0, m1))
        n = w.size
        tau = C.c_double()
        se = C.c_double()
        e1 = np.empty(n) if want_scores else None
        e2 = np.empty(n) if want_scores else None
        self._chk(self._lib.ate_aipw_score(self._h, n, L.ptr(w), L.ptr(y), L.ptr(p), L.ptr(m0), L.ptr(m1),
                                           C.byref(tau), C.byref(se), L.ptr(e1), L.ptr(e2), 0),
                  "ate_aipw_score")
        if want_scores:
            return tau.value, se.value, e1, e2
        return tau.value, se.value

    def aipw_score_device(self, n, w, y, p, m0, m1):
        tau = C.c_double()
        se = C.c_double()
        self._chk(self._lib.ate_aipw_score(self._h, n, L.ptr(w), L.ptr(y), L.ptr(p), L.ptr(m0), L.ptr(m1),
                                           C.byref(tau), C.byref(se), None, None, 1), "ate_aipw_score")
        return tau.value, se.value

    def bootstrap(self, w, y, p, m0, m1, B=1000, idx=None, seed=1991, rep_offset=0, on_device=False, n=None):
        """Bootstrap replicates (ate_functions.R:245-252).  idx: (B, n) 1-based index matrix
        (validation mode) or None for on-device Philox.  Returns (sd, replicates)."""
        if on_device:
            args = [L.ptr(int(a)) for a in (w, y, p, m0, m1)]
        else:
            w, y, p, m0, m1 = map(L.f64, (w, y, p, m0, m1))
            n = w.size
            args = [L.ptr(a) for a in (w, y, p, m0, m1)]
        if idx is not None:
            if on_device:
                ip, B = L.ptr(int(idx)), B
            else:
                idx = np.ascontiguousarray(idx, dtype=np.int32)
                B = idx.shape[0]
                if idx.shape[1] != n:
                    raise ValueError("idx must be (B, n)")
                ip = L.ptr(idx)
            mode = BOOT_INDEX
        else:
            ip, mode = None, BOOT_PHILOX
        boot = np.empty(B)
        se = C.c_double()
        self._chk(self._lib.ate_bootstrap(self._h, n, *args, int(B), mode, ip, int(seed), int(rep_offset),
                                          L.ptr(boot), C.byref(se), int(on_device)), "ate_bootstrap")
        return se.value, boot

    def tau_hat_dr_est(self, w, y, p, tauhat0x, tauhat1x, idx):
        w, y, p, m0, m1 = map(L.f64, (w, y, p, tauhat0x, tauhat1x))
        idx = L.i32(idx)
        tau = C.c_double()
        self._chk(self._lib.ate_tau_hat_dr_est(self._h, w.size, L.ptr(w), L.ptr(y), L.ptr(p), L.ptr(m0),
                                               L.ptr(m1), L.ptr(idx), C.byref(tau)), "ate_tau_hat_dr_est")
        return tau.value

    # ------------------------------------------------------------------ f1: R's RNG
    def r_set_seed(self, seed: int):
        self._chk(self._lib.ate_r_set_seed(self._h, seed & 0xFFFFFFFF), "ate_r_set_seed")

    def r_sample(self, n: int, k: int, rng_kind=RNG_REJECTION):
        out = np.empty(k, dtype=np.int32)
        self._chk(self._lib.ate_r_sample(self._h, n, k, rng_kind, L.ptr(out)), "ate_r_sample")
        return out

    def r_foldid(self, n: int, nfolds: int, rng_kind=RNG_REJECTION):
        out = np.empty(n, dtype=np.int32)
        self._chk(self._lib.ate_r_foldid(self._h, n, nfolds, rng_kind, L.ptr(out)), "ate_r_foldid")
        return out

    # ------------------------------------------------------------------ Gram (tests / roofline)
    def weighted_gram(self, x, v=None, rows=None):
        x = L.fmat(x)
        n, p = x.shape
        v = None if v is None else L.f64(v)
        rows = None if rows is None else L.i32(rows)
        G = np.empty((p, p))
        self._chk(self._lib.ate_weighted_gram(self._h, n, p, L.ptr(x), 0, L.ptr(v), L.ptr(rows),
                                              0 if rows is None else rows.size, L.ptr(G), 0),
                  "ate_weighted_gram")
        return G

    def weighted_gram_device(self, n, p, X_ptr, ld, v_ptr, G_ptr, rows_ptr=None, nrows=None):
        self._chk(self._lib.ate_weighted_gram(self._h, n, p, L.ptr(int(X_ptr)), ld,
                                              L.ptr(int(v_ptr)) if v_ptr else None,
                                              L.ptr(int(rows_ptr)) if rows_ptr else None,
                                              n if nrows is None else nrows, L.ptr(int(G_ptr)), 1),
                  "ate_weighted_gram")

    # ------------------------------------------------------------------ the composite
    def default_opts(self) -> CrossfitOpts:
        o = CrossfitOpts()
        self._lib.ate_crossfit_default_opts(C.byref(o))
        return o

    def crossfit_aipw(self, x_base, w, y, opts: CrossfitOpts | None = None, foldid=None, inner_ps=None,
                      inner_out=None, boot_idx=None, n=None, p0=None, want_nuisances=True, **kw):
        """K-fold cross-fitted AIPW with glmnet nuisances + bootstrap SE (one C-ABI call).

        Host mode: x_base (n, p0) array, w, y vectors.  Device mode (opts.on_device=1):
        x_base / w / y are device pointers (column-major base covariates) and n, p0 are given.
        Keyword arguments override fields of ``opts``."""
        o = opts if opts is not None else self.default_opts()
        for k, v in kw.items():
            if not hasattr(o, k):
                raise TypeError(f"unknown option {k}")
            setattr(o, k, v)
        if o.on_device:
            xp, wp, yp = (L.ptr(int(a)) for a in (x_base, w, y))
        else:
            x_base = L.fmat(x_base)
            n, p0 = x_base.shape
            w = L.f64(w)
            y = L.f64(y)
            xp, wp, yp = L.ptr(x_base), L.ptr(w), L.ptr(y)
        K = o.K
        if o.on_device:   # every input array is a device pointer, the fold ids and boot_idx included
            fid, ips, iout = (None if a is None else int(a) for a in (foldid, inner_ps, inner_out))
            bidx = None if boot_idx is None else int(boot_idx)
        else:
            fid = None if foldid is None else L.i32(foldid)
            ips = None if inner_ps is None else L.i32(inner_ps)
            iout = None if inner_out is None else L.i32(inner_out)
            if ips is not None and ips.size != K * n:
                raise ValueError("inner_ps must hold K*n fold ids")
            bidx = None
            if boot_idx is not None:
                bidx = np.ascontiguousarray(boot_idx, dtype=np.int32)
        nuis = want_nuisances and not o.on_device
        phat = np.empty(n) if nuis else None
        m0 = np.empty(n) if nuis else None
        m1 = np.empty(n) if nuis else None
        lps = np.zeros(K)
        lout = np.zeros(K)
        reps = o.boot_reps_here if o.boot_reps_here > 0 else o.B
        boot = np.empty(max(reps, 1))
        res = CrossfitResult()
        self._chk(self._lib.ate_crossfit_aipw(self._h, C.byref(o), n, p0, xp, wp, yp, L.ptr(fid), L.ptr(ips),
                                              L.ptr(iout), L.ptr(bidx), L.ptr(phat), L.ptr(m0), L.ptr(m1),
                                              L.ptr(lps), L.ptr(lout), L.ptr(boot), C.byref(res)),
                  "ate_crossfit_aipw")
        out = {f: getattr(res, f) for f, _ in CrossfitResult._fields_}
        out.update(phat=phat, m0=m0, m1=m1, lambda_ps=lps, lambda_out=lout, boot=boot[:reps] if o.B > 0 else None)
        return out
