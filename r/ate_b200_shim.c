/*
 * ate_b200_shim.c -- the `.Call` shim a maintainer of ate_functions.R would add.
 *
 * NOT COMPILED IN THIS IMAGE (no R, no Rinternals.h).  Pure marshalling: every wrapper turns
 * SEXPs into the pointers + sizes of include/ate_b200.h and back.  Build inside an R package with
 *     PKG_LIBS = -L<repo>/ate_replication_causalml_b200 -late_b200
 * Rules followed: R objects are touched only from the calling thread;
 * Rf_error() (a longjmp) is raised only after the library call has returned, so no CUDA/C++
 * object is live across it; the ctx lives in an external pointer with a finalizer.
 */
#include <R.h>
#include <Rinternals.h>
#include <R_ext/Rdynload.h>
#include "ate_b200.h"

static void ctx_finalizer(SEXP p) {
    ate_ctx *c = (ate_ctx *)R_ExternalPtrAddr(p);
    if (c) ate_ctx_destroy(c);
    R_ClearExternalPtr(p);
}
static ate_ctx *ctx_of(SEXP p) {
    ate_ctx *c = (ate_ctx *)R_ExternalPtrAddr(p);
    if (!c) Rf_error("ate_b200: context was destroyed");
    return c;
}
#define FAIL_IF(rc, c, what) do { if ((rc) < 0) Rf_error("%s: %s", what, ate_last_error(c)); } while (0)

SEXP C_ate_ctx_create(SEXP device) {
    ate_ctx *c = NULL;
    int rc = ate_ctx_create(Rf_asInteger(device), &c);
    if (rc != 0) Rf_error("ate_b200: no usable CUDA device (%s); there is no CPU fallback", ate_last_error(NULL));
    SEXP p = PROTECT(R_MakeExternalPtr(c, R_NilValue, R_NilValue));
    R_RegisterCFinalizerEx(p, ctx_finalizer, TRUE);
    UNPROTECT(1);
    return p;
}

/* cv.glmnet(x, y, alpha = 1, family, penalty.factor)  -- ate_functions.R:101-103,123-125,139-141,304-305 */
SEXP C_ate_cv_glmnet(SEXP ctx, SEXP family, SEXP x, SEXP y, SEXP pf, SEXP nfolds, SEXP foldid, SEXP seed) {
    ate_ctx *c = ctx_of(ctx);
    const int n = Rf_nrows(x), p = Rf_ncols(x), nlam = 100;
    SEXP lambda = PROTECT(Rf_allocVector(REALSXP, nlam)), a0 = PROTECT(Rf_allocVector(REALSXP, nlam));
    SEXP beta = PROTECT(Rf_allocMatrix(REALSXP, p, nlam)), cvm = PROTECT(Rf_allocVector(REALSXP, nlam));
    SEXP cvsd = PROTECT(Rf_allocVector(REALSXP, nlam)), dev = PROTECT(Rf_allocVector(REALSXP, nlam));
    int lmu = 0, imin = 0, i1se = 0, jerr = 0;
    int rc = ate_cv_glmnet(c, Rf_asInteger(family), n, p, REAL(x), REAL(y),
                           Rf_isNull(pf) ? NULL : REAL(pf), Rf_asInteger(nfolds),
                           Rf_isNull(foldid) ? NULL : INTEGER(foldid), (uint32_t)Rf_asInteger(seed),
                           ATE_RNG_REJECTION, nlam, 1e-7, 100000, &lmu, REAL(lambda), REAL(a0), REAL(beta),
                           REAL(cvm), REAL(cvsd), &imin, &i1se, REAL(dev), &jerr);
    FAIL_IF(rc, c, "ate_cv_glmnet");
    const char *nm[] = {"lambda", "a0", "beta", "cvm", "cvsd", "dev.ratio", "lmu", "idx.min", "idx.1se", "jerr", ""};
    SEXP out = PROTECT(Rf_mkNamed(VECSXP, nm));
    SET_VECTOR_ELT(out, 0, lambda); SET_VECTOR_ELT(out, 1, a0); SET_VECTOR_ELT(out, 2, beta);
    SET_VECTOR_ELT(out, 3, cvm); SET_VECTOR_ELT(out, 4, cvsd); SET_VECTOR_ELT(out, 5, dev);
    SET_VECTOR_ELT(out, 6, Rf_ScalarInteger(lmu)); SET_VECTOR_ELT(out, 7, Rf_ScalarInteger(imin + 1));
    SET_VECTOR_ELT(out, 8, Rf_ScalarInteger(i1se + 1)); SET_VECTOR_ELT(out, 9, Rf_ScalarInteger(jerr));
    UNPROTECT(7);
    return out;
}

/* predict(model, newx, type)  -- ate_functions.R:144 */
SEXP C_ate_predict(SEXP ctx, SEXP x, SEXP a0, SEXP beta, SEXP response) {
    ate_ctx *c = ctx_of(ctx);
    const int n = Rf_nrows(x), p = Rf_ncols(x);
    SEXP out = PROTECT(Rf_allocMatrix(REALSXP, n, 1));
    int rc = ate_predict(c, n, p, REAL(x), Rf_asReal(a0), REAL(beta), Rf_asLogical(response), REAL(out));
    FAIL_IF(rc, c, "ate_predict");
    UNPROTECT(1);
    return out;
}

/* est1/est2/tau_hat and the sandwich se  -- ate_functions.R:241-243, 255-256 */
SEXP C_ate_aipw_score(SEXP ctx, SEXP w, SEXP y, SEXP p, SEXP m0, SEXP m1) {
    ate_ctx *c = ctx_of(ctx);
    double tau, se;
    int rc = ate_aipw_score(c, XLENGTH(w), REAL(w), REAL(y), REAL(p), REAL(m0), REAL(m1), &tau, &se, NULL, NULL, 0);
    FAIL_IF(rc, c, "ate_aipw_score");
    SEXP out = PROTECT(Rf_allocVector(REALSXP, 2));
    REAL(out)[0] = tau; REAL(out)[1] = se;
    UNPROTECT(1);
    return out;
}

/* glm(y ~ ., family = binomial(link = "logit"))  -- ate_functions.R:218-220, 231-234 (also :25-27, :160-162).
 * x is model.matrix(...) without the "(Intercept)" column; returns coef with (Intercept) first. */
SEXP C_ate_glm_binomial(SEXP ctx, SEXP x, SEXP y) {
    ate_ctx *c = ctx_of(ctx);
    const int n = Rf_nrows(x), p = Rf_ncols(x);
    SEXP coef = PROTECT(Rf_allocVector(REALSXP, p + 1));
    double dev;
    int iters, conv;
    int rc = ate_glm_binomial(c, n, p, REAL(x), REAL(y), 25, 1e-8, REAL(coef), &dev, &iters, &conv);
    FAIL_IF(rc, c, "ate_glm_binomial");
    if (!conv) Rf_warning("glm.fit: algorithm did not converge");
    Rf_setAttrib(coef, Rf_install("deviance"), Rf_ScalarReal(dev));
    Rf_setAttrib(coef, Rf_install("iter"), Rf_ScalarInteger(iters));
    UNPROTECT(1);
    return coef;
}

/* the bootstrap loop  -- ate_functions.R:245-252; idx = NULL -> on-device Philox */
SEXP C_ate_bootstrap(SEXP ctx, SEXP w, SEXP y, SEXP p, SEXP m0, SEXP m1, SEXP B, SEXP idx, SEXP seed) {
    ate_ctx *c = ctx_of(ctx);
    const int b = Rf_asInteger(B);
    SEXP boot = PROTECT(Rf_allocVector(REALSXP, b));
    double se;
    int rc = ate_bootstrap(c, XLENGTH(w), REAL(w), REAL(y), REAL(p), REAL(m0), REAL(m1), b,
                           Rf_isNull(idx) ? ATE_BOOT_PHILOX : ATE_BOOT_INDEX, Rf_isNull(idx) ? NULL : INTEGER(idx),
                           (uint64_t)Rf_asReal(seed), 0, REAL(boot), &se, 0);
    FAIL_IF(rc, c, "ate_bootstrap");
    Rf_setAttrib(boot, Rf_install("sd"), Rf_ScalarReal(se));
    UNPROTECT(1);
    return boot;
}

/* tau_hat_dr_est(w, y, p, tauhat0x, tauhat1x) with the index vector sample() produced  -- :267-283 */
SEXP C_ate_tau_hat_dr_est(SEXP ctx, SEXP w, SEXP y, SEXP p, SEXP m0, SEXP m1, SEXP idx) {
    ate_ctx *c = ctx_of(ctx);
    double tau;
    int rc = ate_tau_hat_dr_est(c, XLENGTH(w), REAL(w), REAL(y), REAL(p), REAL(m0), REAL(m1), INTEGER(idx), &tau);
    FAIL_IF(rc, c, "ate_tau_hat_dr_est");
    return Rf_ScalarReal(tau);
}

/* scale()  -- ate_replication.Rmd:72-74 ; belloni's expansion -- ate_functions.R:288-296 */
SEXP C_ate_standardize(SEXP ctx, SEXP x) {
    ate_ctx *c = ctx_of(ctx);
    SEXP out = PROTECT(Rf_duplicate(x));
    int rc = ate_standardize(c, REAL(out), Rf_nrows(x), Rf_ncols(x), NULL, 0, NULL, NULL);
    FAIL_IF(rc, c, "ate_standardize");
    UNPROTECT(1);
    return out;
}
SEXP C_ate_expand_deg2(SEXP ctx, SEXP x) {
    ate_ctx *c = ctx_of(ctx);
    const int n = Rf_nrows(x), p0 = Rf_ncols(x);
    SEXP out = PROTECT(Rf_allocMatrix(REALSXP, n, p0 + p0 * p0));
    int rc = ate_expand_deg2(c, REAL(x), n, p0, 1, REAL(out));
    FAIL_IF(rc, c, "ate_expand_deg2");
    UNPROTECT(1);
    return out;
}

/* the composite: K-fold cross-fitted AIPW + bootstrap   */
/* shard: NULL, or c(rank, count) when the rows of this fit are shared by `count` R processes (one per
 * GPU, each after C_ate_ctx_nccl_init); every process passes the same full inputs and gets the full row */
SEXP C_ate_crossfit_aipw(SEXP ctx, SEXP x, SEXP w, SEXP y, SEXP K, SEXP B, SEXP seed, SEXP foldid,
                         SEXP inner_ps, SEXP inner_out, SEXP boot_idx, SEXP expand, SEXP ncont, SEXP shard) {
    ate_ctx *c = ctx_of(ctx);
    ate_crossfit_opts o;
    ate_crossfit_default_opts(&o);
    o.K = Rf_asInteger(K); o.B = Rf_asInteger(B);
    o.r_seed = (uint32_t)Rf_asInteger(seed); o.boot_seed = (uint64_t)Rf_asInteger(seed);
    o.expand = Rf_asInteger(expand); o.n_continuous = Rf_asInteger(ncont); o.standardize_base = o.n_continuous > 0;
    o.boot_mode = Rf_isNull(boot_idx) ? ATE_BOOT_PHILOX : ATE_BOOT_INDEX;
    if (!Rf_isNull(shard)) { o.shard_rank = INTEGER(shard)[0]; o.shard_count = INTEGER(shard)[1]; }
    ate_crossfit_result r;
    int rc = ate_crossfit_aipw(c, &o, Rf_nrows(x), Rf_ncols(x), REAL(x), REAL(w), REAL(y),
                               Rf_isNull(foldid) ? NULL : INTEGER(foldid),
                               Rf_isNull(inner_ps) ? NULL : INTEGER(inner_ps),
                               Rf_isNull(inner_out) ? NULL : INTEGER(inner_out),
                               Rf_isNull(boot_idx) ? NULL : INTEGER(boot_idx),
                               NULL, NULL, NULL, NULL, NULL, NULL, &r);
    FAIL_IF(rc, c, "ate_crossfit_aipw");
    SEXP out = PROTECT(Rf_allocVector(REALSXP, 5));
    REAL(out)[0] = r.tau; REAL(out)[1] = r.se_sandwich; REAL(out)[2] = r.se_bootstrap;
    REAL(out)[3] = r.lower_ci; REAL(out)[4] = r.upper_ci;
    UNPROTECT(1);
    return out;
}

/* lm() + summary(lm)$coefficients  -- ate_functions.R:25-34, 52-56, 72-81, 320-323.
 * x: model matrix without "(Intercept)"; weights: numeric or NULL.  Returns a (p+1) x 2 matrix
 * (Estimate, Std. Error; NA for aliased columns) with attributes sigma, rss, rank, df, residuals. */
SEXP C_ate_lm(SEXP ctx, SEXP x, SEXP y, SEXP weights, SEXP want_resid) {
    ate_ctx *c = ctx_of(ctx);
    const int n = Rf_nrows(x), p = Rf_ncols(x), wr = Rf_asLogical(want_resid);
    SEXP tab = PROTECT(Rf_allocMatrix(REALSXP, p + 1, 2));
    SEXP res = PROTECT(wr ? Rf_allocVector(REALSXP, n) : R_NilValue);
    double sigma, rss; int rank; long long df;
    int rc = ate_lm(c, n, p, REAL(x), REAL(y), Rf_isNull(weights) ? NULL : REAL(weights), REAL(tab),
                    REAL(tab) + (p + 1), &sigma, &rss, &rank, &df, wr ? REAL(res) : NULL);
    FAIL_IF(rc, c, "ate_lm");
    for (int i = 0; i < 2 * (p + 1); ++i) if (ISNAN(REAL(tab)[i])) REAL(tab)[i] = NA_REAL;
    Rf_setAttrib(tab, Rf_install("sigma"), Rf_ScalarReal(sigma));
    Rf_setAttrib(tab, Rf_install("rss"), Rf_ScalarReal(rss));
    Rf_setAttrib(tab, Rf_install("rank"), Rf_ScalarInteger(rank));
    Rf_setAttrib(tab, Rf_install("df"), Rf_ScalarReal((double)df));
    if (wr) Rf_setAttrib(tab, Rf_install("residuals"), res);
    UNPROTECT(2);
    return tab;
}

/* Multi-GPU (no reference counterpart): rank 0 mints the 128-byte NCCL id, the R driver ships it to
 * the other processes (Rmpi::mpi.bcast of the raw vector, or a file), every process then joins. */
SEXP C_ate_nccl_unique_id(void) {
    SEXP id = PROTECT(Rf_allocVector(RAWSXP, 128));
    int rc = ate_nccl_unique_id((char *)RAW(id));
    UNPROTECT(1);
    if (rc != 0) Rf_error("ate_nccl_unique_id: libnccl.so.2 could not be loaded");
    return id;
}
SEXP C_ate_ctx_nccl_init(SEXP ctx, SEXP id, SEXP rank, SEXP nranks) {
    ate_ctx *c = ctx_of(ctx);
    if (XLENGTH(id) != 128) Rf_error("ate_ctx_nccl_init: id must be the 128 raw bytes of C_ate_nccl_unique_id");
    int rc = ate_ctx_nccl_init(c, (const char *)RAW(id), Rf_asInteger(rank), Rf_asInteger(nranks));
    FAIL_IF(rc, c, "ate_ctx_nccl_init");
    return R_NilValue;
}

static const R_CallMethodDef call_methods[] = {
    {"C_ate_ctx_create", (DL_FUNC)&C_ate_ctx_create, 1},
    {"C_ate_cv_glmnet", (DL_FUNC)&C_ate_cv_glmnet, 8},
    {"C_ate_predict", (DL_FUNC)&C_ate_predict, 5},
    {"C_ate_aipw_score", (DL_FUNC)&C_ate_aipw_score, 6},
    {"C_ate_glm_binomial", (DL_FUNC)&C_ate_glm_binomial, 3},
    {"C_ate_lm", (DL_FUNC)&C_ate_lm, 5},
    {"C_ate_bootstrap", (DL_FUNC)&C_ate_bootstrap, 9},
    {"C_ate_tau_hat_dr_est", (DL_FUNC)&C_ate_tau_hat_dr_est, 7},
    {"C_ate_standardize", (DL_FUNC)&C_ate_standardize, 2},
    {"C_ate_expand_deg2", (DL_FUNC)&C_ate_expand_deg2, 2},
    {"C_ate_crossfit_aipw", (DL_FUNC)&C_ate_crossfit_aipw, 14},
    {"C_ate_nccl_unique_id", (DL_FUNC)&C_ate_nccl_unique_id, 0},
    {"C_ate_ctx_nccl_init", (DL_FUNC)&C_ate_ctx_nccl_init, 4},
    {NULL, NULL, 0}};

void R_init_ateb200(DllInfo *dll) {
    R_registerRoutines(dll, NULL, call_methods, NULL, NULL);
    R_useDynamicSymbols(dll, FALSE);
}
