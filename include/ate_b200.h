/*
 * ate_b200.h -- C ABI of libate_b200.so, the B200 (sm_100a) engine for the cross-fitted
 * AIPW / double-ML ATE path of Zoe187419/ATE_replication_causalML.
 *
 * The reference has no FFI of its own: its boundary is the set of R closures that
 * `source("ate_functions.R")` defines (ate_replication.Rmd:25).  Each entry point below
 * replaces the native work behind one of those closures (the .Fortran("elnet") /
 * .Fortran("lognet") cores of glmnet, base R's sample()/mean()/sd(), scale()), and is what a
 * `.Call` shim binds -- see INTEGRATION.md for the R-side stub of every function.
 *
 * Conventions
 *   - plain C: pointers + sizes, no torch/CUDA types.  `long long` row counts, `int` columns.
 *   - matrices are column-major doubles exactly as R holds them (`as.matrix(...)`,
 *     ate_functions.R:94-95,116-117,135-136,299-301); index vectors are 1-based int32 as in R.
 *   - every function returns 0 on success, <0 on error (ate_last_error(ctx) has the text),
 *     >0 for glmnet-style non-fatal conditions (e.g. path stopped early at lambda m => -m is
 *     glmnet's jerr; we return it as +m in `*jerr` fields and 0 as status).
 *   - a ctx is bound to one GPU; one ctx per thread.  The caller owns every host buffer;
 *     the library owns device memory inside the ctx.  Nothing outlives ate_ctx_destroy().
 *   - there is no CPU fallback: without a CUDA device ate_ctx_create() fails.
 *   - `on_device != 0` (where offered) means the data pointers are device pointers already
 *     resident on the ctx's GPU (used by bench.py's `value` leg and by multi-GPU drivers).
 */
#ifndef ATE_B200_H
#define ATE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ate_ctx ate_ctx;

#if defined(__GNUC__)
#define ATE_API __attribute__((visibility("default")))
#else
#define ATE_API
#endif

#define ATE_FAMILY_GAUSSIAN 0
#define ATE_FAMILY_BINOMIAL 1

#define ATE_BOOT_INDEX 0  /* validation mode: caller supplies the B x n index matrix        */
#define ATE_BOOT_PHILOX 1 /* production mode: indices drawn on device, Philox4x32-10       */

#define ATE_RNG_ROUNDING 0  /* R <= 3.5 sample(): floor(n * unif_rand())                    */
#define ATE_RNG_REJECTION 1 /* R >= 3.6 sample(): rejection sampling on ceil(log2 n) bits   */

/* ---- context ------------------------------------------------------------------------- */
ATE_API int ate_version(void);
/* No reference counterpart (R's GC owns everything there).  device = CUDA ordinal. */
ATE_API int ate_ctx_create(int device, ate_ctx **out);
ATE_API void ate_ctx_destroy(ate_ctx *ctx);
ATE_API const char *ate_last_error(const ate_ctx *ctx);
/* The ctx's cudaStream_t (as void*), so a caller can record its own CUDA events on the stream
 * the kernels are launched on. */
ATE_API void *ate_ctx_stream(const ate_ctx *ctx);
/* Number of kernels this ctx has launched so far (bench.py's `gpu_launches`). */
ATE_API long long ate_launch_count(const ate_ctx *ctx);
/* Device time of the hot kernels accumulated since the last reset, from CUDA events recorded on the
 * ctx stream around each launch: total ms, launch count, algorithmic flop and bytes.  `on`: 0 off
 * (default), 1 the IRLS step kernel (irls_round_kernel), 2 the full-width Gram kernel
 * (gram_dmma_kernel), 3 both. */
ATE_API int ate_profile_enable(ate_ctx *ctx, int on);
ATE_API int ate_profile_read(ate_ctx *ctx, double *gram_ms, long long *gram_launches, double *gram_flops,
                     double *gram_bytes);

/* Multi-GPU (no reference counterpart: the reference is single-threaded).  When one fit is shared
 * by several ranks (ate_crossfit_opts.shard_count > 1; one process and one ctx per GPU) every rank
 * keeps a slice of the ROWS, and per-round partial sums -- the strong-set Grams, X^T r, X^T v and the
 * deviances of every live path fit, the moment matrices of the gaussian fits, the held-out deviances --
 * are summed over the ranks; the nuisances and the bootstrap replicates are all-gathered.
 *
 * ate_ctx_nccl_init: the library opens libnccl.so.2 itself and keeps an NCCL communicator in the
 * ctx; every collective is then enqueued on the ctx stream between the kernels that produce and
 * consume it (no host round trip).  Rank 0 obtains the 128-byte id with ate_nccl_unique_id and
 * distributes it by whatever means the host program has (bench.py: torch.distributed broadcast; an R
 * driver: Rmpi or a file), then every rank calls ate_ctx_nccl_init -- a collective call itself.
 *
 * ate_ctx_set_allreduce: the fallback when no communicator was created: the library calls `fn` to sum
 * a HOST buffer of doubles over all ranks (one device->host->device round trip per collective).  Must
 * return 0 on success.  Used by CPU-side tests with gloo and by hosts that only have MPI. */
ATE_API int ate_nccl_unique_id(char *id128);
ATE_API int ate_ctx_nccl_init(ate_ctx *ctx, const char *id128, int rank, int nranks);
ATE_API int ate_ctx_nccl_destroy(ate_ctx *ctx);
typedef int (*ate_allreduce_fn)(void *user, double *host_buf, long long count);
ATE_API int ate_ctx_set_allreduce(ate_ctx *ctx, ate_allreduce_fn fn, void *user);

/* ---- a1: scale()  (ate_replication.Rmd:72-74) ----------------------------------------- */
/* In place on x[n*p]: columns `cols` (0-based; NULL = all) get (x-mean)/sd, sd with n-1.
 * center/scale (length ncols) receive attr(,"scaled:center") / attr(,"scaled:scale"). */
ATE_API int ate_standardize(ate_ctx *ctx, double *x, long long n, int p, const int *cols, int ncols,
                    double *center, double *scale);

/* ---- a2: degree-2 expansion  (ate_functions.R:288-296) -------------------------------- */
/* out[n * (p0*p0 + (include_main ? p0 : 0))], mains first then c1-major ordered products. */
ATE_API int ate_expand_deg2(ate_ctx *ctx, const double *x, long long n, int p0, int include_main,
                    double *out);

/* ---- a3/a4: glmnet(x, y, family, alpha = 1, ...)  ------------------------------------- */
/* Replaces .Fortran("elnet", ...) (ate_functions.R:101-103,123-125,304-305) and
 * .Fortran("lognet", ...) (ate_functions.R:139-141) together with glmnet()'s R-level
 * defaults (nlambda 100, lambda.min.ratio 1e-4|1e-2, standardize, intercept, fix.lam).
 * lambda_in == NULL: automatic sequence.  penalty_factor NULL = ones.  weights are 1.
 * Outputs: *lmu (<= nlambda), lambda[nlambda], a0[nlambda], beta[p*nlambda] (dense, column l
 * = coefficients at lambda[l] on the caller's x scale), dev_ratio[nlambda], *npasses,
 * *nulldev, *jerr (glmnet's jerr). */
ATE_API int ate_glmnet(ate_ctx *ctx, int family, long long n, int p, const double *x, const double *y,
               const double *penalty_factor, int nlambda, double lambda_min_ratio,
               const double *lambda_in, int nlambda_in, double thresh, int maxit, int *lmu,
               double *lambda, double *a0, double *beta, double *dev_ratio, int *npasses,
               double *nulldev, int *jerr);

/* ---- cv.glmnet(x, y, alpha = 1, family, penalty.factor)  ------------------------------ */
/* The wrapper the reference actually calls (same lines).  foldid[n] in 1..nfolds is what
 * cv.glmnet would draw with sample(rep(seq(nfolds), length = N)); pass it explicitly
 * (validation mode) or NULL to have it drawn with R's Mersenne-Twister (rng_kind = ATE_RNG_*):
 * seed != 0 does set.seed(seed) first; seed == 0 continues the ctx's stream (ate_r_set_seed),
 * the way consecutive cv.glmnet() calls share one stream in an R session.
 * type.measure = "deviance", grouped = TRUE.
 * idx_min / idx_1se: 0-based positions of lambda.min / lambda.1se. */
ATE_API int ate_cv_glmnet(ate_ctx *ctx, int family, long long n, int p, const double *x, const double *y,
                  const double *penalty_factor, int nfolds, const int32_t *foldid, uint32_t seed,
                  int rng_kind, int nlambda, double thresh, int maxit, int *lmu, double *lambda,
                  double *a0, double *beta, double *cvm, double *cvsd, int *idx_min, int *idx_1se,
                  double *dev_ratio, int *jerr);

/* The same cv.glmnet on data that already sits in HBM (the boundary is still ate_functions.R:101-103,
 * :123-125, :139-141, :304-305; the host-matrix form above is what an R shim binds).  X_dev: THIS rank's
 * rows of the library's row-major design, one record of ld doubles per observation laid out
 * [x_1 .. x_p | y | 1 | 0 ...] (ld even, >= p + 2; 32-byte multiples stage fastest); foldid_dev: the
 * fold ids of ALL n rows (device int32, the same on every rank).  shard_count <= 1: X_dev holds all
 * n rows.  shard_count > 1: rank r holds rows [r c, min(n, (r + 1) c)), c = ceil(n / count) rounded up
 * to a multiple of 4; the per-fold moment matrices (gaussian) or the per-round partial sums (binomial)
 * are summed over the ranks inside the call and every rank returns the same model.  penalty_factor and
 * all outputs are host arrays as in ate_cv_glmnet. */
ATE_API int ate_cv_glmnet_device(ate_ctx *ctx, int family, long long n, int p, const double *X_dev, long long ld,
                         const int32_t *foldid_dev, int nfolds, const double *penalty_factor, int nlambda,
                         double thresh, int maxit, int shard_rank, int shard_count, int *lmu, double *lambda,
                         double *a0, double *beta, double *cvm, double *cvsd, int *idx_min, int *idx_1se,
                         double *dev_ratio, int *jerr);

/* ---- f2: glm(formula, family = binomial("logit"))  (ate_functions.R:218-220, :231-234) ----- */
/* The unpenalised logistic nuisance fits of doubly_robust_glm (and :156-158, Rmd:165-168):
 * glm.fit's IRLS from mustart = (y + 0.5)/2, convergence |dev - devold| / (|dev| + 0.1) < epsilon
 * (1e-8), at most maxit (25) steps.  x[n*p] has no intercept column; coef[p+1] comes back in R's
 * order, (Intercept) first.  maxit <= 0 / epsilon <= 0 select glm.control()'s defaults. */
ATE_API int ate_glm_binomial(ate_ctx *ctx, long long n, int p, const double *x, const double *y, int maxit,
                             double epsilon, double *coef, double *deviance, int *iters, int *converged);

/* ---- f3/f4: lm() + summary(lm)$coefficients  (ate_functions.R:25-34 ate_condmean_ols; :52-56
 * prop_score_weight's residual regression; :72-81 prop_score_ols, weighted; :320-323 belloni's
 * post-lasso OLS) ------------------------------------------------------------------------ */
/* Least squares of y on [1, x] (x: n x p col-major host matrix WITHOUT the intercept column;
 * weights: n or NULL, as lm(weights=)).  coef / std_err [p+1] in R's order, (Intercept) first; a column
 * that R's pivoted QR (tol 1e-7) would report as aliased gets NaN in both (R prints NA and
 * summary() drops the row).  sigma = residual standard error = sqrt(rss / df_residual), rss =
 * sum w e^2, rank = number of estimable coefficients, df_residual = #(w != 0) - rank.  resid
 * (optional, host, n): y - fitted (unweighted residuals, residuals(lm)).  Any output but coef may
 * be NULL.  Returns -1 on bad arguments / negative weights. */
ATE_API int ate_lm(ate_ctx *ctx, long long n, int p, const double *x, const double *y, const double *weights,
                   double *coef, double *std_err, double *sigma, double *rss, int *rank,
                   long long *df_residual, double *resid);

/* ---- a5: predict(model, newx, type)  (ate_functions.R:144; coef() at :106,:128,:308) --- */
/* out[n] = a0 + x beta, passed through plogis when response_binomial != 0. */
ATE_API int ate_predict(ate_ctx *ctx, long long n, int p, const double *x, double a0, const double *beta,
                int response_binomial, double *out);

/* ---- a7/a8: AIPW score, ATE, sandwich SE  (ate_functions.R:241-243, :255-256) ---------- */
/* tau = mean(est1, na.rm = TRUE) + mean(est2); se = sqrt(sum(I^2)) / n.  est1/est2 (length n)
 * are optional outputs (NULL to skip).  on_device: pointers are device pointers. */
ATE_API int ate_aipw_score(ate_ctx *ctx, long long n, const double *w, const double *y, const double *p,
                   const double *m0, const double *m1, double *tau, double *se, double *est1,
                   double *est2, int on_device);

/* ---- the same score without any device allocation --------------------------------------- */
/* No counterpart in the reference.  For callers that cannot call cudaMalloc: code running inside
 * a memory pool that already owns the whole GPU, or on a latency-critical path where the device
 * synchronisation hidden in cudaMalloc/cudaFree is not acceptable.  ate_aipw_score() keeps its
 * per-block partial sums in scratch that the ctx grows with cudaMalloc; this variant keeps them in
 * a fixed 48 KB array inside the library (resident on every GPU that loads it, until exit) and
 * calls no allocating routine.  All five inputs are DEVICE pointers of length n; tau and se are
 * host outputs.  Limits and rules:
 *   - n <= 2^28 (268,435,456) rows; larger n is refused with an error, nothing is launched;
 *   - every call on a device shares the one array and finds the previous call's contents in it
 *     (the entry point clears what must start at zero); calls that overlap on the same GPU race
 *     and must be serialised by the caller;
 *   - the first call on each device resolves the array's device address: make that first call
 *     before any stream capture begins. */
ATE_API int ate_aipw_score_static(ate_ctx *ctx, long long n, const double *d_w, const double *d_y,
                                  const double *d_p, const double *d_m0, const double *d_m1,
                                  double *tau, double *se);

/* ---- a9/a10: bootstrap SE with nuisances held fixed  (ate_functions.R:245-252,267-283) - */
/* mode ATE_BOOT_INDEX: idx[B*n] (replicate-major, 1-based) is R's sample(n, n, replace = T)
 * per replicate.  mode ATE_BOOT_PHILOX: indices are drawn on device; replicate b uses the
 * Philox stream keyed by (seed, rep_offset + b), so a replicate's value does not depend on how
 * replicates are sharded over GPUs.  boot[B] receives the replicate estimates and *se their
 * sd (n-1 denominator) -- pass se = NULL when sharding and combine boot[] yourself.  Index mode checks
 * every index: one outside 1..n (R's NA_integer_ included) fails the call.  A host index matrix is
 * streamed to the GPU in blocks of <= 256 MB (it is never resident as a whole); with on_device the
 * five vectors AND idx are device pointers. */
ATE_API int ate_bootstrap(ate_ctx *ctx, long long n, const double *w, const double *y, const double *p,
                  const double *m0, const double *m1, int B, int mode, const int32_t *idx,
                  uint64_t seed, long long rep_offset, double *boot, double *se, int on_device);

/* One replicate, the literal tau_hat_dr_est(w, y, p, tauhat0x, tauhat1x) of :267-283 with the
 * index vector that sample() would have produced. */
ATE_API int ate_tau_hat_dr_est(ate_ctx *ctx, long long n, const double *w, const double *y,
                       const double *p, const double *m0, const double *m1, const int32_t *idx,
                       double *tau);

/* ---- f1: R's sample() on the host, bit-exact with set.seed(seed)  (ate_functions.R:269) - */
/* state lives in the ctx.  ate_r_sample: k draws of sample(n, k, replace = TRUE) (1-based).
 * ate_r_foldid: sample(rep(seq(nfolds), length = n)). */
ATE_API int ate_r_set_seed(ate_ctx *ctx, uint32_t seed);
ATE_API int ate_r_sample(ate_ctx *ctx, int n, long long k, int rng_kind, int32_t *out);
ATE_API int ate_r_foldid(ate_ctx *ctx, int n, int nfolds, int rng_kind, int32_t *out);

/* ---- the composite: K-fold cross-fitted AIPW with glmnet nuisances --- */
typedef struct ate_crossfit_opts {
    int K;              /* outer folds (default 5)                                            */
    int nfolds_inner;   /* cv.glmnet nfolds inside each training set (default 10)             */
    int nlambda;        /* default 100                                                        */
    double thresh;      /* default 1e-7                                                       */
    int maxit;          /* default 100000                                                     */
    int use_1se;        /* 1: lambda.1se (the reference's default `s`), 0: lambda.min         */
    int standardize_base; /* 1: scale() the first n_continuous base columns first (a1)       */
    int n_continuous;
    int expand;         /* 0: design = base columns; 1: products only (p0^2);                 */
                        /* 2: mains + products (what belloni() literally builds)              */
    int B;              /* bootstrap replicates (0 = none)                                    */
    int boot_mode;      /* ATE_BOOT_*                                                         */
    uint64_t boot_seed;
    long long boot_rep_offset; /* first replicate id handled by this call (GPU sharding)      */
    int boot_reps_here; /* replicates computed by this call (<= B; 0 = all B)                 */
    uint32_t r_seed;    /* seed for fold ids when foldid / inner ids are NULL                 */
    int rng_kind;       /* ATE_RNG_*                                                          */
    int on_device;      /* every array argument (x_base, w, y, the fold ids, boot_idx and the  */
                        /*   nuisance outputs) is a device pointer                             */
    int fold_begin;     /* outer folds [fold_begin, fold_end) fitted by this call (1-based,   */
    int fold_end;       /*   0,0 = all); rows of other folds get 0 nuisances (shards can be summed) */
    int skip_score;     /* 1: stop after the nuisances (multi-GPU drivers combine first)      */
    int shard_rank;     /* this process's rank and the number of ranks sharing ONE fit (0,0 = */
    int shard_count;    /*   none): rank r keeps rows [r c, (r + 1) c), c = ceil(n / count)   */
} ate_crossfit_opts;

ATE_API void ate_crossfit_default_opts(ate_crossfit_opts *o);

typedef struct ate_crossfit_result {
    double tau, se_sandwich, se_bootstrap, lower_ci, upper_ci; /* a11: tau +- 1.96 se          */
    int jerr;              /* worst glmnet jerr met                                           */
    int rounds_irls;       /* lock-step IRLS rounds executed                                  */
    long long gram_launches;
    double ms_design, ms_outcome, ms_propensity, ms_score, ms_bootstrap; /* host wall, synced */
    double irls_solves;     /* total IRLS quadratic solves over all propensity path fits          */
    double strong_mean;     /* mean size of glmnet's strong set over those solves                 */
    double strong_frac2;    /* mean of (|strong| / p)^2: the share of Gram flops a solve needs    */
    double ms_folds;        /* building the fold / row lists (device)                              */
    double step_launches;   /* launches of the IRLS step kernel (one per lock-step round)          */
    double row_fits;        /* sum over those launches and their live fits of the rows a fit saw    */
    double rows_staged;     /* sum over those launches and their CTAs of the rows staged            */
    double list_misses;     /* steps repeated because a column joined the strong set unpredicted    */
} ate_crossfit_result;

/* x_base[n*p0] col-major, w[n], y[n].  foldid[n] (1..K) and inner_ps / inner_out [K*n]
 * (row k: cv.glmnet fold ids of outer fold k's training rows in training-row order) may be
 * NULL => drawn with R's RNG in the order R would (k = 1..K: propensity cv.glmnet then outcome
 * cv.glmnet): whenever any ids are drawn the call does set.seed(opts->r_seed) first, so two
 * identical calls draw identical folds.  boot_idx: B*n for ATE_BOOT_INDEX (replicate-major), else
 * NULL.  Outputs (may be NULL): phat, m0, m1 [n] stitched out-of-fold nuisances (host, or device
 * with on_device); lambda_ps, lambda_out [K] and boot [B] (always host).
 * With shard_count > 1 (ate_ctx_nccl_init or ate_ctx_set_allreduce first) every rank passes the SAME
 * full inputs, keeps its slice of the rows on its GPU, and receives the complete result. */
ATE_API int ate_crossfit_aipw(ate_ctx *ctx, const ate_crossfit_opts *opts, long long n, int p0,
                      const double *x_base, const double *w, const double *y,
                      const int32_t *foldid, const int32_t *inner_ps, const int32_t *inner_out,
                      const int32_t *boot_idx, double *phat, double *m0, double *m1,
                      double *lambda_ps, double *lambda_out, double *boot,
                      ate_crossfit_result *res);

/* ---- weighted Gram, exposed for tests, roofline measurement and multi-GPU drivers ------ */
/* x: n x p col-major host matrix (or, with on_device, the library's row-major device layout:
 * n rows of ld doubles, ld % 2 == 0).  G[p*p] = sum_i v_i x_i x_i^T (v NULL = 1), both
 * triangles filled.  Rows are taken from rows[nrows] (0-based; NULL = all n rows). */
ATE_API int ate_weighted_gram(ate_ctx *ctx, long long n, int p, const double *x, long long ld,
                      const double *v, const int32_t *rows, long long nrows, double *G,
                      int on_device);

/* ---- cv.glmnet's summary of per-fold errors (cvstats + getmin; behind ate_functions.R:101-103,
 * :123-125, :139-141, :304-305 and the `lambda.min` / `lambda.1se` reads at :106,128,144,308) ------- */
/* Host only.  cvraw[nfolds x lmu] (row per fold): the fold's mean held-out error at each lambda;
 * fold_weights[nfolds]: held-out observation counts; fold_nlam[nfolds] (or NULL = lmu everywhere):
 * how many lambdas each fold's path reached (folds that stopped early drop out of the later
 * columns).  Writes cvm[lmu] (weighted mean), cvsd[lmu] (sqrt(weighted var / (folds - 1))) and the
 * 0-based indices of lambda.min (largest lambda attaining the minimum) and lambda.1se (largest
 * lambda with cvm <= cvm[min] + cvsd[min]).  ate_cv_glmnet does this internally; the entry point is
 * for callers that hold their own fold errors. */
ATE_API int ate_cv_select(int nfolds, int lmu, const double *cvraw, const double *fold_weights,
                          const int *fold_nlam, const double *lambda, double *cvm, double *cvsd,
                          int *idx_min, int *idx_1se);

/* ---- introspection of the IRLS step kernel's work plan (host only; no device call) ------------- */
/* No counterpart in the reference.  For `nfit` path fits of ONE cv.glmnet problem whose Gram column
 * lists hold list_cols[f] columns (the strong set plus candidates), on a design that stages
 * `staged_cols` columns per row, returns how the step kernel groups them into CTAs: the return value is
 * the number of CTA kinds G (<= max_groups, else -2), and desc[24 g + ...] describes kind g:
 *   [0] fits served (1 to 4), [1 .. 4] their indices, [5] bit s set when this kind evaluates the sums
 *   of its fit s (a list too wide for one CTA is split over several kinds; the first one leads),
 *   [8 + 2 w], [9 + 2 w] for MMA warp w = 0..7: its two halves, each fit slot | segment << 2 |
 *   count << 8 | row0 << 10 | row1 << 18 (a half = up to two tile rows of one 32-column segment of one
 *   fit; count 0: none).
 * Lets tests check on a CPU-only box that every 8x8 tile of every fit's lower triangle is computed
 * exactly once for every mix of widths.  -1 for widths the step cannot take. */
ATE_API int ate_round_plan_describe(int nfit, const int *list_cols, int staged_cols, int max_groups, unsigned *desc);

/* ---- the step kernel alone, on synthetic rows (measurement; no counterpart in the reference) ----- */
/* Times `iters` launches of irls_round_kernel (with its fold kernels) on nprob synthetic problems of
 * nrows rows x ni columns, each with fits_per_prob path fits whose Gram lists hold list_cols scattered
 * columns (fits_per_prob > 1: every fit gives its own 1/fits_per_prob of the rows weight 0, as the
 * cv-fold fits do).  private_copy != 0: the contiguous row layout the binomial driver stages from.
 * skip_mask idles roles of the kernel (1 MMA, 2 helpers, 4 eval's exp / log, 8 eval's dot) to see what
 * paces it; results are then meaningless.  Outputs: ms per launch (CUDA events on the ctx stream), the
 * algorithmic flop and bytes of one launch and the CTAs launched.  tools/round_bench.py. */
ATE_API int ate_round_kernel_bench(ate_ctx *ctx, long long nrows, int ni, int nprob, int fits_per_prob, int list_cols,
                           int private_copy, int iters, int skip_mask, double *ms_per_launch, double *flops,
                           double *alg_bytes, int *ctas);

#ifdef __cplusplus
}
#endif
#endif /* ATE_B200_H */
