// crossfit.cu -- public glmnet / cv.glmnet / predict entry points and the K-fold cross-fitted
// AIPW composite (steps 1-6 of the coverage contract).
#include <math.h>
#include <string.h>

#include <chrono>

#include <algorithm>

#include "folds.cuh"
#include "path.cuh"

namespace {

inline long long round_up(long long v, long long m) { return (v + m - 1) / m * m; }

struct Timer {
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    double lap_ms() {
        auto t1 = std::chrono::steady_clock::now();
        double ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
        t0 = t1;
        return ms;
    }
};

// rows[] (ascending) with fold ids fid[t] in 1..nseg (fid indexed by position t) -> rows grouped
// by fold (stable), seg offsets
void group_by_fold(const std::vector<int32_t> &rows, const int32_t *fid, int nseg,
                   std::vector<int32_t> &grouped, std::vector<long long> &seg) {
    seg.assign(nseg + 1, 0);
    for (size_t t = 0; t < rows.size(); ++t) seg[fid[t]]++;
    for (int f = 0; f < nseg; ++f) seg[f + 1] += seg[f];
    std::vector<long long> cur(seg.begin(), seg.end() - 1);
    grouped.resize(rows.size());
    for (size_t t = 0; t < rows.size(); ++t) grouped[cur[fid[t] - 1]++] = rows[t];
}

int check_foldid(ate_ctx *ctx, const int32_t *fid, long long n, int nfolds, const char *what) {
    for (long long i = 0; i < n; ++i)
        if (fid[i] < 1 || fid[i] > nfolds) return ate_fail(ctx, -1, "%s: fold id %d outside 1..%d", what, fid[i], nfolds);
    return 0;
}

// design [x | y | 1 | 0-pad] from col-major host x, y
int upload_design(ate_ctx *ctx, long long n, int p, const double *x, const double *y,
                  DevBuf<double> &dX, long long *ld_out) {
    const long long ld = round_up(p + 2, 32);
    DevBuf<double> dx;
    ATE_CUDA(ctx, dx.alloc((size_t)n * (p + 1)));
    ATE_CUDA(ctx, dX.alloc((size_t)n * ld));
    ATE_CUDA(ctx, cudaMemcpyAsync(dx.p, x, sizeof(double) * (size_t)n * p, cudaMemcpyHostToDevice, ctx->stream));
    ATE_CUDA(ctx, cudaMemcpyAsync(dx.p + (size_t)n * p, y, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
    ATE_CUDA(ctx, cudaMemsetAsync(dX.p, 0, sizeof(double) * (size_t)n * ld, ctx->stream));
    ATE_TRY(ate_i_transpose_to_rowmajor(ctx, dx.p, n, p + 1, dX.p, ld, 0));
    ATE_TRY(ate_i_fill_column(ctx, dX.p, n, ld, p + 1, 1.0));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // dx goes out of scope
    *ld_out = ld;
    return 0;
}

void copy_out(const CvOut &o, int p, int cap, int *lmu, double *lambda, double *a0, double *beta,
              double *dev) {
    const int L = o.lmu < cap ? o.lmu : cap;
    if (lmu) *lmu = L;
    for (int l = 0; l < L; ++l) {
        if (lambda) lambda[l] = o.lambda[l];
        if (a0) a0[l] = o.a0[l];
        if (dev) dev[l] = o.dev[l];
    }
    if (beta)
        for (size_t e = 0; e < (size_t)p * L; ++e) beta[e] = o.beta[e];
}

}  // namespace

extern "C" int ate_glmnet(ate_ctx *ctx, int family, long long n, int p, const double *x, const double *y,
                          const double *penalty_factor, int nlambda, double lambda_min_ratio,
                          const double *lambda_in, int nlambda_in, double thresh, int maxit, int *lmu,
                          double *lambda, double *a0, double *beta, double *dev_ratio, int *npasses,
                          double *nulldev, int *jerr) {
    if (!ctx) return -1;
    if (!x || !y || n < 2 || p < 1) return ate_fail(ctx, -1, "ate_glmnet: bad arguments");
    if (family != ATE_FAMILY_GAUSSIAN && family != ATE_FAMILY_BINOMIAL)
        return ate_fail(ctx, -1, "ate_glmnet: unknown family %d", family);
    ATE_CUDA(ctx, cudaSetDevice(ctx->device));
    DevBuf<double> dX;
    long long ld = 0;
    ATE_TRY(upload_design(ctx, n, p, x, y, dX, &ld));
    PathOpts po;
    po.nlam = nlambda > 0 ? nlambda : 100;
    po.thr = thresh > 0 ? thresh : 1e-7;
    po.maxit = maxit > 0 ? maxit : 100000;
    if (lambda_in) {
        if (nlambda_in < 1) return ate_fail(ctx, -1, "ate_glmnet: empty lambda sequence");
        po.flmin = 1.0;
        po.ulam = lambda_in;
        po.nulam = nlambda_in;
    } else {
        po.flmin = lambda_min_ratio > 0 ? lambda_min_ratio : (n < p ? 0.01 : 1.0e-4);
        if (po.flmin >= 1.0) return ate_fail(ctx, -1, "ate_glmnet: lambda.min.ratio should be less than 1");
    }
    std::vector<CvProblem> probs(1);
    probs[0].nt = n;
    probs[0].nseg = 0;
    std::vector<CvOut> outs;
    if (family == ATE_FAMILY_GAUSSIAN)
        ATE_TRY(ate_i_cv_gaussian(ctx, dX.p, ld, p, p, p + 1, penalty_factor, po, probs, outs));
    else
        ATE_TRY(ate_i_cv_binomial(ctx, dX.p, ld, p, p, penalty_factor, po, probs, outs, nullptr));
    const int cap = lambda_in ? nlambda_in : po.nlam;
    copy_out(outs[0], p, cap, lmu, lambda, a0, beta, dev_ratio);
    if (npasses) *npasses = outs[0].nlp;
    if (nulldev) *nulldev = outs[0].nulldev;
    if (jerr) *jerr = outs[0].jerr;
    return 0;
}

extern "C" int ate_cv_glmnet(ate_ctx *ctx, int family, long long n, int p, const double *x,
                             const double *y, const double *penalty_factor, int nfolds,
                             const int32_t *foldid, uint32_t seed, int rng_kind, int nlambda,
                             double thresh, int maxit, int *lmu, double *lambda, double *a0,
                             double *beta, double *cvm, double *cvsd, int *idx_min, int *idx_1se,
                             double *dev_ratio, int *jerr) {
    if (!ctx) return -1;
    if (!x || !y || n < 2 || p < 1) return ate_fail(ctx, -1, "ate_cv_glmnet: bad arguments");
    if (nfolds < 3) return ate_fail(ctx, -1, "ate_cv_glmnet: nfolds must be bigger than 3; nfolds=10 recommended");
    ATE_CUDA(ctx, cudaSetDevice(ctx->device));
    std::vector<int32_t> fid((size_t)n);
    if (foldid) {
        ATE_TRY(check_foldid(ctx, foldid, n, nfolds, "ate_cv_glmnet"));
        fid.assign(foldid, foldid + n);
    } else {
        // seed != 0: set.seed(seed) first; seed == 0: continue the ctx's stream, as consecutive
        // cv.glmnet() calls do in one R session (a stream that was never seeded starts at set.seed(0))
        if (seed != 0 || !ctx->rng.seeded) ate_host_rng_seed(&ctx->rng, seed);
        ate_host_foldid(&ctx->rng, (int)n, nfolds, rng_kind, fid.data());
    }
    DevBuf<double> dX;
    long long ld = 0;
    ATE_TRY(upload_design(ctx, n, p, x, y, dX, &ld));
    std::vector<int32_t> rows((size_t)n), grouped;
    for (long long i = 0; i < n; ++i) rows[i] = (int32_t)i;
    std::vector<CvProblem> probs(1);
    group_by_fold(rows, fid.data(), nfolds, grouped, probs[0].seg);
    DevBuf<int32_t> dlist;
    ATE_CUDA(ctx, dlist.alloc((size_t)n));
    ATE_CUDA(ctx, cudaMemcpyAsync(dlist.p, grouped.data(), sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
    probs[0].d_list = dlist.p;
    probs[0].nt = n;
    probs[0].nseg = nfolds;
    PathOpts po;
    po.nlam = nlambda > 0 ? nlambda : 100;
    po.thr = thresh > 0 ? thresh : 1e-7;
    po.maxit = maxit > 0 ? maxit : 100000;
    po.flmin = n < p ? 0.01 : 1.0e-4;
    std::vector<CvOut> outs;
    if (family == ATE_FAMILY_GAUSSIAN)
        ATE_TRY(ate_i_cv_gaussian(ctx, dX.p, ld, p, p, p + 1, penalty_factor, po, probs, outs));
    else if (family == ATE_FAMILY_BINOMIAL)
        ATE_TRY(ate_i_cv_binomial(ctx, dX.p, ld, p, p, penalty_factor, po, probs, outs, nullptr));
    else
        return ate_fail(ctx, -1, "ate_cv_glmnet: unknown family %d", family);
    const CvOut &o = outs[0];
    copy_out(o, p, po.nlam, lmu, lambda, a0, beta, dev_ratio);
    for (int l = 0; l < o.lmu; ++l) {
        if (cvm) cvm[l] = o.cvm[l];
        if (cvsd) cvsd[l] = o.cvsd[l];
    }
    if (idx_min) *idx_min = o.idx_min;
    if (idx_1se) *idx_1se = o.idx_1se;
    if (jerr) *jerr = o.jerr;
    return 0;
}

// cv.glmnet on a design that is already on the GPU in the library's row-major layout (no reference
// counterpart: the device-resident form of ate_cv_glmnet for drivers that hold their data in HBM, and
// the form ranks sharing a fit use -- each passes ITS rows [r c, (r + 1) c), c = round_up(ceil(n / ranks), 4)).
extern "C" int ate_cv_glmnet_device(ate_ctx *ctx, int family, long long n, int p, const double *X_dev, long long ld,
                                    const int32_t *foldid_dev, int nfolds, const double *penalty_factor, int nlambda,
                                    double thresh, int maxit, int shard_rank, int shard_count, int *lmu, double *lambda,
                                    double *a0, double *beta, double *cvm, double *cvsd, int *idx_min, int *idx_1se,
                                    double *dev_ratio, int *jerr) {
    if (!ctx) return -1;
    if (!X_dev || !foldid_dev || n < 2 || p < 1) return ate_fail(ctx, -1, "ate_cv_glmnet_device: bad arguments");
    if (nfolds < 3 || nfolds > 64) return ate_fail(ctx, -1, "ate_cv_glmnet_device: nfolds must be in 3..64");
    if (ld < p + 2 || (ld & 1)) return ate_fail(ctx, -1, "ate_cv_glmnet_device: ld must be even and >= p + 2 ([x | y | 1])");
    if (family != ATE_FAMILY_GAUSSIAN && family != ATE_FAMILY_BINOMIAL) return ate_fail(ctx, -1, "ate_cv_glmnet_device: unknown family %d", family);
    ATE_CUDA(ctx, cudaSetDevice(ctx->device));
    const bool shared = shard_count > 1;
    if (shared && (shard_rank < 0 || shard_rank >= shard_count)) return ate_fail(ctx, -1, "ate_cv_glmnet_device: bad shard rank");
    if (shared && !ctx->nccl_comm && !ctx->allreduce)
        return ate_fail(ctx, -1, "ate_cv_glmnet_device: a shared fit needs ate_ctx_nccl_init or ate_ctx_set_allreduce first");
    const int G = shared ? shard_count : 1, rank = shared ? shard_rank : 0;
    const long long chunk = ((n + G - 1) / G + 3) / 4 * 4;
    const long long lo = std::min(n, (long long)rank * chunk), hi = std::min(n, lo + chunk), nloc = hi - lo;
    if (nloc < 1) return ate_fail(ctx, -1, "ate_cv_glmnet_device: rank %d holds no rows", rank);
    void *blk = nullptr;
    const size_t hist_ints = ate_i_rank_scratch_ints(n, nfolds);
    ATE_TRY(ate_cached(ctx, "cvdev.state", sizeof(int32_t) * (2 * (size_t)nloc) + sizeof(int) * (hist_ints + 8) + sizeof(long long) * 160 + 1024, &blk));
    long long *dtot = (long long *)blk, *dtot_g = dtot + 80;
    int *bad = (int *)(dtot_g + 80);
    int *blockhist = bad + 8;
    int32_t *pos = (int32_t *)(blockhist + hist_ints), *list = pos + nloc;
    ATE_CUDA(ctx, cudaMemsetAsync(bad, 0, sizeof(int) * 8, ctx->stream));
    ATE_TRY(ate_i_group_by_key(ctx, foldid_dev + lo, nloc, nfolds, pos, blockhist, dtot, bad, list));
    if (shared)
        ATE_TRY(ate_i_rank_by_key(ctx, foldid_dev, n, nfolds, nullptr, blockhist, dtot_g, bad));
    std::vector<long long> ht(160, 0);
    int hbad = 0;
    ATE_CUDA(ctx, cudaMemcpyAsync(ht.data(), dtot, sizeof(long long) * 160, cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaMemcpyAsync(&hbad, bad, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (hbad) return ate_fail(ctx, -1, "ate_cv_glmnet_device: fold id outside 1..%d", nfolds);
    std::vector<CvProblem> probs(1);
    CvProblem &c = probs[0];
    c.d_list = list;
    c.nseg = nfolds;
    c.seg.assign(nfolds + 1, 0);
    c.seg_global.assign(nfolds, 0);
    for (int f = 0; f < nfolds; ++f) {
        c.seg[f + 1] = c.seg[f] + ht[1 + f];
        c.seg_global[f] = shared ? ht[80 + 1 + f] : ht[1 + f];
        if (c.seg_global[f] < 1) return ate_fail(ctx, -1, "ate_cv_glmnet_device: fold %d is empty", f + 1);
    }
    c.nt = c.seg[nfolds];
    c.nt_global = n;
    PathOpts po;
    po.nlam = nlambda > 0 ? nlambda : 100;
    po.thr = thresh > 0 ? thresh : 1e-7;
    po.maxit = maxit > 0 ? maxit : 100000;
    po.flmin = n < p ? 0.01 : 1.0e-4;
    if (shared) {
        po.shard_rank = rank;
        po.shard_count = G;
    }
    std::vector<CvOut> outs;
    if (family == ATE_FAMILY_GAUSSIAN)
        ATE_TRY(ate_i_cv_gaussian(ctx, X_dev, ld, p, p, p + 1, penalty_factor, po, probs, outs));
    else
        ATE_TRY(ate_i_cv_binomial(ctx, X_dev, ld, p, p, penalty_factor, po, probs, outs, nullptr));
    const CvOut &o = outs[0];
    copy_out(o, p, po.nlam, lmu, lambda, a0, beta, dev_ratio);
    for (int l = 0; l < o.lmu; ++l) {
        if (cvm) cvm[l] = o.cvm[l];
        if (cvsd) cvsd[l] = o.cvsd[l];
    }
    if (idx_min) *idx_min = o.idx_min;
    if (idx_1se) *idx_1se = o.idx_1se;
    if (jerr) *jerr = o.jerr;
    return 0;
}

extern "C" int ate_predict(ate_ctx *ctx, long long n, int p, const double *x, double a0,
                           const double *beta, int response_binomial, double *out) {
    if (!ctx) return -1;
    if (!x || !beta || !out || n < 1 || p < 1) return ate_fail(ctx, -1, "ate_predict: bad arguments");
    ATE_CUDA(ctx, cudaSetDevice(ctx->device));
    const long long ld = round_up(p, 2);
    DevBuf<double> dx, dX, db, dout;
    ATE_CUDA(ctx, dx.alloc((size_t)n * p));
    ATE_CUDA(ctx, dX.alloc((size_t)n * ld));
    ATE_CUDA(ctx, db.alloc(p));
    ATE_CUDA(ctx, dout.alloc((size_t)n));
    ATE_CUDA(ctx, cudaMemcpyAsync(dx.p, x, sizeof(double) * (size_t)n * p, cudaMemcpyHostToDevice, ctx->stream));
    ATE_CUDA(ctx, cudaMemcpyAsync(db.p, beta, sizeof(double) * p, cudaMemcpyHostToDevice, ctx->stream));
    ATE_CUDA(ctx, cudaMemsetAsync(dX.p, 0, sizeof(double) * (size_t)n * ld, ctx->stream));
    ATE_TRY(ate_i_transpose_to_rowmajor(ctx, dx.p, n, p, dX.p, ld, 0));
    ATE_TRY(ate_i_predict_rows(ctx, dX.p, ld, p, nullptr, 0, n, db.p, a0, 0.0, response_binomial ? 1 : 0, 0, dout.p, nullptr));
    ATE_CUDA(ctx, cudaMemcpyAsync(out, dout.p, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" void ate_crossfit_default_opts(ate_crossfit_opts *o) {
    if (!o) return;
    o->K = 5;
    o->nfolds_inner = 10;
    o->nlambda = 100;
    o->thresh = 1e-7;
    o->maxit = 100000;
    o->use_1se = 1;
    o->standardize_base = 1;
    o->n_continuous = 0;
    o->expand = 0;
    o->B = 1000;  // the reference hard-codes B <- 1000 (ate_functions.R:190,247)
    o->boot_mode = ATE_BOOT_PHILOX;
    o->boot_seed = 1991;
    o->boot_rep_offset = 0;
    o->boot_reps_here = 0;
    o->r_seed = 1991;
    o->rng_kind = ATE_RNG_REJECTION;
    o->on_device = 0;
    o->fold_begin = 0;
    o->fold_end = 0;
    o->skip_score = 0;
    o->shard_rank = 0;
    o->shard_count = 0;
}

namespace {

// [p0 columns, leading dimension n] host or device -> device, all rows (every rank standardises on the
// moments of ALL rows, as scale() does)
struct Carve {
    char *base = nullptr;
    size_t off = 0;
    template <typename T>
    T *take(size_t count) {
        off = (off + 255) / 256 * 256;
        T *p = base ? reinterpret_cast<T *>(base + off) : nullptr;
        off += sizeof(T) * (count ? count : 1);
        return p;
    }
};

}  // namespace

extern "C" int ate_crossfit_aipw(ate_ctx *ctx, const ate_crossfit_opts *opts, long long n, int p0,
                                 const double *x_base, const double *w, const double *y,
                                 const int32_t *foldid, const int32_t *inner_ps,
                                 const int32_t *inner_out, const int32_t *boot_idx, double *phat,
                                 double *m0, double *m1, double *lambda_ps, double *lambda_out,
                                 double *boot, ate_crossfit_result *res) {
    if (!ctx) return -1;
    if (!opts || !x_base || !w || !y || !res || n < 10 || p0 < 1)
        return ate_fail(ctx, -1, "ate_crossfit_aipw: bad arguments");
    const ate_crossfit_opts &o = *opts;
    if (o.K < 2 || o.K > 64) return ate_fail(ctx, -1, "ate_crossfit_aipw: K must be in 2..64");
    if (o.nfolds_inner < 3 || o.nfolds_inner > 64) return ate_fail(ctx, -1, "ate_crossfit_aipw: nfolds_inner must be in 3..64");
    if (o.B > 0 && o.boot_mode == ATE_BOOT_INDEX && !boot_idx)
        return ate_fail(ctx, -1, "ate_crossfit_aipw: ATE_BOOT_INDEX needs boot_idx");
    if (n > 2147483647LL) return ate_fail(ctx, -1, "ate_crossfit_aipw: n exceeds 32-bit row ids");
    ATE_CUDA(ctx, cudaSetDevice(ctx->device));
    memset(res, 0, sizeof(*res));
    res->tau = res->se_sandwich = res->se_bootstrap = res->lower_ci = res->upper_ci = NAN;
    const long long launches0 = ctx->prof_launches;
    Timer tm;
    const size_t N = (size_t)n;
    const int K = o.K, F = o.nfolds_inner;
    int kb = o.fold_begin > 0 ? o.fold_begin : 1, ke = o.fold_end > 0 ? o.fold_end : K + 1;
    if (kb < 1 || ke > K + 1 || kb > ke) return ate_fail(ctx, -1, "ate_crossfit_aipw: bad fold range");
    const int nk = ke - kb;
    // ranks sharing the fit: rank r keeps rows [r * chunk, min(n, (r + 1) * chunk))
    const bool shared = o.shard_count > 1;
    if (shared && (o.shard_rank < 0 || o.shard_rank >= o.shard_count)) return ate_fail(ctx, -1, "ate_crossfit_aipw: bad shard rank");
    if (shared && !ctx->nccl_comm && !ctx->allreduce)
        return ate_fail(ctx, -1, "ate_crossfit_aipw: a shared fit needs ate_ctx_nccl_init or ate_ctx_set_allreduce first");
    if (shared && nk != K) return ate_fail(ctx, -1, "ate_crossfit_aipw: fold ranges and row sharding do not combine");
    const int G = shared ? o.shard_count : 1, rank = shared ? o.shard_rank : 0;
    const long long chunk = ((n + G - 1) / G + 3) / 4 * 4;
    const long long lo = std::min(n, (long long)rank * chunk), hi = std::min(n, lo + chunk), nloc = hi - lo;
    if (nloc < 1) return ate_fail(ctx, -1, "ate_crossfit_aipw: rank %d holds no rows", rank);
    const int p = o.expand == 0 ? p0 : (o.expand == 1 ? p0 * p0 : p0 + p0 * p0);
    const long long ld = round_up(p + 3, 32);
    const bool ids_on_device = o.on_device != 0;
    const bool draw_outer = (foldid == nullptr);
    if ((inner_ps == nullptr) != (inner_out == nullptr))
        return ate_fail(ctx, -1, "ate_crossfit_aipw: pass both inner fold matrices or neither");
    const bool draw_inner = (inner_ps == nullptr);
    if (ids_on_device && draw_outer != draw_inner)
        return ate_fail(ctx, -1, "ate_crossfit_aipw: with on_device, pass all fold ids (as device pointers) or none");

    // ---- device state of the call: one cached block ----
    Carve cv;
    double *dbase, *dwy, *dcs, *dnu_loc, *dnu, *dbeta, *dboot, *dpsi;
    int32_t *dfold, *dinner, *dlists;
    long long *dtot;
    FoldScratch sc;
    const size_t full = (size_t)chunk * G;  // >= n: the all-gathered vectors
    for (int pass = 0; pass < 2; ++pass) {
        cv.off = 0;
        dbase = cv.take<double>(N * p0);
        dwy = cv.take<double>(N * 2);
        dcs = cv.take<double>(2 * (size_t)p0);
        dnu_loc = cv.take<double>((size_t)chunk * 3);     // this rank's phat, m0, m1
        dnu = cv.take<double>(full * 5);                  // phat, m0, m1 over all rows; est1, est2
        dpsi = cv.take<double>(N);                        // est1 + est2: the table the bootstrap gathers
        dbeta = cv.take<double>(((size_t)p + 2) * 2 * K);
        dboot = cv.take<double>((size_t)(o.B > 0 ? o.B : 1) + G);
        dfold = cv.take<int32_t>(N);
        dinner = cv.take<int32_t>(ids_on_device && !draw_inner ? 1 : 2 * N);
        dlists = cv.take<int32_t>((size_t)nk * 3 * (size_t)nloc);
        dtot = cv.take<long long>((size_t)K * (4 * F + 2));
        sc.key_n = cv.take<int32_t>(N);
        sc.pos_n = cv.take<int32_t>(N);
        sc.key_a = cv.take<int32_t>(N);
        sc.key_b = cv.take<int32_t>(N);
        sc.key_h = cv.take<int32_t>(N);
        sc.pos_loc = cv.take<int32_t>((size_t)nloc);
        sc.blockhist = cv.take<int>(ate_i_rank_scratch_ints(n, F > K ? F : K));
        sc.totals = cv.take<long long>(80);
        sc.bad = cv.take<int>(4);
        if (pass == 0) {
            void *blk = nullptr;
            ATE_TRY(ate_cached(ctx, "crossfit.state", cv.off + 256, &blk));
            cv.base = (char *)blk;
        }
    }

    // ---- 1. inputs on the device; a1 standardisation (moments over ALL rows); a2 design of the local rows ----
    const cudaMemcpyKind kind = o.on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    ATE_CUDA(ctx, cudaMemcpyAsync(dbase, x_base, sizeof(double) * N * p0, kind, ctx->stream));
    ATE_CUDA(ctx, cudaMemcpyAsync(dwy, w, sizeof(double) * N, kind, ctx->stream));
    ATE_CUDA(ctx, cudaMemcpyAsync(dwy + N, y, sizeof(double) * N, kind, ctx->stream));
    if (o.standardize_base && o.n_continuous > 0) {
        const int nc = o.n_continuous < p0 ? o.n_continuous : p0;
        ATE_TRY(ate_i_standardize(ctx, dbase, n, p0, nullptr, nc, dcs, dcs + nc));
    }
    double *dX = nullptr;
    {
        void *pool = nullptr;
        ATE_TRY(ate_cached(ctx, "crossfit.design", sizeof(double) * (size_t)nloc * (size_t)ld, &pool));
        dX = (double *)pool;
    }
    ATE_TRY(ate_i_build_design(ctx, dbase + lo, n, nloc, p0, o.expand, dwy + lo, dwy + N + lo, dX, ld));
    if (getenv("ATE_TIMING")) ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    res->ms_design = tm.lap_ms();

    // ---- 2. fold structure: outer folds, inner cv.glmnet folds, row lists (device; folds.cu) ----
    std::vector<int32_t> h_fold, h_inner;
    if (draw_outer || draw_inner) ate_host_rng_seed(&ctx->rng, o.r_seed);  // ids are drawn: always from r_seed
    if (draw_outer) {
        h_fold.resize(N);
        ate_host_foldid(&ctx->rng, (int)n, K, o.rng_kind, h_fold.data());
        ATE_CUDA(ctx, cudaMemcpyAsync(dfold, h_fold.data(), sizeof(int32_t) * N, cudaMemcpyHostToDevice, ctx->stream));
    } else {
        ATE_CUDA(ctx, cudaMemcpyAsync(dfold, foldid, sizeof(int32_t) * N, ids_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice,
                                      ctx->stream));
    }
    ATE_CUDA(ctx, cudaMemsetAsync(sc.bad, 0, sizeof(int) * 4, ctx->stream));
    std::vector<long long> htot((size_t)K * (4 * F + 2), 0);
    const int32_t *fold_host = draw_outer ? h_fold.data() : (ids_on_device ? nullptr : foldid);
    for (int k = 1; k <= K; ++k) {
        const int32_t *da, *db;
        if (draw_inner) {
            // R's order inside the loop over k: the propensity cv.glmnet draws first, then the outcome one
            long long nt = 0;
            if (fold_host) {
                for (long long i = 0; i < n; ++i) nt += (fold_host[i] != k);
            } else {
                return ate_fail(ctx, -1, "ate_crossfit_aipw: inner ids can only be drawn when the outer ids are on the host");
            }
            h_inner.resize(2 * N);
            ate_host_foldid(&ctx->rng, (int)nt, F, o.rng_kind, h_inner.data());
            ate_host_foldid(&ctx->rng, (int)nt, F, o.rng_kind, h_inner.data() + N);
            // (a pageable source is staged before cudaMemcpyAsync returns: h_inner may be redrawn next turn)
            ATE_CUDA(ctx, cudaMemcpyAsync(dinner, h_inner.data(), sizeof(int32_t) * nt, cudaMemcpyHostToDevice, ctx->stream));
            ATE_CUDA(ctx, cudaMemcpyAsync(dinner + N, h_inner.data() + N, sizeof(int32_t) * nt, cudaMemcpyHostToDevice, ctx->stream));
            da = dinner;
            db = dinner + N;
        } else if (ids_on_device) {
            da = inner_ps + (size_t)(k - 1) * N;
            db = inner_out + (size_t)(k - 1) * N;
        } else {
            ATE_CUDA(ctx, cudaMemcpyAsync(dinner, inner_ps + (size_t)(k - 1) * N, sizeof(int32_t) * N, cudaMemcpyHostToDevice, ctx->stream));
            ATE_CUDA(ctx, cudaMemcpyAsync(dinner + N, inner_out + (size_t)(k - 1) * N, sizeof(int32_t) * N, cudaMemcpyHostToDevice, ctx->stream));
            da = dinner;
            db = dinner + N;
        }
        if (k < kb || k >= ke) continue;
        const int q = k - kb;
        int32_t *la = dlists + (size_t)(3 * q) * nloc, *lb = la + nloc, *lh = lb + nloc;
        ATE_TRY(ate_i_build_fold_lists(ctx, dfold, n, lo, hi, k, K, da, db, F, sc, la, lb, lh, dtot + (size_t)(k - 1) * (4 * F + 2)));
    }
    int hbad[4] = {0, 0, 0, 0};
    ATE_CUDA(ctx, cudaMemcpyAsync(htot.data(), dtot, sizeof(long long) * htot.size(), cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaMemcpyAsync(hbad, sc.bad, sizeof(int) * 4, cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (hbad[0] & 1) return ate_fail(ctx, -1, "ate_crossfit_aipw(foldid): fold id outside 1..%d", K);
    if (hbad[0] & 2) return ate_fail(ctx, -1, "ate_crossfit_aipw(inner_ps / inner_out): fold id outside 1..%d", F);
    std::vector<CvProblem> prob_ps(nk), prob_out(nk);
    std::vector<long long> nheld(nk);
    for (int q = 0; q < nk; ++q) {
        const long long *t = htot.data() + (size_t)(kb - 1 + q) * (4 * F + 2);
        const long long nt_glob = t[0];
        if (nt_glob < 2 * F || nt_glob >= n) return ate_fail(ctx, -1, "ate_crossfit_aipw: fold %d is degenerate", kb + q);
        CvProblem *pr[2] = {&prob_ps[q], &prob_out[q]};
        for (int which = 0; which < 2; ++which) {
            CvProblem &c = *pr[which];
            c.nseg = F;
            c.seg.assign(F + 1, 0);
            c.seg_global.assign(F, 0);
            for (int f = 0; f < F; ++f) {
                c.seg[f + 1] = c.seg[f] + t[1 + which * F + f];
                c.seg_global[f] = t[2 + 2 * F + which * F + f];
                if (c.seg_global[f] < 1) return ate_fail(ctx, -1, "ate_crossfit_aipw: an inner fold of outer fold %d is empty", kb + q);
            }
            c.nt = c.seg[F];
            c.nt_global = nt_glob;
            c.d_list = dlists + (size_t)(3 * q + which) * nloc;
        }
        nheld[q] = t[1 + 2 * F];
    }
    res->ms_folds = tm.lap_ms();

    // nuisance vectors; rows of folds not handled here stay 0 so shards can be summed
    ATE_CUDA(ctx, cudaMemsetAsync(dnu_loc, 0, sizeof(double) * (size_t)chunk * 3, ctx->stream));
    double *dphat_l = dnu_loc, *dm0_l = dnu_loc + chunk, *dm1_l = dnu_loc + 2 * chunk;
    double *dphat = dnu, *dm0 = dnu + full, *dm1 = dnu + 2 * full, *de1 = dnu + 3 * full, *de2 = dnu + 4 * full;
    PathOpts po;
    po.nlam = o.nlambda > 0 ? o.nlambda : 100;
    po.thr = o.thresh > 0 ? o.thresh : 1e-7;
    po.maxit = o.maxit > 0 ? o.maxit : 100000;
    po.flmin = 1.0e-4;  // nobs >= nvars on every training set we accept
    if (shared) {
        po.shard_rank = rank;
        po.shard_count = G;
    }
    int worst = 0;
    const size_t bstride = (size_t)p + 2;
    std::vector<double> hbeta(bstride * 2 * K, 0.0);

    // ---- 3a. outcome model: gaussian lasso on [covariates, W], penalty.factor = c(1..1, 0)
    //          (ate_functions.R:91-103), predicted at W := 1 and W := 0 (:221-228) ----
    {
        std::vector<double> vp((size_t)p + 1, 1.0);
        vp[p] = 0.0;
        std::vector<CvOut> outs;
        ATE_TRY(ate_i_cv_gaussian(ctx, dX, ld, p + 1, p + 1, p + 2, vp.data(), po, prob_out, outs));
        std::vector<double> a0s(nk), bws(nk);
        for (int q = 0; q < nk; ++q) {
            const CvOut &c = outs[q];
            if (c.jerr > 0 || c.lmu < 1) return ate_fail(ctx, -3, "outcome fit of fold %d failed (jerr %d)", kb + q, c.jerr);
            if (c.jerr && !worst) worst = c.jerr;
            const int s = o.use_1se ? c.idx_1se : c.idx_min;
            double *b = hbeta.data() + bstride * (size_t)q;
            for (int j = 0; j <= p; ++j) b[j] = c.beta[(size_t)s * (p + 1) + j];
            a0s[q] = c.a0[s];
            bws[q] = b[p];
            if (lambda_out) lambda_out[kb - 1 + q] = c.lambda[s];
        }
        ATE_CUDA(ctx, cudaMemcpyAsync(dbeta, hbeta.data(), sizeof(double) * bstride * nk, cudaMemcpyHostToDevice, ctx->stream));
        for (int q = 0; q < nk; ++q)
            ATE_TRY(ate_i_predict_rows(ctx, dX, ld, p, dlists + (size_t)(3 * q + 2) * nloc, 0, nheld[q], dbeta + bstride * (size_t)q,
                                       a0s[q], bws[q], 2, 1, dm0_l, dm1_l));
        if (getenv("ATE_TIMING")) ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    res->ms_outcome = tm.lap_ms();

    // ---- 3b. propensity model: L1-logistic on the covariates (ate_functions.R:139-141),
    //          type = "response" at the selected lambda (:144) ----
    {
        std::vector<CvOut> outs;
        IrlsStats st;
        ATE_TRY(ate_i_cv_binomial(ctx, dX, ld, p, p, nullptr, po, prob_ps, outs, &st));
        res->rounds_irls = st.rounds;
        res->irls_solves = st.solves;
        res->step_launches = st.step_launches;
        res->row_fits = st.row_fits;
        res->rows_staged = st.rows_staged;
        res->list_misses = st.list_misses;
        if (st.solves > 0) {
            res->strong_mean = st.sum_strong / st.solves;
            res->strong_frac2 = st.sum_strong2 / st.solves / ((double)p * p);
        }
        std::vector<double> a0s(nk);
        double *hb2 = hbeta.data() + bstride * (size_t)K;
        for (int q = 0; q < nk; ++q) {
            const CvOut &c = outs[q];
            if (c.jerr > 0 || c.lmu < 1) return ate_fail(ctx, -3, "propensity fit of fold %d failed (jerr %d)", kb + q, c.jerr);
            if (c.jerr && !worst) worst = c.jerr;
            const int s = o.use_1se ? c.idx_1se : c.idx_min;
            for (int j = 0; j < p; ++j) hb2[bstride * (size_t)q + j] = c.beta[(size_t)s * p + j];
            a0s[q] = c.a0[s];
            if (lambda_ps) lambda_ps[kb - 1 + q] = c.lambda[s];
        }
        double *db2 = dbeta + bstride * (size_t)K;
        ATE_CUDA(ctx, cudaMemcpyAsync(db2, hb2, sizeof(double) * bstride * nk, cudaMemcpyHostToDevice, ctx->stream));
        for (int q = 0; q < nk; ++q)
            ATE_TRY(ate_i_predict_rows(ctx, dX, ld, p, dlists + (size_t)(3 * q + 2) * nloc, 0, nheld[q], db2 + bstride * (size_t)q,
                                       a0s[q], 0.0, 1, 1, dphat_l, nullptr));
    }
    // every rank ends with the nuisances of ALL rows (rank-major slices of `chunk` rows)
    if (shared) {
        ATE_TRY(ate_i_allgather(ctx, dphat_l, dphat, chunk, rank, G));
        ATE_TRY(ate_i_allgather(ctx, dm0_l, dm0, chunk, rank, G));
        ATE_TRY(ate_i_allgather(ctx, dm1_l, dm1, chunk, rank, G));
    } else {
        ATE_CUDA(ctx, cudaMemcpyAsync(dphat, dphat_l, sizeof(double) * N, cudaMemcpyDeviceToDevice, ctx->stream));
        ATE_CUDA(ctx, cudaMemcpyAsync(dm0, dm0_l, sizeof(double) * N, cudaMemcpyDeviceToDevice, ctx->stream));
        ATE_CUDA(ctx, cudaMemcpyAsync(dm1, dm1_l, sizeof(double) * N, cudaMemcpyDeviceToDevice, ctx->stream));
    }
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    res->ms_propensity = tm.lap_ms();
    res->jerr = worst;
    res->gram_launches = ctx->prof_launches - launches0;
    auto fetch = [&](double *dst, const double *src) -> int {
        if (!dst) return 0;
        ATE_CUDA(ctx, cudaMemcpyAsync(dst, src, sizeof(double) * N, o.on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost,
                                      ctx->stream));
        return 0;
    };
    ATE_TRY(fetch(phat, dphat));
    ATE_TRY(fetch(m0, dm0));
    ATE_TRY(fetch(m1, dm1));
    if (o.skip_score || nk != K) {
        ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        return 0;
    }

    // ---- 4. a7/a8 score; 5. a9/a10 bootstrap; 6. a11 interval ----
    double n_ok = 0.0;
    ATE_TRY(ate_i_score(ctx, n, dwy, dwy + N, dphat, dm0, dm1, de1, de2, &res->tau, &res->se_sandwich, dpsi, &n_ok));
    res->ms_score = tm.lap_ms();
    double se = res->se_sandwich;
    if (o.B > 0) {
        // ranks sharing the fit split the replicates by id (the Philox stream of a replicate depends on
        // (seed, id) only) and gather the estimates; a caller may instead ask for a slice of its own
        long long rep0 = o.boot_rep_offset;
        int reps = o.boot_reps_here > 0 ? o.boot_reps_here : o.B;
        const bool split = shared && o.boot_reps_here <= 0 && o.boot_mode == ATE_BOOT_PHILOX;
        const int per = (o.B + G - 1) / G;
        if (split) {
            rep0 = (long long)rank * per;
            reps = (int)std::max(0LL, std::min((long long)o.B, rep0 + per) - rep0);
        }
        std::vector<double> hb((size_t)(split ? per * G : reps), 0.0);
        const int32_t *dix = nullptr;
        if (o.boot_mode == ATE_BOOT_INDEX) dix = boot_idx;  // host or device; ate_i_bootstrap streams host matrices in blocks
        ATE_TRY(ate_i_bootstrap(ctx, n, de1, de2, reps, o.boot_mode, dix, o.boot_mode == ATE_BOOT_INDEX && !o.on_device,
                                o.boot_seed, rep0, split ? dboot : nullptr, hb.data() + (split ? (size_t)rank * per : 0), dpsi,
                                n_ok != (double)n));
        if (split) {
            // dboot holds this rank's `reps` estimates; gather `per` from everyone
            void *gbuf = nullptr;
            ATE_TRY(ate_cached(ctx, "crossfit.bootgather", sizeof(double) * (size_t)per * (G + 1), &gbuf));
            double *gsend = (double *)gbuf, *grecv = gsend + per;
            ATE_CUDA(ctx, cudaMemsetAsync(gsend, 0, sizeof(double) * per, ctx->stream));
            ATE_CUDA(ctx, cudaMemcpyAsync(gsend, dboot, sizeof(double) * reps, cudaMemcpyDeviceToDevice, ctx->stream));
            ATE_TRY(ate_i_allgather(ctx, gsend, grecv, per, rank, G));
            ATE_CUDA(ctx, cudaMemcpyAsync(hb.data(), grecv, sizeof(double) * (size_t)per * G, cudaMemcpyDeviceToHost, ctx->stream));
            ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            // rank r's slice starts at r * per; the tail of the last slice is padding
            std::vector<double> all((size_t)o.B);
            for (int b = 0; b < o.B; ++b) all[b] = hb[b];
            hb.swap(all);
            reps = o.B;
        }
        if (boot)
            for (int b = 0; b < reps; ++b) boot[b] = hb[b];
        if (reps == o.B) {
            res->se_bootstrap = ate_host_sd(hb.data(), reps);
            se = res->se_bootstrap;
        }
        res->ms_bootstrap = tm.lap_ms();
    }
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    res->lower_ci = res->tau - 1.96 * se;
    res->upper_ci = res->tau + 1.96 * se;
    return 0;
}
