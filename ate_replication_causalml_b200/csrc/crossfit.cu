// crossfit.cu -- public glmnet / cv.glmnet / predict entry points and the K-fold cross-fitted
// AIPW composite (steps 1-6 of the coverage contract).
#include <math.h>
#include <string.h>

#include <chrono>

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
    if (o.K < 2) return ate_fail(ctx, -1, "ate_crossfit_aipw: K must be >= 2");
    if (o.nfolds_inner < 3) return ate_fail(ctx, -1, "ate_crossfit_aipw: nfolds_inner must be >= 3");
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

    // ---- 1. inputs on the device; a1 standardisation; a2 design ----
    DevBuf<double> dbase, dwy, dcs;
    ATE_CUDA(ctx, dbase.alloc(N * p0));
    ATE_CUDA(ctx, dwy.alloc(N * 2));
    const cudaMemcpyKind kind = o.on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    ATE_CUDA(ctx, cudaMemcpyAsync(dbase.p, x_base, sizeof(double) * N * p0, kind, ctx->stream));
    ATE_CUDA(ctx, cudaMemcpyAsync(dwy.p, w, sizeof(double) * N, kind, ctx->stream));
    ATE_CUDA(ctx, cudaMemcpyAsync(dwy.p + N, y, sizeof(double) * N, kind, ctx->stream));
    if (o.standardize_base && o.n_continuous > 0) {
        const int nc = o.n_continuous < p0 ? o.n_continuous : p0;
        ATE_CUDA(ctx, dcs.alloc(2 * (size_t)nc));
        ATE_TRY(ate_i_standardize(ctx, dbase.p, n, p0, nullptr, nc, dcs.p, dcs.p + nc));
    }
    const int p = o.expand == 0 ? p0 : (o.expand == 1 ? p0 * p0 : p0 + p0 * p0);
    const long long ld = round_up(p + 3, 32);
    struct { double *p; } dX;
    {
        void *pool = nullptr;
        ATE_TRY(ate_cached(ctx, "crossfit.design", sizeof(double) * N * (size_t)ld, &pool));
        dX.p = (double *)pool;
    }
    ATE_TRY(ate_i_build_design(ctx, dbase.p, n, p0, o.expand, dwy.p, dwy.p + N, dX.p, ld));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    dbase.release();
    res->ms_design = tm.lap_ms();

    // ---- 2. fold structure (host): outer folds, inner cv.glmnet folds, row lists ----
    std::vector<int32_t> fid(N);
    const bool draw = (foldid == nullptr);
    if (draw)
        ate_host_rng_seed(&ctx->rng, o.r_seed);
    if (foldid) {
        ATE_TRY(check_foldid(ctx, foldid, n, K, "ate_crossfit_aipw(foldid)"));
        fid.assign(foldid, foldid + n);
    } else {
        ate_host_foldid(&ctx->rng, (int)n, K, o.rng_kind, fid.data());
    }
    if ((inner_ps == nullptr) != (inner_out == nullptr))
        return ate_fail(ctx, -1, "ate_crossfit_aipw: pass both inner fold matrices or neither");
    if (!inner_ps && !ctx->rng.seeded) ate_host_rng_seed(&ctx->rng, o.r_seed);
    const int nk = ke - kb;
    std::vector<CvProblem> prob_ps(nk), prob_out(nk);
    std::vector<std::vector<int32_t>> lists;  // per k: ps-grouped, out-grouped, held-out
    std::vector<long long> offs;
    size_t total = 0;
    std::vector<int32_t> tmp_ids;
    for (int k = 1; k <= K; ++k) {
        std::vector<int32_t> train, held;
        for (long long i = 0; i < n; ++i) (fid[i] == k ? held : train).push_back((int32_t)i);
        const long long nt = (long long)train.size();
        if (nt < 2 * F || held.empty()) return ate_fail(ctx, -1, "ate_crossfit_aipw: fold %d is degenerate", k);
        const int32_t *ips, *iout;
        std::vector<int32_t> dps, dout_;
        if (inner_ps) {
            ips = inner_ps + (size_t)(k - 1) * N;
            iout = inner_out + (size_t)(k - 1) * N;
            ATE_TRY(check_foldid(ctx, ips, nt, F, "ate_crossfit_aipw(inner_ps)"));
            ATE_TRY(check_foldid(ctx, iout, nt, F, "ate_crossfit_aipw(inner_out)"));
        } else {
            // R's order inside the loop over k: propensity cv.glmnet draws first, then outcome
            dps.resize(nt);
            dout_.resize(nt);
            ate_host_foldid(&ctx->rng, (int)nt, F, o.rng_kind, dps.data());
            ate_host_foldid(&ctx->rng, (int)nt, F, o.rng_kind, dout_.data());
            ips = dps.data();
            iout = dout_.data();
        }
        if (k < kb || k >= ke) continue;
        const int q = k - kb;
        std::vector<int32_t> gps, gout;
        group_by_fold(train, ips, F, gps, prob_ps[q].seg);
        group_by_fold(train, iout, F, gout, prob_out[q].seg);
        prob_ps[q].nt = prob_out[q].nt = nt;
        prob_ps[q].nseg = prob_out[q].nseg = F;
        offs.push_back((long long)total); total += gps.size();
        offs.push_back((long long)total); total += gout.size();
        offs.push_back((long long)total); total += held.size();
        lists.push_back(std::move(gps));
        lists.push_back(std::move(gout));
        lists.push_back(std::move(held));
    }
    DevBuf<int32_t> dlists;
    ATE_CUDA(ctx, dlists.alloc(total));
    for (size_t l = 0; l < lists.size(); ++l)
        ATE_CUDA(ctx, cudaMemcpyAsync(dlists.p + offs[l], lists[l].data(), sizeof(int32_t) * lists[l].size(), cudaMemcpyHostToDevice, ctx->stream));
    for (int q = 0; q < nk; ++q) {
        prob_ps[q].d_list = dlists.p + offs[3 * q];
        prob_out[q].d_list = dlists.p + offs[3 * q + 1];
    }
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));

    // nuisance vectors; rows of folds not handled here stay 0 so shards can be summed
    DevBuf<double> dnu;
    ATE_CUDA(ctx, dnu.alloc(N * 5));  // phat, m0, m1, est1, est2
    ATE_CUDA(ctx, cudaMemsetAsync(dnu.p, 0, sizeof(double) * N * 3, ctx->stream));
    double *dphat = dnu.p, *dm0 = dnu.p + N, *dm1 = dnu.p + 2 * N, *de1 = dnu.p + 3 * N, *de2 = dnu.p + 4 * N;
    PathOpts po;
    po.nlam = o.nlambda > 0 ? o.nlambda : 100;
    po.thr = o.thresh > 0 ? o.thresh : 1e-7;
    po.maxit = o.maxit > 0 ? o.maxit : 100000;
    po.flmin = 1.0e-4;  // nobs >= nvars on every training set we accept
    if (o.shard_count > 1) {
        if (o.shard_rank < 0 || o.shard_rank >= o.shard_count) return ate_fail(ctx, -1, "ate_crossfit_aipw: bad shard rank");
        po.shard_rank = o.shard_rank;
        po.shard_count = o.shard_count;
    }
    DevBuf<double> dbeta;
    ATE_CUDA(ctx, dbeta.alloc((size_t)p + 1));
    int worst = 0;

    // ---- 3a. outcome model: gaussian lasso on [covariates, W], penalty.factor = c(1..1, 0)
    //          (ate_functions.R:91-103), predicted at W := 1 and W := 0 (:221-228) ----
    {
        std::vector<double> vp((size_t)p + 1, 1.0);
        vp[p] = 0.0;
        // Selected model of every fold: jerr, lmu, lambda, a0, beta[p + 1].  Ranks sharing the fit
        // solve the folds round-robin and sum the zero-filled records through the all-reduce hook.
        const size_t rec = (size_t)p + 5;
        std::vector<double> sel(rec * (size_t)nk, 0.0);
        const bool shared = o.shard_count > 1;
        if (shared && !ctx->allreduce) return ate_fail(ctx, -1, "ate_crossfit_aipw: sharded fit but no all-reduce hook set");
        std::vector<CvProblem> mine;
        std::vector<int> mine_q;
        for (int q = 0; q < nk; ++q)
            if (!shared || q % o.shard_count == o.shard_rank) {
                mine.push_back(prob_out[q]);
                mine_q.push_back(q);
            }
        if (!mine.empty()) {
            std::vector<CvOut> outs;
            ATE_TRY(ate_i_cv_gaussian(ctx, dX.p, ld, p + 1, p + 1, p + 2, vp.data(), po, mine, outs));
            for (size_t m = 0; m < mine.size(); ++m) {
                const CvOut &c = outs[m];
                double *r = sel.data() + rec * (size_t)mine_q[m];
                r[0] = c.jerr;
                r[1] = c.lmu;
                if (c.lmu < 1) continue;
                const int s = o.use_1se ? c.idx_1se : c.idx_min;
                r[2] = c.lambda[s];
                r[3] = c.a0[s];
                for (int j = 0; j <= p; ++j) r[4 + j] = c.beta[(size_t)s * (p + 1) + j];
            }
        }
        if (shared && ctx->allreduce(ctx->allreduce_user, sel.data(), (long long)sel.size()) != 0)
            return ate_fail(ctx, -4, "ate_crossfit_aipw: all-reduce hook failed");
        for (int q = 0; q < nk; ++q) {
            const double *r = sel.data() + rec * (size_t)q;
            const int jerr = (int)r[0], lmu = (int)r[1];
            if (jerr > 0 || lmu < 1) return ate_fail(ctx, -3, "outcome fit of fold %d failed (jerr %d)", kb + q, jerr);
            if (jerr && !worst) worst = jerr;
            const double *b = r + 4;
            ATE_CUDA(ctx, cudaMemcpyAsync(dbeta.p, b, sizeof(double) * (p + 1), cudaMemcpyHostToDevice, ctx->stream));
            ATE_TRY(ate_i_predict_rows(ctx, dX.p, ld, p, dlists.p + offs[3 * q + 2], 0, (long long)lists[3 * q + 2].size(),
                                       dbeta.p, r[3], b[p], 2, 1, dm0, dm1));
            ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            if (lambda_out) lambda_out[kb - 1 + q] = r[2];
        }
    }
    res->ms_outcome = tm.lap_ms();

    // ---- 3b. propensity model: L1-logistic on the covariates (ate_functions.R:139-141),
    //          type = "response" at the selected lambda (:144) ----
    {
        std::vector<CvOut> outs;
        IrlsStats st;
        ATE_TRY(ate_i_cv_binomial(ctx, dX.p, ld, p, p, nullptr, po, prob_ps, outs, &st));
        res->rounds_irls = st.rounds;
        res->irls_solves = st.solves;
        if (st.solves > 0) {
            res->strong_mean = st.sum_strong / st.solves;
            res->strong_frac2 = st.sum_strong2 / st.solves / ((double)p * p);
        }
        for (int q = 0; q < nk; ++q) {
            const CvOut &c = outs[q];
            if (c.jerr > 0 || c.lmu < 1) return ate_fail(ctx, -3, "propensity fit of fold %d failed (jerr %d)", kb + q, c.jerr);
            if (c.jerr && !worst) worst = c.jerr;
            const int s = o.use_1se ? c.idx_1se : c.idx_min;
            const double *b = c.beta.data() + (size_t)s * p;
            ATE_CUDA(ctx, cudaMemcpyAsync(dbeta.p, b, sizeof(double) * p, cudaMemcpyHostToDevice, ctx->stream));
            ATE_TRY(ate_i_predict_rows(ctx, dX.p, ld, p, dlists.p + offs[3 * q + 2], 0, (long long)lists[3 * q + 2].size(),
                                       dbeta.p, c.a0[s], 0.0, 1, 1, dphat, nullptr));
            ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            if (lambda_ps) lambda_ps[kb - 1 + q] = c.lambda[s];
        }
    }
    res->ms_propensity = tm.lap_ms();
    res->jerr = worst;
    res->gram_launches = ctx->prof_launches - launches0;
    auto fetch = [&](double *dst, const double *src) -> int {
        if (!dst) return 0;
        if (o.on_device)
            ATE_CUDA(ctx, cudaMemcpyAsync(dst, src, sizeof(double) * N, cudaMemcpyDeviceToDevice, ctx->stream));
        else
            ATE_CUDA(ctx, cudaMemcpyAsync(dst, src, sizeof(double) * N, cudaMemcpyDeviceToHost, ctx->stream));
        return 0;
    };
    ATE_TRY(fetch(phat, dphat));
    ATE_TRY(fetch(m0, dm0));
    ATE_TRY(fetch(m1, dm1));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (o.skip_score || nk != K) return 0;

    // ---- 4. a7/a8 score; 5. a9/a10 bootstrap; 6. a11 interval ----
    ATE_TRY(ate_i_score(ctx, n, dwy.p, dwy.p + N, dphat, dm0, dm1, de1, de2, &res->tau, &res->se_sandwich));
    res->ms_score = tm.lap_ms();
    double se = res->se_sandwich;
    if (o.B > 0) {
        const int reps = o.boot_reps_here > 0 ? o.boot_reps_here : o.B;
        std::vector<double> hb((size_t)reps);
        DevBuf<int32_t> didx;
        const int32_t *dix = nullptr;
        if (o.boot_mode == ATE_BOOT_INDEX) {
            ATE_CUDA(ctx, didx.alloc(N * (size_t)reps));
            ATE_CUDA(ctx, cudaMemcpyAsync(didx.p, boot_idx, sizeof(int32_t) * N * (size_t)reps, cudaMemcpyHostToDevice, ctx->stream));
            dix = didx.p;
        }
        ATE_TRY(ate_i_bootstrap(ctx, n, de1, de2, reps, o.boot_mode, dix, o.boot_seed, o.boot_rep_offset, hb.data()));
        if (boot)
            for (int b = 0; b < reps; ++b) boot[b] = hb[b];
        if (reps == o.B) {
            res->se_bootstrap = ate_host_sd(hb.data(), reps);
            se = res->se_bootstrap;
        }
        res->ms_bootstrap = tm.lap_ms();
    }
    res->lower_ci = res->tau - 1.96 * se;
    res->upper_ci = res->tau + 1.96 * se;
    return 0;
}
