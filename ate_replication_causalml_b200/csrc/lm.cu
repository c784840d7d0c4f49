// lm.cu -- rows f3/f4: stats::lm (+ summary.lm) as the reference's regression estimators use it:
//   ate_condmean_ols   lm(Y ~ .)                      ate_functions.R:25-29, coefficient table :31-34
//   prop_score_weight  lm(tau_hat ~ .) residuals      ate_functions.R:52-56
//   prop_score_ols     lm(Y ~ W, weights = weights)   ate_functions.R:72-76, table :78-81
//   belloni            lm(y ~ x_restricted)           ate_functions.R:320, table :322-323
//
// R factors the n x (p+1) model matrix by Householder QR (dqrdc2, limited pivoting, tol 1e-7).
// Here the only O(n) work is two streaming passes over the row-major design [x | 1 | y]: ONE
// weighted Gram on the DMMA pipe (X'WX, X'Wy, y'Wy together) and one residual pass (sum w e^2
// accumulated from the residuals themselves, not from y'y - b'X'y, which cancels).  The (p+1)^2
// factorisation runs on the host in long double, columns taken in R's order ((Intercept) first);
// a column whose remaining norm falls below tol * its original norm is set aside exactly where
// dqrdc2 would move it to the end, and its coefficient is reported as NaN (R's NA).
#include <math.h>

#include <vector>

#include "common.cuh"

namespace {

constexpr int LM_THREADS = 256;

// e_i = y_i - b0 - x_i . b ;  partials: sum v e^2, count(v != 0).  One warp per row.
template <int ITER>
__global__ void __launch_bounds__(LM_THREADS)
lm_resid_kernel(const double *__restrict__ X, long long ld, long long n, int p, const double *__restrict__ coef,
                const double *__restrict__ v, double *__restrict__ resid, double *__restrict__ partials) {
    __shared__ double red[64];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    double b[2 * ITER];
#pragma unroll
    for (int i = 0; i < ITER; ++i) {
        const int c = 2 * lane + 64 * i;
        b[2 * i] = c < p ? coef[c] : 0.0;
        b[2 * i + 1] = c + 1 < p ? coef[c + 1] : 0.0;
    }
    const double b0 = coef[p];
    double rss = 0.0, cnt = 0.0;
    const long long wstride = (long long)gridDim.x * nw;
    for (long long t = (long long)blockIdx.x * nw + wid; t < n; t += wstride) {
        const double *xr = X + (size_t)t * (size_t)ld;
        double dot = 0.0;
#pragma unroll
        for (int i = 0; i < ITER; ++i) {
            const int c = 2 * lane + 64 * i;
            if (c + 1 < ld) {
                const double2 t2 = *reinterpret_cast<const double2 *>(xr + c);
                dot += t2.x * b[2 * i] + t2.y * b[2 * i + 1];
            }
        }
        dot = warp_sum(dot);
        if (lane == 0) {
            const double e = xr[p + 1] - b0 - dot;
            const double vi = v ? v[t] : 1.0;
            if (resid) resid[t] = e;
            if (vi != 0.0) {
                rss += vi * e * e;
                cnt += 1.0;
            }
        }
    }
    double acc[2] = {rss, cnt};
    block_sum<2>(acc, red);
    if (threadIdx.x == 0) {
        partials[2 * blockIdx.x] = acc[0];
        partials[2 * blockIdx.x + 1] = acc[1];
    }
}

template <int ITER>
void launch_resid(ate_ctx *ctx, const double *X, long long ld, long long n, int p, const double *coef, const double *v,
                  double *resid, double *part, int blocks) {
    lm_resid_kernel<ITER><<<blocks, LM_THREADS, 0, ctx->stream>>>(X, ld, n, p, coef, v, resid, part);
}

// Sequential Cholesky of the q x q moment matrix A (row-major, both triangles) with dqrdc2's
// set-aside rule.  kept[] lists the retained columns in order; L is the factor over kept columns.
struct LmFactor {
    std::vector<int> kept;
    std::vector<long double> L;  // kept.size() x kept.size(), row-major lower
    int r = 0;
};

void lm_factor(int q, const double *A, double tol, LmFactor *f) {
    f->kept.clear();
    f->L.assign((size_t)q * q, 0.0L);
    std::vector<long double> row(q);
    int r = 0;
    for (int j = 0; j < q; ++j) {
        // project column j on the kept ones: row[k] = L[j,k]
        long double d = A[(size_t)j * q + j];
        for (int k = 0; k < r; ++k) {
            long double s = A[(size_t)j * q + f->kept[k]];
            for (int m = 0; m < k; ++m) s -= row[m] * f->L[(size_t)k * q + m];
            row[k] = s / f->L[(size_t)k * q + k];
            d -= row[k] * row[k];
        }
        const long double orig = A[(size_t)j * q + j];
        // remaining norm < tol * original norm  (norms, hence tol^2 on the squares); an all-zero
        // column has nothing remaining either
        if (!(d > (long double)tol * tol * orig) || !(orig > 0.0L)) continue;
        for (int k = 0; k < r; ++k) f->L[(size_t)r * q + k] = row[k];
        f->L[(size_t)r * q + r] = sqrtl(d);
        f->kept.push_back(j);
        ++r;
    }
    f->r = r;
}

inline double nan_value() { return nan(""); }

}  // namespace

extern "C" int ate_lm(ate_ctx *ctx, long long n, int p, const double *x, const double *y, const double *weights,
                      double *coef, double *std_err, double *sigma, double *rss_out, int *rank,
                      long long *df_residual, double *resid) {
    if (!ctx) return -1;
    if ((!x && p > 0) || !y || !coef || n < 1 || p < 0) return ate_fail(ctx, -1, "ate_lm: bad arguments");
    if (p + 2 > 2048) return ate_fail(ctx, -1, "ate_lm: p = %d is beyond the host factorisation this entry point uses", p);
    if (weights)
        for (long long i = 0; i < n; ++i)
            if (!(weights[i] >= 0.0)) return ate_fail(ctx, -1, "ate_lm: missing or negative weights not allowed");
    ATE_CUDA(ctx, cudaSetDevice(ctx->device));
    const int q = p + 1;                           // unknowns: x columns, then the intercept
    const int g = p + 2;                           // Gram columns: [x | 1 | y]
    const long long ld = (g + 31) / 32 * 32;
    DevBuf<double> dx, dX, dv, dG, dcoef, dpart, dres;
    ATE_CUDA(ctx, dx.alloc((size_t)n * (p > 0 ? p : 1)));
    ATE_CUDA(ctx, dX.alloc((size_t)n * ld));
    ATE_CUDA(ctx, dG.alloc((size_t)g * g));
    ATE_CUDA(ctx, dcoef.alloc((size_t)q));
    int blocks = ctx->num_sms * 4;
    if ((n + 7) / 8 < blocks) blocks = (int)((n + 7) / 8);
    ATE_CUDA(ctx, dpart.alloc((size_t)2 * blocks));
    ATE_CUDA(ctx, cudaMemsetAsync(dX.p, 0, sizeof(double) * (size_t)n * ld, ctx->stream));
    if (p > 0) {
        ATE_CUDA(ctx, cudaMemcpyAsync(dx.p, x, sizeof(double) * (size_t)n * p, cudaMemcpyHostToDevice, ctx->stream));
        ATE_TRY(ate_i_transpose_to_rowmajor(ctx, dx.p, n, p, dX.p, ld, 0));
    }
    ATE_TRY(ate_i_fill_column(ctx, dX.p, n, ld, p, 1.0));
    ATE_CUDA(ctx, dv.alloc((size_t)n));
    ATE_CUDA(ctx, cudaMemcpyAsync(dv.p, y, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
    ATE_TRY(ate_i_transpose_to_rowmajor(ctx, dv.p, n, 1, dX.p, ld, p + 1));  // y column
    if (weights)
        ATE_CUDA(ctx, cudaMemcpyAsync(dv.p, weights, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
    const double *v = weights ? dv.p : nullptr;

    std::vector<double> hG((size_t)g * g), A((size_t)q * q), sol(q, 0.0), hpart((size_t)2 * blocks);
    ATE_TRY(ate_i_gram(ctx, dX.p, ld, g, nullptr, 0, n, v, dG.p));
    ATE_CUDA(ctx, cudaMemcpyAsync(hG.data(), dG.p, sizeof(double) * hG.size(), cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    // R's column order: (Intercept), then x.  perm[a] = Gram index of model column a.
    std::vector<int> perm(q);
    perm[0] = p;
    for (int a = 0; a < p; ++a) perm[1 + a] = a;
    std::vector<long double> rhs(q);
    for (int a = 0; a < q; ++a)
        for (int c = 0; c < q; ++c) A[(size_t)a * q + c] = hG[(size_t)perm[a] * g + perm[c]];
    LmFactor f;
    lm_factor(q, A.data(), 1e-7, &f);
    const int r = f.r;
    // forward/back substitution over the kept columns
    std::vector<long double> zv(r), bv(r);
    for (int k = 0; k < r; ++k) {
        long double s = hG[(size_t)perm[f.kept[k]] * g + (p + 1)];
        for (int m = 0; m < k; ++m) s -= f.L[(size_t)k * q + m] * zv[m];
        zv[k] = s / f.L[(size_t)k * q + k];
    }
    for (int k = r - 1; k >= 0; --k) {
        long double s = zv[k];
        for (int m = k + 1; m < r; ++m) s -= f.L[(size_t)m * q + k] * bv[m];
        bv[k] = s / f.L[(size_t)k * q + k];
    }
    // diag of (X'WX)^-1 = column sums of squares of L^-1
    std::vector<long double> Li((size_t)r * r, 0.0L), dinv(r, 0.0L);
    for (int c = 0; c < r; ++c) {
        Li[(size_t)c * r + c] = 1.0L / f.L[(size_t)c * q + c];
        for (int a = c + 1; a < r; ++a) {
            long double s = 0.0L;
            for (int m = c; m < a; ++m) s -= f.L[(size_t)a * q + m] * Li[(size_t)m * r + c];
            Li[(size_t)a * r + c] = s / f.L[(size_t)a * q + a];
        }
    }
    for (int c = 0; c < r; ++c)
        for (int a = c; a < r; ++a) dinv[c] += Li[(size_t)a * r + c] * Li[(size_t)a * r + c];

    const double nan = nan_value();
    for (int a = 0; a < q; ++a) coef[a] = nan;
    for (int k = 0; k < r; ++k) coef[f.kept[k]] = (double)bv[k];
    // device coefficient vector in design order, aliased columns contribute nothing
    for (int a = 0; a < q; ++a) {
        const double c = coef[a];
        sol[perm[a]] = (c == c) ? c : 0.0;
    }
    ATE_CUDA(ctx, cudaMemcpyAsync(dcoef.p, sol.data(), sizeof(double) * q, cudaMemcpyHostToDevice, ctx->stream));
    if (resid) ATE_CUDA(ctx, dres.alloc((size_t)n));
    const int it = (int)((ld + 63) / 64);
    if (it <= 1) launch_resid<1>(ctx, dX.p, ld, n, p, dcoef.p, v, dres.p, dpart.p, blocks);
    else if (it <= 2) launch_resid<2>(ctx, dX.p, ld, n, p, dcoef.p, v, dres.p, dpart.p, blocks);
    else if (it <= 4) launch_resid<4>(ctx, dX.p, ld, n, p, dcoef.p, v, dres.p, dpart.p, blocks);
    else if (it <= 8) launch_resid<8>(ctx, dX.p, ld, n, p, dcoef.p, v, dres.p, dpart.p, blocks);
    else if (it <= 16) launch_resid<16>(ctx, dX.p, ld, n, p, dcoef.p, v, dres.p, dpart.p, blocks);
    else if (it <= 32) launch_resid<32>(ctx, dX.p, ld, n, p, dcoef.p, v, dres.p, dpart.p, blocks);
    else return ate_fail(ctx, -1, "ate_lm: a %lld-column row is wider than the residual kernel takes", ld);
    ATE_LAUNCHED(ctx);
    ATE_CUDA(ctx, cudaMemcpyAsync(hpart.data(), dpart.p, sizeof(double) * 2 * blocks, cudaMemcpyDeviceToHost, ctx->stream));
    if (resid)
        ATE_CUDA(ctx, cudaMemcpyAsync(resid, dres.p, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    double rss = 0.0, cnt = 0.0;
    for (int b = 0; b < blocks; ++b) {
        rss += hpart[2 * b];
        cnt += hpart[2 * b + 1];
    }
    const long long rdf = (long long)cnt - r;  // zero-weight rows do not count (lm.wfit drops them)
    const double s2 = rdf > 0 ? rss / (double)rdf : nan;
    if (std_err) {
        for (int a = 0; a < q; ++a) std_err[a] = nan;
        for (int k = 0; k < r; ++k) std_err[f.kept[k]] = sqrt(s2 * (double)dinv[k]);
    }
    if (sigma) *sigma = sqrt(s2);
    if (rss_out) *rss_out = rss;
    if (rank) *rank = r;
    if (df_residual) *df_residual = rdf;
    return 0;
}
