// glm.cu -- f2: stats::glm(family = binomial("logit")) as doubly_robust_glm uses it for both
// nuisances (ate_functions.R:218-220 outcome, :231-234 propensity; also :156-158 and
// ate_replication.Rmd:165-168).  glm.fit's IRLS: mustart = (y + 0.5)/2, one weighted least-squares
// solve per step, stop when |dev - devold| / (|dev| + 0.1) < epsilon, at most maxit steps.
//
// Each step is the same two pieces as the penalised path: a row pass (eta, mu, working weight
// w = mu(1-mu), working response z, deviance; one warp per row) and ONE weighted Gram
// [x 1 z]' W [x 1 z] on the DMMA pipe, which holds X'WX and X'Wz at once.  R solves the step by
// QR (Cdqrls); here the (p+1)-dimensional normal equations are solved on the host by a long-double
// Cholesky (p <= ~400), which agrees to rounding on the well-conditioned designs of this path.
#include <math.h>

#include <vector>

#include "common.cuh"

namespace {

constexpr int GLM_THREADS = 256;

// binomial()$linkinv as R's C code computes it (family.c: logit_linkinv)
__device__ __forceinline__ double r_logit_linkinv(double eta) {
    const double eps = 2.220446049250313e-16;
    const double t = (eta < -30.0) ? eps : ((eta > 30.0) ? 1.0 / eps : exp(eta));
    return t / (1.0 + t);
}

// design row: [x (p) | 1 | z | y | pad].  init != 0: glm.fit's mustart; else eta from coef.
template <int ITER>
__global__ void __launch_bounds__(GLM_THREADS)
glm_update_kernel(double *__restrict__ X, long long ld, long long n, int p, const double *__restrict__ coef,
                  int init, double *__restrict__ w_out, double *__restrict__ dev_partials) {
    __shared__ double red[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    double b[2 * ITER];
#pragma unroll
    for (int i = 0; i < ITER; ++i) {
        const int c = 2 * lane + 64 * i;
        b[2 * i] = (!init && c < p) ? coef[c] : 0.0;
        b[2 * i + 1] = (!init && c + 1 < p) ? coef[c + 1] : 0.0;
    }
    const double b0 = init ? 0.0 : coef[p];
    double dev = 0.0;
    const long long wstride = (long long)gridDim.x * nw;
    for (long long t = (long long)blockIdx.x * nw + wid; t < n; t += wstride) {
        double *xr = X + (size_t)t * (size_t)ld;
        const double yi = xr[p + 2];
        double eta, mu;
        if (init) {
            mu = (yi + 0.5) / 2.0;
            eta = log(mu / (1.0 - mu));
        } else {
            double dot = 0.0;
#pragma unroll
            for (int i = 0; i < ITER; ++i) {
                const int c = 2 * lane + 64 * i;
                if (c + 1 < ld) {
                    const double2 t2 = *reinterpret_cast<const double2 *>(xr + c);
                    dot += t2.x * b[2 * i] + t2.y * b[2 * i + 1];
                }
            }
            eta = b0 + warp_sum(dot);
            mu = r_logit_linkinv(eta);
        }
        if (lane == 0) {
            const double w = mu * (1.0 - mu);
            xr[p + 1] = eta + (yi - mu) / w;  // working response
            w_out[t] = w;
            // binomial()$dev.resids for y in {0, 1} (the y log y terms vanish)
            const double a = yi > 0.0 ? yi * log(yi / mu) : 0.0;
            const double c = yi < 1.0 ? (1.0 - yi) * log((1.0 - yi) / (1.0 - mu)) : 0.0;
            dev += 2.0 * (a + c);
        }
    }
    double acc[1] = {dev};
    block_sum<1>(acc, red);
    if (threadIdx.x == 0) dev_partials[blockIdx.x] = acc[0];
}

// solve A x = b for symmetric positive definite A (q x q, row-major, both triangles) in long double
bool chol_solve(int q, const double *A_in, const double *b_in, double *x) {
    std::vector<long double> A((size_t)q * q), b(q);
    for (size_t e = 0; e < (size_t)q * q; ++e) A[e] = A_in[e];
    for (int a = 0; a < q; ++a) b[a] = b_in[a];
    for (int a = 0; a < q; ++a)
        for (int c = 0; c <= a; ++c) {
            long double s = A[(size_t)a * q + c];
            for (int k = 0; k < c; ++k) s -= A[(size_t)a * q + k] * A[(size_t)c * q + k];
            if (a == c) {
                if (!(s > 0.0L)) return false;
                A[(size_t)a * q + a] = sqrtl(s);
            } else
                A[(size_t)a * q + c] = s / A[(size_t)c * q + c];
        }
    for (int a = 0; a < q; ++a) {
        long double s = b[a];
        for (int k = 0; k < a; ++k) s -= A[(size_t)a * q + k] * b[k];
        b[a] = s / A[(size_t)a * q + a];
    }
    for (int a = q - 1; a >= 0; --a) {
        long double s = b[a];
        for (int k = a + 1; k < q; ++k) s -= A[(size_t)k * q + a] * b[k];
        b[a] = s / A[(size_t)a * q + a];
    }
    for (int a = 0; a < q; ++a) x[a] = (double)b[a];
    return true;
}

template <int ITER>
int launch_update(ate_ctx *ctx, double *X, long long ld, long long n, int p, const double *coef, int init,
                  double *w, double *part, int blocks) {
    glm_update_kernel<ITER><<<blocks, GLM_THREADS, 0, ctx->stream>>>(X, ld, n, p, coef, init, w, part);
    return 0;
}

}  // namespace

extern "C" int ate_glm_binomial(ate_ctx *ctx, long long n, int p, const double *x, const double *y, int maxit,
                                double epsilon, double *coef, double *deviance, int *iters, int *converged) {
    if (!ctx) return -1;
    if (!x || !y || !coef || n < 2 || p < 0) return ate_fail(ctx, -1, "ate_glm_binomial: bad arguments");
    if (p + 3 > 2048) return ate_fail(ctx, -1, "ate_glm_binomial: p = %d is beyond the host solve this entry point uses", p);
    ATE_CUDA(ctx, cudaSetDevice(ctx->device));
    if (maxit <= 0) maxit = 25;
    if (!(epsilon > 0)) epsilon = 1e-8;
    const int q = p + 1;                      // unknowns: x columns, then the intercept
    const long long ld = (p + 3 + 31) / 32 * 32;  // [x | 1 | z | y | 0...]
    DevBuf<double> dx, dX, dw, dG, dcoef, dpart;
    ATE_CUDA(ctx, dx.alloc((size_t)n * (p > 0 ? p : 1)));
    ATE_CUDA(ctx, dX.alloc((size_t)n * ld));
    ATE_CUDA(ctx, dw.alloc((size_t)n));
    ATE_CUDA(ctx, dG.alloc((size_t)(p + 2) * (p + 2)));
    ATE_CUDA(ctx, dcoef.alloc((size_t)q));
    int blocks = ctx->num_sms * 4;
    if ((n + 7) / 8 < blocks) blocks = (int)((n + 7) / 8);
    ATE_CUDA(ctx, dpart.alloc((size_t)blocks));
    ATE_CUDA(ctx, cudaMemsetAsync(dX.p, 0, sizeof(double) * (size_t)n * ld, ctx->stream));
    if (p > 0) {
        ATE_CUDA(ctx, cudaMemcpyAsync(dx.p, x, sizeof(double) * (size_t)n * p, cudaMemcpyHostToDevice, ctx->stream));
        ATE_TRY(ate_i_transpose_to_rowmajor(ctx, dx.p, n, p, dX.p, ld, 0));
    }
    ATE_TRY(ate_i_fill_column(ctx, dX.p, n, ld, p, 1.0));
    ATE_CUDA(ctx, cudaMemcpyAsync(dw.p, y, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
    ATE_TRY(ate_i_transpose_to_rowmajor(ctx, dw.p, n, 1, dX.p, ld, p + 2));  // y column

    std::vector<double> hpart(blocks), hG((size_t)(p + 2) * (p + 2)), A((size_t)q * q), rhs(q), sol(q);
    auto update = [&](int init) -> int {
        const int it = (int)((ld + 63) / 64);
        if (it <= 1) launch_update<1>(ctx, dX.p, ld, n, p, dcoef.p, init, dw.p, dpart.p, blocks);
        else if (it <= 2) launch_update<2>(ctx, dX.p, ld, n, p, dcoef.p, init, dw.p, dpart.p, blocks);
        else if (it <= 4) launch_update<4>(ctx, dX.p, ld, n, p, dcoef.p, init, dw.p, dpart.p, blocks);
        else if (it <= 8) launch_update<8>(ctx, dX.p, ld, n, p, dcoef.p, init, dw.p, dpart.p, blocks);
        else if (it <= 16) launch_update<16>(ctx, dX.p, ld, n, p, dcoef.p, init, dw.p, dpart.p, blocks);
        else if (it <= 32) launch_update<32>(ctx, dX.p, ld, n, p, dcoef.p, init, dw.p, dpart.p, blocks);
        else return ate_fail(ctx, -1, "ate_glm_binomial: a %lld-column row is wider than the update kernel takes", ld);
        ATE_LAUNCHED(ctx);
        ATE_CUDA(ctx, cudaMemcpyAsync(hpart.data(), dpart.p, sizeof(double) * blocks, cudaMemcpyDeviceToHost, ctx->stream));
        ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        return 0;
    };
    auto dev_of = [&]() {
        double s = 0.0;
        for (int b = 0; b < blocks; ++b) s += hpart[b];
        return s;
    };
    ATE_TRY(update(1));
    double devold = dev_of(), dev = devold;
    int conv = 0, it = 0;
    for (it = 1; it <= maxit; ++it) {
        // [x 1 z]' W [x 1 z]: the (p+1) x (p+1) leading block is X'WX, the last column X'Wz
        ATE_TRY(ate_i_gram(ctx, dX.p, ld, p + 2, nullptr, 0, n, dw.p, dG.p));
        ATE_CUDA(ctx, cudaMemcpyAsync(hG.data(), dG.p, sizeof(double) * hG.size(), cudaMemcpyDeviceToHost, ctx->stream));
        ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        for (int a = 0; a < q; ++a) {
            for (int c = 0; c < q; ++c) A[(size_t)a * q + c] = hG[(size_t)a * (p + 2) + c];
            rhs[a] = hG[(size_t)a * (p + 2) + p + 1];
        }
        if (!chol_solve(q, A.data(), rhs.data(), sol.data()))
            return ate_fail(ctx, -3, "ate_glm_binomial: X'WX is not positive definite at step %d (collinear columns?)", it);
        ATE_CUDA(ctx, cudaMemcpyAsync(dcoef.p, sol.data(), sizeof(double) * q, cudaMemcpyHostToDevice, ctx->stream));
        ATE_TRY(update(0));
        dev = dev_of();
        if (fabs(dev - devold) / (fabs(dev) + 0.1) < epsilon) {
            conv = 1;
            break;
        }
        devold = dev;
    }
    coef[0] = sol[p];  // R's order: (Intercept) first
    for (int a = 0; a < p; ++a) coef[1 + a] = sol[a];
    if (deviance) *deviance = dev;
    if (iters) *iters = it > maxit ? maxit : it;
    if (converged) *converged = conv;
    return 0;
}
