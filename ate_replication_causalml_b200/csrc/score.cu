// score.cu -- a7/a8: fused AIPW score, ATE and sandwich SE (ate_functions.R:241-243,255-256).
//
// One HBM pass over the five n-vectors (w, y, p, m0, m1): 5*n*8 B read, plus 2*n*8 B written
// when est1/est2 are materialised for the bootstrap.  Each block leaves six partial sums in
// scratch; the last block to finish folds them in block order (deterministic) and writes
// tau and se, so score + fold reduction are a single launch.
#include <math.h>

#include "common.cuh"

namespace {

constexpr int SCORE_THREADS = 256;
constexpr int SCORE_NV = 6;  // sum est1 (non-NaN), count, sum est2, sum J, sum J^2, (spare)

__global__ void __launch_bounds__(SCORE_THREADS)
aipw_score_kernel(long long n, const double *__restrict__ w, const double *__restrict__ y,
                  const double *__restrict__ p, const double *__restrict__ m0,
                  const double *__restrict__ m1, double *__restrict__ est1,
                  double *__restrict__ est2, double *__restrict__ partials,
                  unsigned int *__restrict__ ticket, double *__restrict__ out) {
    __shared__ double red[SCORE_NV * 32];
    __shared__ bool is_last;
    double acc[SCORE_NV] = {0, 0, 0, 0, 0, 0};
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const double wi = __ldg(w + i), yi = __ldg(y + i), pi = __ldg(p + i);
        const double a0 = __ldg(m0 + i), a1 = __ldg(m1 + i);
        const double e1 = wi * (yi - a1) / pi + (1.0 - wi) * (yi - a0) / (1.0 - pi);
        const double e2 = a1 - a0;
        // I_i + tau of ate_functions.R:255, term by term as written there
        const double J = (wi * yi) / pi - a1 * (wi - pi) / pi -
                         (((1.0 - wi) * yi / (1.0 - pi)) + (a0 * (wi - pi) / (1.0 - pi)));
        if (est1) est1[i] = e1;
        if (est2) est2[i] = e2;
        if (!isnan(e1)) {  // mean(est1, na.rm = TRUE): NaN dropped, +-Inf kept
            acc[0] += e1;
            acc[1] += 1.0;
        }
        acc[2] += e2;
        acc[3] += J;
        acc[4] += J * J;
    }
    block_sum<SCORE_NV>(acc, red);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < SCORE_NV; ++k) partials[(size_t)blockIdx.x * SCORE_NV + k] = acc[k];
        __threadfence();
        const unsigned int t = atomicAdd(ticket, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    // fold the per-block partials in block order
    double tot[SCORE_NV] = {0, 0, 0, 0, 0, 0};
    if (threadIdx.x < SCORE_NV) {
        double s = 0.0;
        for (unsigned int b = 0; b < gridDim.x; ++b)
            s += ((volatile double *)partials)[(size_t)b * SCORE_NV + threadIdx.x];
        red[threadIdx.x] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < SCORE_NV; ++k) tot[k] = red[k];
        const double dn = (double)n;
        const double tau = tot[0] / tot[1] + tot[2] / dn;
        // sum (J - tau)^2 = sum J^2 - 2 tau sum J + n tau^2
        double ss = tot[4] - 2.0 * tau * tot[3] + dn * tau * tau;
        if (ss < 0.0) ss = 0.0;
        out[0] = tau;
        out[1] = sqrt(ss) / dn;
        out[2] = tot[0];
        out[3] = tot[1];
        out[4] = tot[2];
        *ticket = 0u;  // ready for the next launch
    }
}

}  // namespace

// device pointers in; tau/se are host outputs (one small D2H)
int ate_i_score(ate_ctx *ctx, long long n, const double *w, const double *y, const double *p,
                const double *m0, const double *m1, double *d_est1, double *d_est2, double *tau,
                double *se) {
    if (n <= 0) return ate_fail(ctx, -1, "aipw score: n must be positive");
    int blocks = ctx->num_sms * 8;
    long long need = (n + SCORE_THREADS - 1) / SCORE_THREADS;
    if (need < blocks) blocks = (int)need;
    // scratch layout: [ticket (16 B)] [out 8 doubles] [partials blocks*NV]
    void *s = nullptr;
    ATE_TRY(ate_scratch(ctx, 16 + 64 + sizeof(double) * (size_t)blocks * SCORE_NV, &s));
    unsigned int *ticket = (unsigned int *)s;
    double *out = (double *)((char *)s + 16);
    double *partials = out + 8;
    ATE_CUDA(ctx, cudaMemsetAsync(ticket, 0, 16, ctx->stream));
    aipw_score_kernel<<<blocks, SCORE_THREADS, 0, ctx->stream>>>(n, w, y, p, m0, m1, d_est1, d_est2,
                                                                 partials, ticket, out);
    ATE_LAUNCHED(ctx);
    double h[2];
    ATE_CUDA(ctx, cudaMemcpyAsync(h, out, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (tau) *tau = h[0];
    if (se) *se = h[1];
    return 0;
}

extern "C" int ate_aipw_score(ate_ctx *ctx, long long n, const double *w, const double *y,
                              const double *p, const double *m0, const double *m1, double *tau,
                              double *se, double *est1, double *est2, int on_device) {
    if (!ctx) return -1;
    if (!w || !y || !p || !m0 || !m1 || n <= 0)
        return ate_fail(ctx, -1, "ate_aipw_score: NULL input or n <= 0");
    ATE_CUDA(ctx, cudaSetDevice(ctx->device));
    if (on_device) return ate_i_score(ctx, n, w, y, p, m0, m1, est1, est2, tau, se);
    DevBuf<double> d;
    const size_t N = (size_t)n;
    ATE_CUDA(ctx, d.alloc(N * 7));
    const double *src[5] = {w, y, p, m0, m1};
    for (int k = 0; k < 5; ++k)
        ATE_CUDA(ctx, cudaMemcpyAsync(d.p + k * N, src[k], N * sizeof(double),
                                      cudaMemcpyHostToDevice, ctx->stream));
    double *de1 = est1 ? d.p + 5 * N : nullptr, *de2 = est2 ? d.p + 6 * N : nullptr;
    ATE_TRY(ate_i_score(ctx, n, d.p, d.p + N, d.p + 2 * N, d.p + 3 * N, d.p + 4 * N, de1, de2, tau,
                        se));
    if (est1) ATE_CUDA(ctx, cudaMemcpyAsync(est1, de1, N * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    if (est2) ATE_CUDA(ctx, cudaMemcpyAsync(est2, de2, N * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}
