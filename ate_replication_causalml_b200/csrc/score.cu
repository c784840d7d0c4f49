// score.cu -- a7/a8: fused AIPW score, ATE and sandwich SE (ate_functions.R:241-243,255-256).
//
// One HBM pass over the five n-vectors (w, y, p, m0, m1): 5*n*8 B read, plus 2*n*8 B written
// when est1/est2 are materialised for the bootstrap.  Each block leaves six partial sums in
// scratch; the last block to finish folds them in block order (deterministic) and writes
// tau and se, so score + fold reduction are a single launch.
#include <math.h>

#include "score_core.cuh"

namespace {

constexpr int SCORE_THREADS = 256;

template <bool VEC>
__global__ void __launch_bounds__(SCORE_THREADS)
aipw_score_kernel(long long n, const double *__restrict__ w, const double *__restrict__ y,
                  const double *__restrict__ p, const double *__restrict__ m0,
                  const double *__restrict__ m1, double *__restrict__ est1,
                  double *__restrict__ est2, double *__restrict__ psi, double *__restrict__ partials,
                  unsigned int *__restrict__ ticket, double *__restrict__ out) {
    __shared__ double red[score_core::NV * 32];
    __shared__ bool is_last;
    double acc[score_core::NV] = {0, 0, 0, 0, 0, 0};
    score_core::accumulate<VEC>(n, w, y, p, m0, m1, est1, est2, psi, acc);
    score_core::finish(acc, red, &is_last, n, partials, ticket, out);
}

}  // namespace

// device pointers in; tau/se are host outputs (one small D2H).  d_psi (optional): est1 + est2, the
// single table the bootstrap gathers when no est1 is NaN; *n_ok (optional): count of non-NaN est1.
int ate_i_score(ate_ctx *ctx, long long n, const double *w, const double *y, const double *p,
                const double *m0, const double *m1, double *d_est1, double *d_est2, double *tau,
                double *se, double *d_psi, double *n_ok) {
    if (n <= 0) return ate_fail(ctx, -1, "aipw score: n must be positive");
    int blocks = ctx->num_sms * 8;
    long long need = (n + 2 * SCORE_THREADS - 1) / (2 * SCORE_THREADS);
    if (need < blocks) blocks = (int)(need < 1 ? 1 : need);
    // scratch layout: [ticket (16 B)] [out 8 doubles] [partials blocks*NV]
    void *s = nullptr;
    ATE_TRY(ate_scratch(ctx, 16 + 64 + sizeof(double) * (size_t)blocks * score_core::NV, &s));
    unsigned int *ticket = (unsigned int *)s;
    double *out = (double *)((char *)s + 16);
    double *partials = out + 8;
    ATE_CUDA(ctx, cudaMemsetAsync(ticket, 0, 16, ctx->stream));
    if (score_core::aligned16(w, y, p, m0, m1, d_est1, d_est2, d_psi))
        aipw_score_kernel<true><<<blocks, SCORE_THREADS, 0, ctx->stream>>>(n, w, y, p, m0, m1, d_est1, d_est2, d_psi,
                                                                           partials, ticket, out);
    else
        aipw_score_kernel<false><<<blocks, SCORE_THREADS, 0, ctx->stream>>>(n, w, y, p, m0, m1, d_est1, d_est2, d_psi,
                                                                            partials, ticket, out);
    ATE_LAUNCHED(ctx);
    double h[4];
    ATE_CUDA(ctx, cudaMemcpyAsync(h, out, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (tau) *tau = h[0];
    if (se) *se = h[1];
    if (n_ok) *n_ok = h[3];
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
    const size_t N = ((size_t)n + 1) / 2 * 2;  // vectors start 16-byte aligned
    ATE_CUDA(ctx, d.alloc(N * 7));
    const double *src[5] = {w, y, p, m0, m1};
    for (int k = 0; k < 5; ++k)
        ATE_CUDA(ctx, cudaMemcpyAsync(d.p + k * N, src[k], (size_t)n * sizeof(double),
                                      cudaMemcpyHostToDevice, ctx->stream));
    double *de1 = est1 ? d.p + 5 * N : nullptr, *de2 = est2 ? d.p + 6 * N : nullptr;
    ATE_TRY(ate_i_score(ctx, n, d.p, d.p + N, d.p + 2 * N, d.p + 3 * N, d.p + 4 * N, de1, de2, tau,
                        se));
    if (est1) ATE_CUDA(ctx, cudaMemcpyAsync(est1, de1, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    if (est2) ATE_CUDA(ctx, cudaMemcpyAsync(est2, de2, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}
