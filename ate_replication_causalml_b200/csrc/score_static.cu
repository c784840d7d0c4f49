// score_static.cu -- ate_aipw_score_static(): the AIPW score / ATE / sandwich-SE kernel for
// callers that must not allocate device memory.
//
// Who that is: code that runs with the whole GPU already handed to someone else's allocator
// (a framework memory pool sized to the device, an R session sharing the card with a job that
// reserved it), and code on a latency-critical path that cannot afford cudaMalloc/cudaFree,
// which synchronise the device.  ate_aipw_score() keeps its block partials in a scratch buffer
// that the context grows with cudaMalloc on demand; this variant keeps them in a fixed array
// that is part of the library image instead, so a call performs no allocation of any kind.
//
// The price, stated once:
//   * the array below (ATE_STATIC_SCORE_BYTES, 48.1 KB) is resident on every GPU that loads the
//     library, until the process exits;
//   * n is limited to ATE_STATIC_SCORE_MAX_N rows (2^28): the kernel runs at most
//     ATE_STATIC_SCORE_MAX_BLOCKS blocks, each leaving six partial sums in the array;
//   * all calls on a device share the one array.  It keeps whatever the previous call left, so the
//     entry point clears the completion ticket before every launch; two calls that overlap on
//     the same GPU (different streams or threads) race on it and must be serialised by the caller;
//   * the first call on each device looks up the array's device address; make that first call
//     before any stream capture begins.
#include <math.h>

#include "score_core.cuh"

#define ATE_STATIC_SCORE_MAX_BLOCKS 1024
#define ATE_STATIC_SCORE_MAX_N (1LL << 28)
#define ATE_STATIC_SCORE_NV 6  /* = score_core::NV */
// layout: [ticket: 16 B][result: 8 doubles][partials: MAX_BLOCKS x NV doubles]
#define ATE_STATIC_SCORE_BYTES (16 + 64 + 8 * ATE_STATIC_SCORE_NV * ATE_STATIC_SCORE_MAX_BLOCKS)

// In host code this name is only a handle for cudaGetSymbolAddress(); it is not a device address.
static __device__ __align__(256) unsigned char g_score_scratch[ATE_STATIC_SCORE_BYTES];

namespace {

constexpr int ST_THREADS = 256;
constexpr int ST_MAX_DEVICES = 64;
void *g_scratch_addr[ST_MAX_DEVICES] = {nullptr};  // device address of the array, per device

// the same device code as aipw_score_kernel (score_core.cuh); only where the partials live differs
__global__ void __launch_bounds__(ST_THREADS)
aipw_score_static_kernel(long long n, const double *__restrict__ w, const double *__restrict__ y,
                         const double *__restrict__ p, const double *__restrict__ m0,
                         const double *__restrict__ m1, double *__restrict__ partials,
                         unsigned int *__restrict__ ticket, double *__restrict__ out) {
    __shared__ double red[score_core::NV * 32];
    __shared__ bool is_last;
    double acc[score_core::NV] = {0, 0, 0, 0, 0, 0};
    score_core::accumulate<false>(n, w, y, p, m0, m1, nullptr, nullptr, nullptr, acc);
    score_core::finish(acc, red, &is_last, n, partials, ticket, out);
}

}  // namespace

extern "C" int ate_aipw_score_static(ate_ctx *ctx, long long n, const double *d_w, const double *d_y,
                                     const double *d_p, const double *d_m0, const double *d_m1,
                                     double *tau, double *se) {
    if (!ctx) return -1;
    if (!d_w || !d_y || !d_p || !d_m0 || !d_m1 || !tau || !se || n <= 0)
        return ate_fail(ctx, -1, "ate_aipw_score_static: NULL argument or n <= 0");
    if (n > ATE_STATIC_SCORE_MAX_N)
        return ate_fail(ctx, -1,
                        "ate_aipw_score_static: n = %lld exceeds the static-scratch limit of %lld rows; "
                        "use ate_aipw_score (which allocates) for larger inputs",
                        n, (long long)ATE_STATIC_SCORE_MAX_N);
    if (ctx->device < 0 || ctx->device >= ST_MAX_DEVICES)
        return ate_fail(ctx, -1, "ate_aipw_score_static: device ordinal %d not supported", ctx->device);
    ATE_CUDA(ctx, cudaSetDevice(ctx->device));
    // one-time set-up per device, with that device current: resolve the array's device address
    if (!g_scratch_addr[ctx->device]) {
        void *addr = nullptr;
        ATE_CUDA(ctx, cudaGetSymbolAddress(&addr, g_score_scratch));
        if (!addr) return ate_fail(ctx, -100, "ate_aipw_score_static: cudaGetSymbolAddress returned NULL");
        g_scratch_addr[ctx->device] = addr;
    }
    unsigned char *base = static_cast<unsigned char *>(g_scratch_addr[ctx->device]);
    unsigned int *ticket = reinterpret_cast<unsigned int *>(base);
    double *out = reinterpret_cast<double *>(base + 16);
    double *partials = out + 8;
    int blocks = ctx->num_sms * 8;
    if (blocks > ATE_STATIC_SCORE_MAX_BLOCKS) blocks = ATE_STATIC_SCORE_MAX_BLOCKS;
    const long long need = (n + ST_THREADS - 1) / ST_THREADS;
    if (need < blocks) blocks = (int)need;
    // the array still holds the last call's ticket and results: the ticket must start at zero
    ATE_CUDA(ctx, cudaMemsetAsync(ticket, 0, 16 + 64, ctx->stream));
    aipw_score_static_kernel<<<blocks, ST_THREADS, 0, ctx->stream>>>(n, d_w, d_y, d_p, d_m0, d_m1, partials,
                                                                      ticket, out);
    ATE_LAUNCHED(ctx);
    double h[2];
    ATE_CUDA(ctx, cudaMemcpyAsync(h, out, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *tau = h[0];
    *se = h[1];
    return 0;
}
