// bootstrap.cu -- a9/a10: bootstrap replicates of the AIPW mean with the nuisances held fixed
// (ate_functions.R:245-252 driver, :267-283 replicate).
//
// Because tau_hat_dr_est() does not refit anything, a replicate is
//   mean(est1[idx], na.rm = TRUE) + mean(est2[idx])
// so only the two score tables (est1, est2; 2*n*8 B, L2-resident) are gathered, not the five
// input vectors.  One CTA per replicate.
//   ATE_BOOT_INDEX : streams the caller's index matrix (B*n*4 B, the HBM-bound term).
//   ATE_BOOT_PHILOX: draws the indices on the fly (Philox4x32-10, counter = (pair, 0, rep)),
//                    two indices per Philox block, idx = mulhi64(r64, n).  The stream of a
//                    replicate depends only on (seed, replicate id).
#include <math.h>

#include "common.cuh"

namespace {

constexpr int BOOT_THREADS = 512;

struct Philox4 {
    uint32_t x, y, z, w;
};

__device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                 uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0;
        c1 = lo1;
        c2 = n2;
        c3 = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return Philox4{c0, c1, c2, c3};
}

template <bool TWO_TABLES>
__device__ __forceinline__ void take(const double *__restrict__ t1, const double *__restrict__ t2,
                                     long long i, double &s1, double &c1, double &s2) {
    if (TWO_TABLES) {
        const double a = __ldg(t1 + i), b = __ldg(t2 + i);
        if (!isnan(a)) {
            s1 += a;
            c1 += 1.0;
        }
        s2 += b;
    } else {
        s1 += __ldg(t1 + i);  // t1 = est1 + est2, no NaN present
    }
}

// MODE 0: index matrix, MODE 1: Philox
template <int MODE, bool TWO_TABLES>
__global__ void __launch_bounds__(BOOT_THREADS)
bootstrap_kernel(long long n, const double *__restrict__ t1, const double *__restrict__ t2,
                 const int32_t *__restrict__ idx, uint64_t seed, long long rep0,
                 double *__restrict__ boot) {
    __shared__ double red[3 * 32];
    const long long rep = rep0 + blockIdx.x;
    double acc[3] = {0.0, 0.0, 0.0};
    if (MODE == 0) {
        const int32_t *row = idx + (size_t)blockIdx.x * (size_t)n;
        long long t = threadIdx.x;
        // 4 independent gathers in flight per thread
        for (; t + 3LL * BOOT_THREADS < n; t += 4LL * BOOT_THREADS) {
            const int i0 = __ldg(row + t), i1 = __ldg(row + t + BOOT_THREADS);
            const int i2 = __ldg(row + t + 2 * BOOT_THREADS), i3 = __ldg(row + t + 3 * BOOT_THREADS);
            take<TWO_TABLES>(t1, t2, i0 - 1, acc[0], acc[1], acc[2]);
            take<TWO_TABLES>(t1, t2, i1 - 1, acc[0], acc[1], acc[2]);
            take<TWO_TABLES>(t1, t2, i2 - 1, acc[0], acc[1], acc[2]);
            take<TWO_TABLES>(t1, t2, i3 - 1, acc[0], acc[1], acc[2]);
        }
        for (; t < n; t += BOOT_THREADS) take<TWO_TABLES>(t1, t2, __ldg(row + t) - 1, acc[0], acc[1], acc[2]);
    } else {
        const long long npairs = (n + 1) >> 1;
        const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
        const uint32_t r0 = (uint32_t)rep, r1 = (uint32_t)((unsigned long long)rep >> 32);
        for (long long c = threadIdx.x; c < npairs; c += BOOT_THREADS) {
            const Philox4 q = philox4x32_10((uint32_t)c, (uint32_t)((unsigned long long)c >> 32), r0, r1, k0, k1);
            const unsigned long long ra = ((unsigned long long)q.y << 32) | q.x;
            const unsigned long long rb = ((unsigned long long)q.w << 32) | q.z;
            const long long ia = (long long)__umul64hi(ra, (unsigned long long)n);
            take<TWO_TABLES>(t1, t2, ia, acc[0], acc[1], acc[2]);
            if (2 * c + 1 < n) {
                const long long ib = (long long)__umul64hi(rb, (unsigned long long)n);
                take<TWO_TABLES>(t1, t2, ib, acc[0], acc[1], acc[2]);
            }
        }
    }
    block_sum<3>(acc, red);
    if (threadIdx.x == 0) {
        const double dn = (double)n;
        boot[blockIdx.x] = TWO_TABLES ? (acc[0] / acc[1] + acc[2] / dn) : acc[0] / dn;
    }
}

__global__ void psi_sum_kernel(long long n, const double *__restrict__ e1,
                               const double *__restrict__ e2, double *__restrict__ psi,
                               unsigned int *__restrict__ nan_flag) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    bool bad = false;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const double a = e1[i];
        bad |= isnan(a);
        psi[i] = a + e2[i];
    }
    if (bad) atomicOr(nan_flag, 1u);
}

}  // namespace

// d_est1/d_est2/d_idx are device pointers; boot_host[B] is a host buffer.
int ate_i_bootstrap(ate_ctx *ctx, long long n, const double *d_est1, const double *d_est2, int B,
                    int mode, const int32_t *d_idx, uint64_t seed, long long rep_offset,
                    double *boot_host) {
    if (B <= 0) return 0;
    if (mode == ATE_BOOT_INDEX && !d_idx) return ate_fail(ctx, -1, "bootstrap: index mode needs idx");
    DevBuf<double> psi, dboot;
    DevBuf<unsigned int> flag;
    ATE_CUDA(ctx, psi.alloc((size_t)n));
    ATE_CUDA(ctx, dboot.alloc((size_t)B));
    ATE_CUDA(ctx, flag.alloc(1));
    ATE_CUDA(ctx, cudaMemsetAsync(flag.p, 0, sizeof(unsigned int), ctx->stream));
    int blocks = ctx->num_sms * 4;
    psi_sum_kernel<<<blocks, 256, 0, ctx->stream>>>(n, d_est1, d_est2, psi.p, flag.p);
    ATE_LAUNCHED(ctx);
    unsigned int has_nan = 0;
    ATE_CUDA(ctx, cudaMemcpyAsync(&has_nan, flag.p, sizeof(has_nan), cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (mode == ATE_BOOT_INDEX) {
        if (has_nan)
            bootstrap_kernel<0, true><<<B, BOOT_THREADS, 0, ctx->stream>>>(n, d_est1, d_est2, d_idx, seed, rep_offset, dboot.p);
        else
            bootstrap_kernel<0, false><<<B, BOOT_THREADS, 0, ctx->stream>>>(n, psi.p, nullptr, d_idx, seed, rep_offset, dboot.p);
    } else {
        if (has_nan)
            bootstrap_kernel<1, true><<<B, BOOT_THREADS, 0, ctx->stream>>>(n, d_est1, d_est2, nullptr, seed, rep_offset, dboot.p);
        else
            bootstrap_kernel<1, false><<<B, BOOT_THREADS, 0, ctx->stream>>>(n, psi.p, nullptr, nullptr, seed, rep_offset, dboot.p);
    }
    ATE_LAUNCHED(ctx);
    ATE_CUDA(ctx, cudaMemcpyAsync(boot_host, dboot.p, sizeof(double) * (size_t)B, cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" int ate_bootstrap(ate_ctx *ctx, long long n, const double *w, const double *y,
                             const double *p, const double *m0, const double *m1, int B, int mode,
                             const int32_t *idx, uint64_t seed, long long rep_offset, double *boot,
                             double *se, int on_device) {
    if (!ctx) return -1;
    if (!w || !y || !p || !m0 || !m1 || !boot || n <= 0 || B <= 0)
        return ate_fail(ctx, -1, "ate_bootstrap: NULL input, n <= 0 or B <= 0");
    if (mode != ATE_BOOT_INDEX && mode != ATE_BOOT_PHILOX)
        return ate_fail(ctx, -1, "ate_bootstrap: unknown mode %d", mode);
    if (mode == ATE_BOOT_INDEX && !idx)
        return ate_fail(ctx, -1, "ate_bootstrap: ATE_BOOT_INDEX needs the B x n index matrix");
    ATE_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t N = (size_t)n;
    DevBuf<double> d, e;
    DevBuf<int32_t> di;
    const double *dw = w, *dy = y, *dp = p, *d0 = m0, *d1 = m1;
    const int32_t *didx = idx;
    if (!on_device) {
        ATE_CUDA(ctx, d.alloc(N * 5));
        const double *src[5] = {w, y, p, m0, m1};
        for (int k = 0; k < 5; ++k)
            ATE_CUDA(ctx, cudaMemcpyAsync(d.p + k * N, src[k], N * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
        dw = d.p; dy = d.p + N; dp = d.p + 2 * N; d0 = d.p + 3 * N; d1 = d.p + 4 * N;
        if (mode == ATE_BOOT_INDEX) {
            ATE_CUDA(ctx, di.alloc(N * (size_t)B));
            ATE_CUDA(ctx, cudaMemcpyAsync(di.p, idx, N * (size_t)B * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
            didx = di.p;
        }
    }
    ATE_CUDA(ctx, e.alloc(N * 2));
    double tau, se0;
    ATE_TRY(ate_i_score(ctx, n, dw, dy, dp, d0, d1, e.p, e.p + N, &tau, &se0));
    ATE_TRY(ate_i_bootstrap(ctx, n, e.p, e.p + N, B, mode, didx, seed, rep_offset, boot));
    if (se) *se = ate_host_sd(boot, B);
    return 0;
}

extern "C" int ate_tau_hat_dr_est(ate_ctx *ctx, long long n, const double *w, const double *y,
                                  const double *p, const double *m0, const double *m1,
                                  const int32_t *idx, double *tau) {
    if (!tau) return ate_fail(ctx, -1, "ate_tau_hat_dr_est: NULL output");
    return ate_bootstrap(ctx, n, w, y, p, m0, m1, 1, ATE_BOOT_INDEX, idx, 0, 0, tau, nullptr, 0);
}
