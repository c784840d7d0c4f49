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

#include <algorithm>

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

// one validated gather: a 1-based index outside 1..n (R would give NA for it) raises the flag and
// contributes nothing
template <bool TWO_TABLES>
__device__ __forceinline__ void take_checked(const double *__restrict__ t1, const double *__restrict__ t2, int idx1,
                                             long long n, double &s1, double &c1, double &s2, int &bad) {
    const long long i = (long long)idx1 - 1;
    if (i < 0 || i >= n)
        bad = 1;
    else
        take<TWO_TABLES>(t1, t2, i, s1, c1, s2);
}

// MODE 0: index matrix (HBM-bound: the B x n int32 stream is read once, 16 bytes per load where the
// replicate's row is 16-byte aligned; the score tables are L2-resident), MODE 1: Philox
template <int MODE, bool TWO_TABLES>
__global__ void __launch_bounds__(BOOT_THREADS)
bootstrap_kernel(long long n, const double *__restrict__ t1, const double *__restrict__ t2,
                 const int32_t *__restrict__ idx, uint64_t seed, long long rep0,
                 double *__restrict__ boot, int *__restrict__ bad_flag) {
    __shared__ double red[3 * 32];
    const long long rep = rep0 + blockIdx.x;
    double acc[3] = {0.0, 0.0, 0.0};
    if (MODE == 0) {
        const int32_t *row = idx + (size_t)blockIdx.x * (size_t)n;
        int bad = 0;
        // peel to a 16-byte boundary, then int4 loads (two in flight per thread), then the tail
        long long head = (long long)(((16 - ((uintptr_t)row & 15)) & 15) >> 2);
        if (head > n) head = n;
        for (long long t = threadIdx.x; t < head; t += BOOT_THREADS)
            take_checked<TWO_TABLES>(t1, t2, __ldg(row + t), n, acc[0], acc[1], acc[2], bad);
        const int4 *row4 = reinterpret_cast<const int4 *>(row + head);
        const long long n4 = (n - head) >> 2;
        long long q = threadIdx.x;
        for (; q + BOOT_THREADS < n4; q += 2LL * BOOT_THREADS) {
            const int4 a = __ldg(row4 + q), b = __ldg(row4 + q + BOOT_THREADS);
            take_checked<TWO_TABLES>(t1, t2, a.x, n, acc[0], acc[1], acc[2], bad);
            take_checked<TWO_TABLES>(t1, t2, a.y, n, acc[0], acc[1], acc[2], bad);
            take_checked<TWO_TABLES>(t1, t2, a.z, n, acc[0], acc[1], acc[2], bad);
            take_checked<TWO_TABLES>(t1, t2, a.w, n, acc[0], acc[1], acc[2], bad);
            take_checked<TWO_TABLES>(t1, t2, b.x, n, acc[0], acc[1], acc[2], bad);
            take_checked<TWO_TABLES>(t1, t2, b.y, n, acc[0], acc[1], acc[2], bad);
            take_checked<TWO_TABLES>(t1, t2, b.z, n, acc[0], acc[1], acc[2], bad);
            take_checked<TWO_TABLES>(t1, t2, b.w, n, acc[0], acc[1], acc[2], bad);
        }
        for (; q < n4; q += BOOT_THREADS) {
            const int4 a = __ldg(row4 + q);
            take_checked<TWO_TABLES>(t1, t2, a.x, n, acc[0], acc[1], acc[2], bad);
            take_checked<TWO_TABLES>(t1, t2, a.y, n, acc[0], acc[1], acc[2], bad);
            take_checked<TWO_TABLES>(t1, t2, a.z, n, acc[0], acc[1], acc[2], bad);
            take_checked<TWO_TABLES>(t1, t2, a.w, n, acc[0], acc[1], acc[2], bad);
        }
        for (long long t = head + 4 * n4 + threadIdx.x; t < n; t += BOOT_THREADS)
            take_checked<TWO_TABLES>(t1, t2, __ldg(row + t), n, acc[0], acc[1], acc[2], bad);
        if (bad) atomicOr(bad_flag, 1);
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

// d_est1 / d_est2: device score tables.  idx (index mode): the B x n matrix of 1-based indices, on the
// device, or on the host (idx_on_host) in which case it is streamed through two device blocks of
// <= 256 MB, the upload of one overlapping the kernel on the other -- the matrix (8 GB at C1, 40 GB at
// C2) is never resident as a whole.  Results: d_out (device, optional) and boot_host (host, optional).
// Index mode validates every index: one outside 1..n fails the call.
// d_psi (optional): est1 + est2 as the score kernel already wrote it, with has_nan_known = whether any
// est1 is NaN; without it one extra pass builds the table here.
int ate_i_bootstrap(ate_ctx *ctx, long long n, const double *d_est1, const double *d_est2, int B,
                    int mode, const int32_t *idx, int idx_on_host, uint64_t seed, long long rep_offset,
                    double *d_out, double *boot_host, const double *d_psi, int has_nan_known) {
    if (B <= 0) return 0;
    if (mode == ATE_BOOT_INDEX && !idx) return ate_fail(ctx, -1, "bootstrap: index mode needs idx");
    void *blk = nullptr;
    ATE_TRY(ate_cached(ctx, "boot.state", sizeof(double) * ((size_t)n + (size_t)B) + 256, &blk));
    double *psi_own = (double *)blk, *dboot = psi_own + n;
    unsigned int *flag = (unsigned int *)(dboot + B);
    int *bad = (int *)(flag + 1);
    ATE_CUDA(ctx, cudaMemsetAsync(flag, 0, 2 * sizeof(int), ctx->stream));
    unsigned int has_nan = 0;
    const double *psi = d_psi;
    if (d_psi) {
        has_nan = has_nan_known ? 1u : 0u;
    } else {
        psi_sum_kernel<<<ctx->num_sms * 4, 256, 0, ctx->stream>>>(n, d_est1, d_est2, psi_own, flag);
        ATE_LAUNCHED(ctx);
        ATE_CUDA(ctx, cudaMemcpyAsync(&has_nan, flag, sizeof(has_nan), cudaMemcpyDeviceToHost, ctx->stream));
        ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        psi = psi_own;
    }
    auto launch = [&](int nb, const int32_t *didx, long long rep0, double *out, cudaStream_t st) {
        if (mode == ATE_BOOT_INDEX) {
            if (has_nan)
                bootstrap_kernel<0, true><<<nb, BOOT_THREADS, 0, st>>>(n, d_est1, d_est2, didx, seed, rep0, out, bad);
            else
                bootstrap_kernel<0, false><<<nb, BOOT_THREADS, 0, st>>>(n, psi, nullptr, didx, seed, rep0, out, bad);
        } else {
            if (has_nan)
                bootstrap_kernel<1, true><<<nb, BOOT_THREADS, 0, st>>>(n, d_est1, d_est2, nullptr, seed, rep0, out, bad);
            else
                bootstrap_kernel<1, false><<<nb, BOOT_THREADS, 0, st>>>(n, psi, nullptr, nullptr, seed, rep0, out, bad);
        }
    };
    if (mode == ATE_BOOT_INDEX && idx_on_host) {
        long long per = (256LL << 20) / (4 * n);
        if (per < 1) per = 1;
        if (per > B) per = B;
        void *ib = nullptr;
        ATE_TRY(ate_cached(ctx, "boot.idx", sizeof(int32_t) * (size_t)n * (size_t)per * 2, &ib));
        int32_t *dblk[2] = {(int32_t *)ib, (int32_t *)ib + (size_t)n * per};
        cudaStream_t cs = nullptr;
        cudaEvent_t up[2] = {nullptr, nullptr}, used[2] = {nullptr, nullptr};
        ATE_CUDA(ctx, cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
        for (int k = 0; k < 2; ++k) {
            cudaEventCreateWithFlags(&up[k], cudaEventDisableTiming);
            cudaEventCreateWithFlags(&used[k], cudaEventDisableTiming);
        }
        int rc = 0, turn = 0;
        for (long long b0 = 0; b0 < B && rc == 0; b0 += per, turn ^= 1) {
            const int nb = (int)std::min<long long>(per, B - b0);
            if (b0 >= 2 * per) cudaStreamWaitEvent(cs, used[turn], 0);  // the kernel that read this block is done
            if (cudaMemcpyAsync(dblk[turn], idx + (size_t)b0 * (size_t)n, sizeof(int32_t) * (size_t)n * nb, cudaMemcpyHostToDevice, cs) != cudaSuccess) rc = -1;
            cudaEventRecord(up[turn], cs);
            cudaStreamWaitEvent(ctx->stream, up[turn], 0);
            launch(nb, dblk[turn], rep_offset + b0, dboot + b0, ctx->stream);
            ctx->launches++;
            if (cudaGetLastError() != cudaSuccess) rc = -1;
            cudaEventRecord(used[turn], ctx->stream);
        }
        cudaStreamSynchronize(cs);
        cudaStreamSynchronize(ctx->stream);
        for (int k = 0; k < 2; ++k) {
            cudaEventDestroy(up[k]);
            cudaEventDestroy(used[k]);
        }
        cudaStreamDestroy(cs);
        if (rc != 0) return ate_fail(ctx, -100, "bootstrap: streaming the index matrix failed");
    } else {
        launch(B, idx, rep_offset, dboot, ctx->stream);
        ATE_LAUNCHED(ctx);
    }
    int hbad = 0;
    if (mode == ATE_BOOT_INDEX) ATE_CUDA(ctx, cudaMemcpyAsync(&hbad, bad, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    if (d_out) ATE_CUDA(ctx, cudaMemcpyAsync(d_out, dboot, sizeof(double) * (size_t)B, cudaMemcpyDeviceToDevice, ctx->stream));
    if (boot_host) ATE_CUDA(ctx, cudaMemcpyAsync(boot_host, dboot, sizeof(double) * (size_t)B, cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (hbad) return ate_fail(ctx, -1, "bootstrap: an index lies outside 1..%lld (R's sample() never produces one; NA_integer_ is not accepted)", n);
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
    void *blk = nullptr;
    ATE_TRY(ate_cached(ctx, "boot.inputs", sizeof(double) * N * 7, &blk));
    double *d = (double *)blk, *e = d + 5 * N;
    const double *dw = w, *dy = y, *dp = p, *d0 = m0, *d1 = m1;
    if (!on_device) {
        const double *src[5] = {w, y, p, m0, m1};
        for (int k = 0; k < 5; ++k)
            ATE_CUDA(ctx, cudaMemcpyAsync(d + k * N, src[k], N * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
        dw = d; dy = d + N; dp = d + 2 * N; d0 = d + 3 * N; d1 = d + 4 * N;
    }
    double tau, se0;
    ATE_TRY(ate_i_score(ctx, n, dw, dy, dp, d0, d1, e, e + N, &tau, &se0));
    // boot is a host buffer in both modes
    ATE_TRY(ate_i_bootstrap(ctx, n, e, e + N, B, mode, idx, !on_device, seed, rep_offset, nullptr, boot, nullptr, 0));
    if (se) *se = ate_host_sd(boot, B);
    return 0;
}

extern "C" int ate_tau_hat_dr_est(ate_ctx *ctx, long long n, const double *w, const double *y,
                                  const double *p, const double *m0, const double *m1,
                                  const int32_t *idx, double *tau) {
    if (!tau) return ate_fail(ctx, -1, "ate_tau_hat_dr_est: NULL output");
    return ate_bootstrap(ctx, n, w, y, p, m0, m1, 1, ATE_BOOT_INDEX, idx, 0, 0, tau, nullptr, 0);
}
