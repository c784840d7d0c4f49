// folds.cu -- the fold structure of the cross-fitted composite, built on the device.
//
// cv.glmnet works on `foldid`; the composite needs, for every outer
// fold k, the training rows grouped by their inner cv.glmnet fold (one list per nuisance) and the
// held-out rows.  The inner ids are given per TRAINING-ROW POSITION (row k of inner_ps / inner_out:
// "the ids cv.glmnet would draw for the training rows of outer fold k, in training-row order"), so
// row i's inner id sits at the number of training rows before i -- a prefix count.  Everything here is
// one primitive, a stable counting sort by a small key (rank_by_key: for every element the number of
// earlier elements with the same key), applied to "is a training row" and to the inner ids.
// Integer work, HBM-bound, a few passes over n ints; the host only reads the class totals.
#include "folds.cuh"

namespace {

constexpr int RK_THREADS = 256;
constexpr int RK_SUB = 16;                      // sub-rounds per block
constexpr int RK_BLOCK = RK_THREADS * RK_SUB;   // elements per block
constexpr int RK_MAXC = 64;                     // classes 1..C, C <= RK_MAXC

__global__ void __launch_bounds__(RK_THREADS)
rk_hist_kernel(const int32_t *__restrict__ keys, long long n, int C, int *__restrict__ blockhist, int *__restrict__ bad) {
    __shared__ int h[RK_MAXC + 1];
    for (int c = threadIdx.x; c <= C; c += blockDim.x) h[c] = 0;
    __syncthreads();
    const long long base = (long long)blockIdx.x * RK_BLOCK;
    for (int s = 0; s < RK_SUB; ++s) {
        const long long i = base + (long long)s * RK_THREADS + threadIdx.x;
        if (i < n) {
            const int k = keys[i];
            if (k < 0 || k > C)
                atomicOr(bad, 1);
            else
                atomicAdd(&h[k], 1);
        }
    }
    __syncthreads();
    for (int c = threadIdx.x; c <= C; c += blockDim.x) blockhist[(size_t)blockIdx.x * (C + 1) + c] = h[c];
}

// exclusive prefix over the blocks, per class; totals[c] = class size
__global__ void rk_scan_kernel(int *__restrict__ blockhist, int nblocks, int C, long long *__restrict__ totals) {
    const int c = threadIdx.x;
    if (c > C) return;
    long long run = 0;
    for (int b = 0; b < nblocks; ++b) {
        const int v = blockhist[(size_t)b * (C + 1) + c];
        blockhist[(size_t)b * (C + 1) + c] = (int)run;
        run += v;
    }
    totals[c] = run;
}

__global__ void __launch_bounds__(RK_THREADS)
rk_rank_kernel(const int32_t *__restrict__ keys, long long n, int C, const int *__restrict__ blockbase,
               int32_t *__restrict__ pos) {
    __shared__ int base[RK_MAXC + 1];
    __shared__ int wcnt[RK_THREADS / 32][RK_MAXC + 1];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int c = threadIdx.x; c <= C; c += blockDim.x) base[c] = blockbase[(size_t)blockIdx.x * (C + 1) + c];
    const long long b0 = (long long)blockIdx.x * RK_BLOCK;
    for (int s = 0; s < RK_SUB; ++s) {
        for (int e = threadIdx.x; e < (RK_THREADS / 32) * (RK_MAXC + 1); e += blockDim.x) (&wcnt[0][0])[e] = 0;
        __syncthreads();
        const long long i = b0 + (long long)s * RK_THREADS + threadIdx.x;
        int k = -1;
        if (i < n) {
            k = keys[i];
            if (k < 0 || k > C) k = -1;
        }
        const unsigned peers = __match_any_sync(0xffffffffu, k);
        const int myrank = __popc(peers & ((1u << lane) - 1u));
        if (k >= 0 && myrank == 0) wcnt[w][k] = __popc(peers);
        __syncthreads();
        if (k >= 0) {
            int off = base[k] + myrank;
            for (int q = 0; q < w; ++q) off += wcnt[q][k];
            pos[i] = off;
        }
        __syncthreads();
        for (int c = threadIdx.x; c <= C; c += blockDim.x) {
            int t = 0;
            for (int q = 0; q < RK_THREADS / 32; ++q) t += wcnt[q][c];
            base[c] += t;
        }
        __syncthreads();
    }
}

// key[i] = (fold[i] != k) ? 1 : 0 over all rows; flags ids outside 1..K
__global__ void train_key_kernel(const int32_t *__restrict__ fold, long long n, int k, int K, int32_t *__restrict__ key,
                                 int *__restrict__ bad) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int f = fold[i];
        if (f < 1 || f > K) atomicOr(bad, 1);
        key[i] = (f != k) ? 1 : 0;
    }
}

// for every row: key = inner id of the row (read at its training position) or 0 for held-out rows;
// key_h = 1 for held-out rows
__global__ void inner_key_kernel(const int32_t *__restrict__ fold, const int32_t *__restrict__ tpos, long long n, int k,
                                 const int32_t *__restrict__ inner_a, const int32_t *__restrict__ inner_b, int F,
                                 int32_t *__restrict__ key_a, int32_t *__restrict__ key_b, int32_t *__restrict__ key_h,
                                 int *__restrict__ bad) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const bool train = fold[i] != k;
        int a = 0, b = 0;
        if (train) {
            const long long t = tpos[i];
            a = inner_a[t];
            b = inner_b[t];
            if (a < 1 || a > F || b < 1 || b > F) {
                atomicOr(bad, 2);
                a = b = 0;
            }
        }
        key_a[i] = a;
        key_b[i] = b;
        key_h[i] = train ? 0 : 1;
    }
}

// list[segoff[key] + pos] = local row index, for keys 1..C
__global__ void scatter_kernel(const int32_t *__restrict__ key, const int32_t *__restrict__ pos, long long n, int C,
                               const long long *__restrict__ totals, int32_t *__restrict__ list) {
    __shared__ long long off[RK_MAXC + 2];
    if (threadIdx.x == 0) {
        long long run = 0;
        for (int c = 1; c <= C; ++c) {
            off[c] = run;
            run += totals[c];
        }
    }
    __syncthreads();
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int k = key[i];
        if (k >= 1 && k <= C) list[off[k] + pos[i]] = (int32_t)i;
    }
}

}  // namespace

int ate_i_rank_by_key(ate_ctx *ctx, const int32_t *d_keys, long long n, int C, int32_t *d_pos, int *d_blockhist,
                      long long *d_totals, int *d_bad) {
    if (C < 1 || C > RK_MAXC) return ate_fail(ctx, -1, "rank_by_key: %d classes unsupported", C);
    if (n <= 0) {
        ATE_CUDA(ctx, cudaMemsetAsync(d_totals, 0, sizeof(long long) * (C + 1), ctx->stream));
        return 0;
    }
    const int nb = (int)((n + RK_BLOCK - 1) / RK_BLOCK);
    rk_hist_kernel<<<nb, RK_THREADS, 0, ctx->stream>>>(d_keys, n, C, d_blockhist, d_bad);
    ATE_LAUNCHED(ctx);
    rk_scan_kernel<<<1, RK_MAXC + 1, 0, ctx->stream>>>(d_blockhist, nb, C, d_totals);
    ATE_LAUNCHED(ctx);
    if (d_pos) {
        rk_rank_kernel<<<nb, RK_THREADS, 0, ctx->stream>>>(d_keys, n, C, d_blockhist, d_pos);
        ATE_LAUNCHED(ctx);
    }
    return 0;
}

// stable group-by: list = indices 0..n-1 ordered by key (keys 1..C; key 0 skipped), d_totals[c] = class sizes
int ate_i_group_by_key(ate_ctx *ctx, const int32_t *d_keys, long long n, int C, int32_t *d_pos, int *d_blockhist,
                       long long *d_totals, int *d_bad, int32_t *d_list) {
    ATE_TRY(ate_i_rank_by_key(ctx, d_keys, n, C, d_pos, d_blockhist, d_totals, d_bad));
    if (n > 0) {
        scatter_kernel<<<ctx->num_sms * 4, 256, 0, ctx->stream>>>(d_keys, d_pos, n, C, d_totals, d_list);
        ATE_LAUNCHED(ctx);
    }
    return 0;
}

size_t ate_i_rank_scratch_ints(long long n, int C) {
    const long long nb = (n + RK_BLOCK - 1) / RK_BLOCK + 1;
    return (size_t)nb * (size_t)(C + 1);
}

// Fold structure of outer fold k for the local rows [lo, hi).  d_fold: n outer ids; d_inner_a / _b: the
// two nuisances' inner ids of fold k's training rows (training-row order, length >= nt).
// Out (device): list_a / list_b (local row indices of the local training rows, grouped by inner fold),
// list_h (local row indices of the local held-out rows).  d_tot[0 .. 4F + 2) (device; the caller copies it
// back): [0] training rows over all ranks, [1 .. F] local sizes of list_a's segments, [F + 1 .. 2F] of
// list_b's, [2F + 1] local held-out rows, [2F + 2 .. 3F + 1] inner-fold sizes over all ranks (a),
// [3F + 2 .. 4F + 1] (b).
int ate_i_build_fold_lists(ate_ctx *ctx, const int32_t *d_fold, long long n, long long lo, long long hi, int k, int K,
                           const int32_t *d_inner_a, const int32_t *d_inner_b, int F, FoldScratch &sc,
                           int32_t *list_a, int32_t *list_b, int32_t *list_h, long long *d_tot) {
    const long long nloc = hi - lo;
    const int blocks = ctx->num_sms * 4;
    const bool whole = (lo == 0 && hi == n);
    auto keep = [&](long long *dst, int count) -> int {
        ATE_CUDA(ctx, cudaMemcpyAsync(dst, sc.totals + 1, sizeof(long long) * count, cudaMemcpyDeviceToDevice, ctx->stream));
        return 0;
    };
    train_key_kernel<<<blocks, 256, 0, ctx->stream>>>(d_fold, n, k, K, sc.key_n, sc.bad);
    ATE_LAUNCHED(ctx);
    ATE_TRY(ate_i_rank_by_key(ctx, sc.key_n, n, 1, sc.pos_n, sc.blockhist, sc.totals, sc.bad));
    ATE_TRY(keep(d_tot, 1));
    inner_key_kernel<<<blocks, 256, 0, ctx->stream>>>(d_fold, sc.pos_n, n, k, d_inner_a, d_inner_b, F, sc.key_a, sc.key_b,
                                                       sc.key_h, sc.bad);
    ATE_LAUNCHED(ctx);
    const int32_t *keys[2] = {sc.key_a, sc.key_b};
    int32_t *lists[2] = {list_a, list_b};
    for (int q = 0; q < 2; ++q) {
        ATE_TRY(ate_i_rank_by_key(ctx, keys[q] + lo, nloc, F, sc.pos_loc, sc.blockhist, sc.totals, sc.bad));
        scatter_kernel<<<blocks, 256, 0, ctx->stream>>>(keys[q] + lo, sc.pos_loc, nloc, F, sc.totals, lists[q]);
        ATE_LAUNCHED(ctx);
        ATE_TRY(keep(d_tot + 1 + q * F, F));
        if (!whole) ATE_TRY(ate_i_rank_by_key(ctx, keys[q], n, F, nullptr, sc.blockhist, sc.totals, sc.bad));
        ATE_TRY(keep(d_tot + 2 + 2 * F + q * F, F));
    }
    ATE_TRY(ate_i_rank_by_key(ctx, sc.key_h + lo, nloc, 1, sc.pos_loc, sc.blockhist, sc.totals, sc.bad));
    scatter_kernel<<<blocks, 256, 0, ctx->stream>>>(sc.key_h + lo, sc.pos_loc, nloc, 1, sc.totals, list_h);
    ATE_LAUNCHED(ctx);
    ATE_TRY(keep(d_tot + 1 + 2 * F, 1));
    return 0;
}
