// common.cuh -- internal plumbing shared by the translation units of libate_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/ate_b200.h"

#define ATE_NUM_SMS_FALLBACK 148

struct ate_rng_state {
    uint32_t mt[624];
    int mti;
    bool seeded;
};

struct ate_ctx {
    int device = 0;
    int num_sms = ATE_NUM_SMS_FALLBACK;
    cudaStream_t stream = nullptr;
    std::string err;
    long long launches = 0;
    // per-launch CUDA-event timing: bit 1 = the IRLS step kernel (irls_round.cu), bit 2 = the Gram kernel
    int profile = 0;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events;
    std::vector<std::pair<int, long long>> prof_shape;  // (columns in the Gram, rows) per timed launch
    double prof_flops = 0.0, prof_bytes = 0.0;
    long long prof_launches = 0;
    // small reusable device scratch (block partials, flags)
    void *scratch = nullptr;
    size_t scratch_bytes = 0;
    // gram partial-sum scratch
    void *gram_scratch = nullptr;
    size_t gram_scratch_bytes = 0;
    ate_rng_state rng;
    // large work buffers kept across calls (cudaMalloc/cudaFree of multi-GB blocks cost ~1 s per fit)
    std::map<std::string, std::pair<void *, size_t>> cache;
    std::map<std::string, std::pair<void *, size_t>> pinned;  // page-locked host staging, same lifetime
    // multi-GPU: host-buffer sum over the ranks that share one fit
    ate_allreduce_fn allreduce = nullptr;
    void *allreduce_user = nullptr;
    // multi-GPU: NCCL communicator owned by the ctx (comm.cu); collectives run on `stream`
    int debug_skip = 0;  // ate_round_kernel_bench only: roles of the step kernel to idle
    void *nccl_comm = nullptr;
    int nccl_rank = 0, nccl_nranks = 1;
};

// ---- error handling ------------------------------------------------------------------
int ate_fail(ate_ctx *ctx, int code, const char *fmt, ...);

#define ATE_CUDA(ctx, call)                                                                   \
    do {                                                                                      \
        cudaError_t e__ = (call);                                                             \
        if (e__ != cudaSuccess)                                                               \
            return ate_fail((ctx), -100, "%s:%d: %s -> %s", __FILE__, __LINE__, #call,        \
                            cudaGetErrorString(e__));                                         \
    } while (0)

#define ATE_TRY(expr)                 \
    do {                              \
        int rc__ = (expr);            \
        if (rc__ < 0) return rc__;    \
    } while (0)

#define ATE_LAUNCHED(ctx)                                                                     \
    do {                                                                                      \
        (ctx)->launches++;                                                                    \
        cudaError_t e__ = cudaGetLastError();                                                 \
        if (e__ != cudaSuccess)                                                               \
            return ate_fail((ctx), -101, "%s:%d: kernel launch -> %s", __FILE__, __LINE__,    \
                            cudaGetErrorString(e__));                                         \
    } while (0)

// ---- RAII device buffer (freed when the entry point returns) ------------------------------
template <typename T>
struct DevBuf {
    T *p = nullptr;
    size_t n = 0;
    DevBuf() {}
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    ~DevBuf() { release(); }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        n = 0;
    }
    cudaError_t alloc(size_t count) {
        release();
        n = count;
        cudaError_t e = cudaMalloc((void **)&p, sizeof(T) * (count ? count : 1));
        // ATE_POISON=1 (debugging): fill fresh buffers with 0xFF (NaN doubles, -1 ints) so that a
        // read of memory the library never wrote shows up deterministically
        if (e == cudaSuccess && getenv("ATE_POISON")) {
            e = cudaMemset(p, 0xFF, sizeof(T) * (count ? count : 1));
            cudaDeviceSynchronize();  // the ctx stream does not order against the legacy stream
        }
        return e;
    }
};

int ate_scratch(ate_ctx *ctx, size_t bytes, void **out);
// persistent, grow-only buffer identified by `tag`; released by ate_ctx_destroy
int ate_cached(ate_ctx *ctx, const char *tag, size_t bytes, void **out);
int ate_gram_scratch(ate_ctx *ctx, size_t bytes, void **out);
// persistent, grow-only page-locked HOST buffer (contents are lost when it grows)
int ate_cached_pinned(ate_ctx *ctx, const char *tag, size_t bytes, void **out);

// ---- device helpers ---------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// block-wide sum of NV values per thread; result valid in thread 0 (smem: NV*32 doubles)
template <int NV>
__device__ __forceinline__ void block_sum(double (&v)[NV], double *smem) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int i = 0; i < NV; ++i) v[i] = warp_sum(v[i]);
    __syncthreads();
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < NV; ++i) smem[i * 32 + wid] = v[i];
    }
    __syncthreads();
    if (wid == 0) {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            double t = lane < nw ? smem[i * 32 + lane] : 0.0;
            v[i] = warp_sum(t);
        }
    }
}

// ---- internal entry points used across translation units (device pointers) ---------------
// row-major design layout: n rows of ld doubles.
int ate_i_gram(ate_ctx *ctx, const double *X, long long ld, int p, const int32_t *rows,
               long long row_begin, long long nrows, const double *v, double *G /*p*p*/,
               long long hole_begin = 0, long long hole_len = 0, const int *colidx = nullptr,
               int ncload = 0, int ldg = 0);
// fold the row-chunk partials of a block-lower-triangular Gram into G (both triangles)
int ate_i_gram_fold(ate_ctx *ctx, const double *partials, int nchunks, int nblk, int p, int ldg, double *G);
int ate_i_transpose_to_rowmajor(ate_ctx *ctx, const double *x_colmajor, long long n, int p,
                                double *X, long long ld, int col_offset);
int ate_i_standardize(ate_ctx *ctx, double *d_x, long long n, int p, const int *d_cols, int ncols,
                      double *d_center, double *d_scale);
// d_xbase: column-major base covariates with leading dimension ldx (>= n rows are built from its top)
int ate_i_build_design(ate_ctx *ctx, const double *d_xbase, long long ldx, long long n, int p0, int expand,
                       const double *d_w, const double *d_y, double *d_X, long long ld);
int ate_i_gather_rows(ate_ctx *ctx, const double *X, long long ld, const int32_t *list, long long nt,
                      int ncs, double *Xp, long long ldp);
int ate_i_fill_column(ate_ctx *ctx, double *X, long long n, long long ld, int col, double value);
int ate_i_score(ate_ctx *ctx, long long n, const double *w, const double *y, const double *p,
                const double *m0, const double *m1, double *d_est1, double *d_est2,
                double *tau, double *se, double *d_psi = nullptr, double *n_ok = nullptr);
int ate_i_bootstrap(ate_ctx *ctx, long long n, const double *d_est1, const double *d_est2, int B,
                    int mode, const int32_t *idx, int idx_on_host, uint64_t seed, long long rep_offset,
                    double *d_out, double *boot_host, const double *d_psi, int has_nan_known);
double ate_host_sd(const double *x, long long n);
void ate_host_rng_seed(ate_rng_state *r, uint32_t seed);
void ate_host_sample_replace(ate_rng_state *r, int n, long long k, int kind, int32_t *out);
void ate_host_foldid(ate_rng_state *r, int n, int nfolds, int kind, int32_t *out);
