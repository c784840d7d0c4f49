// gram.cu -- the weighted Gram  G = sum_i v_i x_i x_i^T  in fp64 on the DMMA tensor pipe.
//
// This is the dominant kernel of the path: every IRLS step of the L1-logistic propensity
// fits (ate_functions.R:139-141 -> glmnet's lognet) needs a fresh X^T V X, and the gaussian
// lasso fits (ate_functions.R:101-103) need X^T X per inner fold.  glmnet itself works with
// O(n) column dot products; the engine moves all O(n) work into this one contraction and runs
// coordinate descent on the p x p result.
//
// Shape of the kernel (sm_100a):
//   * the design is row-major, one contiguous record per observation, so a stage of 32
//     observations is staged by 32 one-dimensional TMA bulk copies (cp.async.bulk, SASS
//     UBLKCP) issued by a producer warp -- one per lane -- into a 3..4 deep mbarrier ring.
//     Row lists (training rows of a fold) cost nothing extra: each copy takes its own row id.
//   * fp64 tensor cores on sm_100a are reached through mma.sync.m8n8k4.f64 (DMMA.8x8x4);
//     tcgen05 has no f64 kind.  The k dimension of the MMA is the observation index.
//   * the lower triangle of G is cut into 32x32 blocks; each consumer warp owns one block
//     (16 accumulators of 8x8 = 64 registers) and reads both operands from the same staged
//     rows: A = X[rows, bi-panel]^T, B = v * X[rows, bj-panel].  A padded row pitch
//     (ncols + 4 doubles) makes the 64-bit fragment loads bank-conflict free.
//   * split over observations: CTA (g, c) handles block group g on row chunk c and leaves its
//     partial blocks in scratch; a second small kernel folds the chunks in fixed order, so
//     the result is bit-reproducible run to run.
//
// Algorithmic work per launch: nrows * p * (p + 1) flop (lower triangle incl.
// diagonal), nrows * (p + 1) * 8 bytes.
#include "common.cuh"
#include "common.cuh"

namespace {

constexpr int G_ROWS = 32;    // observations per pipeline stage
constexpr int G_CWARPS = 12;  // consumer (MMA) warps per CTA
constexpr int G_THREADS = (G_CWARPS + 1) * 32;
constexpr int G_PAD = 4;      // doubles of padding per staged row
constexpr int G_MAX_STAGES = 4;
constexpr int G_PANEL = 96;   // column panel of the wide-design path: 3 x 3 blocks per pair = 9 of the 12 MMA warps
constexpr size_t G_SMEM_LIMIT = 227 * 1024;

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!done);
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// block id -> (bi, bj), bi >= bj, lower triangle in row order
__device__ __forceinline__ void block_coords(int q, int &bi, int &bj) {
    int i = (int)((sqrtf(8.0f * (float)q + 1.0f) - 1.0f) * 0.5f);
    while ((i + 1) * (i + 2) / 2 <= q) ++i;
    while (i * (i + 1) / 2 > q) --i;
    bi = i;
    bj = q - i * (i + 1) / 2;
}

__global__ void __launch_bounds__(G_THREADS, 1)
gram_dmma_kernel(const double *__restrict__ X, long long ld, int ncols, const int32_t *__restrict__ rows,
                 long long row_begin, long long nrows, long long hole_begin, long long hole_len,
                 long long rows_per_chunk, const double *__restrict__ v, int nblk, int stages,
                 const int *__restrict__ colidx, double *__restrict__ partials, int wA, long long offB) {
    // Panel-pair mode (wA > 0; designs too wide to stage a whole row): a staged row is
    // [X[r, 0 : wA) | X[r, offB : offB + ncols - wA)] (two bulk copies) and the blocks are the
    // rectangle (second panel) x (first panel), block q = (q / (wA/32), q % (wA/32)).
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int pitch = ncols + G_PAD;                       // doubles
    const size_t stage_doubles = (size_t)G_ROWS * pitch;
    double *sx = reinterpret_cast<double *>(smem_raw);
    double *sv = sx + stage_doubles * stages;              // [stages][G_ROWS]
    uint64_t *full = reinterpret_cast<uint64_t *>(sv + (size_t)stages * G_ROWS);
    uint64_t *empty = full + stages;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long c_begin = (long long)blockIdx.y * rows_per_chunk;
    long long c_rows = nrows - c_begin;
    if (c_rows > rows_per_chunk) c_rows = rows_per_chunk;
    if (c_rows < 0) c_rows = 0;
    const int nit = (int)((c_rows + G_ROWS - 1) / G_ROWS);
    const uint32_t row_bytes = (uint32_t)ncols * 8u;

    for (int e = threadIdx.x; e < stages * G_ROWS * G_PAD; e += blockDim.x)
        sx[(size_t)(e / G_PAD) * pitch + ncols + (e % G_PAD)] = 0.0;
    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(full + s, 1);
            mbar_init(empty + s, G_CWARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp == G_CWARPS) {
        // ---------------- producer warp: one staged observation per lane ----------------
        // The row id and weight of stage it+1 are fetched before waiting for stage it's slot, so
        // the dependent global loads never sit between "slot free" and "copy issued".
        const long long c_end = c_begin + c_rows;
        auto fetch = [&](int it, long long &r, double &vv) -> bool {
            const long long t = c_begin + (long long)it * G_ROWS + lane;
            if (t >= c_end) return false;
            // logical position t skips the hole (the held-out inner fold of a CV fit)
            const long long tp = row_begin + t + (t >= hole_begin ? hole_len : 0);
            r = rows ? (long long)__ldg(rows + tp) : tp;
            vv = v ? __ldg(v + t) : 1.0;
            return true;
        };
        long long cur_r = 0, nxt_r = 0;
        double cur_v = 0.0, nxt_v = 0.0;
        bool cur_ok = nit > 0 ? fetch(0, cur_r, cur_v) : false, nxt_ok = false;
        for (int it = 0; it < nit; ++it) {
            const int s = it % stages;
            const uint32_t ph = (uint32_t)(it / stages) & 1u;
            nxt_ok = (it + 1 < nit) ? fetch(it + 1, nxt_r, nxt_v) : false;
            mbar_wait(empty + s, ph ^ 1u);
            double *dst = sx + stage_doubles * s + (size_t)lane * pitch;
            sv[s * G_ROWS + lane] = cur_ok ? cur_v : 0.0;
            if (!cur_ok)
                for (int c = 0; c < ncols; ++c) dst[c] = 0.0;
            const unsigned vmask = __ballot_sync(0xffffffffu, cur_ok);
            __syncwarp();
            if (lane == 0) mbar_arrive_expect_tx(full + s, row_bytes * (uint32_t)__popc(vmask));
            __syncwarp();
            if (cur_ok) {
                const double *src = X + (size_t)cur_r * (size_t)ld;
                if (wA > 0) {
                    bulk_g2s(dst, src, (uint32_t)wA * 8u, full + s);
                    bulk_g2s(dst + wA, src + offB, row_bytes - (uint32_t)wA * 8u, full + s);
                } else {
                    bulk_g2s(dst, src, row_bytes, full + s);
                }
            }
            cur_ok = nxt_ok;
            cur_r = nxt_r;
            cur_v = nxt_v;
        }
    } else {
        // ---------------- consumer warps: one 32x32 block of the lower triangle each -----
        const int q = blockIdx.x * G_CWARPS + warp;
        const bool active = q < nblk;
        int bi = 0, bj = 0;
        if (active) {
            if (wA > 0) {
                const int nbA = wA >> 5;
                bi = nbA + q / nbA;  // staged columns of the second panel start at wA
                bj = q % nbA;
            } else {
                block_coords(q, bi, bj);
            }
        }
        const bool diag = (bi == bj);
        double acc[4][4][2];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;
        const int g = lane >> 2, k4 = lane & 3;
        // operand columns: the block's own 32 columns, or -- for a column-restricted Gram -- the
        // design columns the list names (a padded list points at the zeroed pad slot `ncols`)
        int offA[4], offB[4];
#pragma unroll
        for (int m = 0; m < 4; ++m) {
            const int ia = bi * 32 + m * 8 + g, ib = bj * 32 + m * 8 + g;
            offA[m] = (colidx && active) ? __ldg(colidx + ia) : ia;
            offB[m] = (colidx && active) ? __ldg(colidx + ib) : ib;
        }
        for (int it = 0; it < nit; ++it) {
            const int s = it % stages;
            const uint32_t ph = (uint32_t)(it / stages) & 1u;
            mbar_wait(full + s, ph);
            if (active) {
                const double *base = sx + stage_doubles * s;
                const double *vs = sv + s * G_ROWS;
#pragma unroll
                for (int kk = 0; kk < G_ROWS / 4; ++kk) {
                    const int r = kk * 4 + k4;
                    const double *rowp = base + (size_t)r * pitch;
                    const double vv = vs[r];
                    double a[4], b[4];
#pragma unroll
                    for (int m = 0; m < 4; ++m) a[m] = rowp[offA[m]];
#pragma unroll
                    for (int m = 0; m < 4; ++m) b[m] = rowp[offB[m]] * vv;
#pragma unroll
                    for (int m = 0; m < 4; ++m)
#pragma unroll
                        for (int nn = 0; nn < 4; ++nn)
                            if (!diag || nn <= m)  // a diagonal block only needs its lower 8x8 tiles
                                dmma884(acc[m][nn][0], acc[m][nn][1], a[m], b[nn]);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty + s);
        }
        if (active) {
            double *out = partials + ((size_t)blockIdx.y * nblk + q) * 1024;
#pragma unroll
            for (int m = 0; m < 4; ++m)
#pragma unroll
                for (int nn = 0; nn < 4; ++nn) {
                    const int r = m * 8 + g, c = nn * 8 + 2 * k4;
                    *reinterpret_cast<double2 *>(out + r * 32 + c) = make_double2(acc[m][nn][0], acc[m][nn][1]);
                }
        }
    }
}

// fold the row chunks in order and scatter into the full symmetric p x p matrix
__global__ void __launch_bounds__(256)
gram_fold_kernel(const double *__restrict__ partials, int nchunks, int nblk, int p, int ldg,
                 double *__restrict__ G) {
    // grid = (blocks of the lower triangle, 4): one element per thread, many chunk loads in flight;
    // the additions stay in chunk order (bit-reproducible)
    const int q = blockIdx.x;
    int bi, bj;
    block_coords(q, bi, bj);
    const int e = blockIdx.y * 256 + threadIdx.x;
    const double *src = partials + (size_t)q * 1024 + e;
    const size_t cstride = (size_t)nblk * 1024;
    double s = 0.0;
    int c = 0;
    for (; c + 16 <= nchunks; c += 16) {
        double t[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) t[k] = __ldg(src + (size_t)(c + k) * cstride);
#pragma unroll
        for (int k = 0; k < 16; ++k) s += t[k];
    }
    for (; c < nchunks; ++c) s += __ldg(src + (size_t)c * cstride);
    const int r = bi * 32 + (e >> 5), col = bj * 32 + (e & 31);
    if (r < p && col < p && col <= r) {
        G[(size_t)r * ldg + col] = s;
        G[(size_t)col * ldg + r] = s;
    }
}

// fold for the panel-pair mode: block q is (q / nbA, q % nbA) of the rectangle G[rowoff.., coloff..]
__global__ void __launch_bounds__(256)
gram_fold_rect_kernel(const double *__restrict__ partials, int nchunks, int nblk, int nbA, int rowoff, int coloff, int p,
                      int ldg, double *__restrict__ G) {
    const int q = blockIdx.x;
    const int e = blockIdx.y * 256 + threadIdx.x;
    const double *src = partials + (size_t)q * 1024 + e;
    const size_t cstride = (size_t)nblk * 1024;
    double s = 0.0;
    for (int c = 0; c < nchunks; ++c) s += __ldg(src + (size_t)c * cstride);
    const int r = rowoff + (q / nbA) * 32 + (e >> 5), col = coloff + (q % nbA) * 32 + (e & 31);
    if (r < p && col < p) {
        G[(size_t)r * ldg + col] = s;
        G[(size_t)col * ldg + r] = s;
    }
}

// the opt-in to more than 48 KB of dynamic shared memory is per device: remember it per device, not
// per process (one process may hold contexts on several GPUs)
int gram_allow_big_smem(ate_ctx *ctx) {
    static bool done[64] = {false};
    const int d = ctx->device;
    if (d >= 0 && d < 64 && done[d]) return 0;
    ATE_CUDA(ctx, cudaFuncSetAttribute(gram_dmma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G_SMEM_LIMIT));
    if (d >= 0 && d < 64) done[d] = true;
    return 0;
}

}  // namespace

int ate_i_gram_fold(ate_ctx *ctx, const double *partials, int nchunks, int nblk, int p, int ldg, double *G) {
    gram_fold_kernel<<<dim3(nblk, 4), 256, 0, ctx->stream>>>(partials, nchunks, nblk, p, ldg, G);
    ATE_LAUNCHED(ctx);
    return 0;
}

// X: row-major device design (ld doubles per row; columns [0, round_up(p,32)) must be readable
// and zero beyond the real columns).  rows: device row list or NULL.  v: device weights indexed
// by position in the list (or NULL).  G: device p*p output.  The nrows logical positions map to
// list entries row_begin + t, skipping hole_len entries from logical position hole_begin on.
// Column-restricted form: colidx (device, padded to a multiple of 32 with the value `ncload`) names
// the p design columns that enter the Gram; `ncload` columns of every row are staged; G is written
// compactly with leading dimension ldg.  colidx == NULL: columns 0..p-1, ncload = round_up(p, 32).
int ate_i_gram(ate_ctx *ctx, const double *X, long long ld, int p, const int32_t *rows,
               long long row_begin, long long nrows, const double *v, double *G, long long hole_begin,
               long long hole_len, const int *colidx, int ncload, int ldg) {
    if (hole_len <= 0) {
        hole_begin = nrows;
        hole_len = 0;
    }
    if (p <= 0 || nrows < 0) return ate_fail(ctx, -1, "gram: bad shape");
    const int pp = (p + 31) / 32 * 32;
    const int ncols = colidx ? ncload : pp;
    if (!colidx) ldg = p;
    if (colidx && (ncload % 2 != 0 || ncload <= 0)) return ate_fail(ctx, -1, "gram: bad ncload");
    if (ncols > ld) return ate_fail(ctx, -1, "gram: ld (%lld) < padded column count (%d)", ld, ncols);
    if ((ld & 1) != 0) return ate_fail(ctx, -1, "gram: ld must be even (16-byte aligned rows)");
    const int nb = pp / 32, nblk = nb * (nb + 1) / 2;
    const size_t stage_bytes = sizeof(double) * (size_t)G_ROWS * (ncols + G_PAD);
    int stages = (int)((G_SMEM_LIMIT - 1024) / (stage_bytes + G_ROWS * sizeof(double)));
    if (stages > G_MAX_STAGES) stages = G_MAX_STAGES;
    if (stages < 2) {
        // A row of this design does not fit the staging ring: cut the columns into panels of
        // G_PANEL and run one launch per panel (triangular) and per panel pair (rectangular).  Every
        // 32x32 block is still computed exactly once; rows are re-read once per panel pair, which at
        // these widths (>= 2 * 96^2 flop per 1.5 KB staged) stays far below the DMMA time.
        if (colidx) return ate_fail(ctx, -1, "gram: a %d-column list does not fit the full-row kernel", ncload);
        if (pp > ld) return ate_fail(ctx, -1, "gram: ld (%lld) < padded column count (%d)", ld, pp);
        ATE_TRY(gram_allow_big_smem(ctx));
        auto launch = [&](const double *Xb, int nc, int nblk_l, int wA, long long offB, double **scr_out, int *nchunks_out,
                          double flops) -> int {
            const size_t sb = sizeof(double) * (size_t)G_ROWS * (nc + G_PAD);
            int st = (int)((G_SMEM_LIMIT - 1024) / (sb + G_ROWS * sizeof(double)));
            if (st > G_MAX_STAGES) st = G_MAX_STAGES;
            const size_t sm = st * (sb + G_ROWS * sizeof(double)) + 2 * st * sizeof(uint64_t);
            const int ngr = (nblk_l + G_CWARPS - 1) / G_CWARPS;
            int nch = ctx->num_sms / ngr;
            if (nch < 1) nch = 1;
            long long mc = (nrows + 4 * G_ROWS - 1) / (4 * G_ROWS);
            if (mc < 1) mc = 1;
            if (nch > mc) nch = (int)mc;
            long long rpc_l = (nrows + nch - 1) / nch;
            rpc_l = (rpc_l + G_ROWS - 1) / G_ROWS * G_ROWS;
            if (rpc_l < G_ROWS) rpc_l = G_ROWS;
            nch = (int)((nrows + rpc_l - 1) / rpc_l);
            if (nch < 1) nch = 1;
            void *scr_l = nullptr;
            ATE_TRY(ate_gram_scratch(ctx, sizeof(double) * 1024 * (size_t)nblk_l * nch, &scr_l));
            cudaEvent_t e0 = nullptr, e1 = nullptr;
            if (ctx->profile & 2) {
                ATE_CUDA(ctx, cudaEventCreate(&e0));
                ATE_CUDA(ctx, cudaEventCreate(&e1));
                ATE_CUDA(ctx, cudaEventRecord(e0, ctx->stream));
            }
            gram_dmma_kernel<<<dim3(ngr, nch), G_THREADS, sm, ctx->stream>>>(Xb, ld, nc, rows, row_begin, nrows, hole_begin,
                                                                            hole_len, rpc_l, v, nblk_l, st, nullptr,
                                                                            (double *)scr_l, wA, offB);
            ATE_LAUNCHED(ctx);
            if (ctx->profile & 2) {
                ATE_CUDA(ctx, cudaEventRecord(e1, ctx->stream));
                ctx->prof_events.emplace_back(e0, e1);
                ctx->prof_shape.emplace_back(-nc, nrows);
                ctx->prof_launches++;
                ctx->prof_flops += flops;
                ctx->prof_bytes += (double)nrows * (nc + 1.0) * 8.0;
            }
            *scr_out = (double *)scr_l;
            *nchunks_out = nch;
            return 0;
        };
        const int npan = (pp + G_PANEL - 1) / G_PANEL;
        for (int I = 0; I < npan; ++I) {
            const int cI = I * G_PANEL, wI = (pp - cI < G_PANEL) ? pp - cI : G_PANEL, nbI = wI / 32;
            double *scr_l = nullptr;
            int nch = 0;
            const int pI = (p - cI < wI) ? p - cI : wI;  // real columns of this panel
            ATE_TRY(launch(X + cI, wI, nbI * (nbI + 1) / 2, 0, 0, &scr_l, &nch, (double)nrows * pI * (pI + 1.0)));
            gram_fold_kernel<<<dim3(nbI * (nbI + 1) / 2, 4), 256, 0, ctx->stream>>>(scr_l, nch, nbI * (nbI + 1) / 2, pI, p,
                                                                                   G + (size_t)cI * p + cI);
            ATE_LAUNCHED(ctx);
            for (int J = 0; J < I; ++J) {
                const int cJ = J * G_PANEL, nbJ = G_PANEL / 32;
                ATE_TRY(launch(X + cJ, G_PANEL + wI, nbI * nbJ, G_PANEL, (long long)(cI - cJ), &scr_l, &nch,
                               2.0 * (double)nrows * pI * G_PANEL));
                gram_fold_rect_kernel<<<dim3(nbI * nbJ, 4), 256, 0, ctx->stream>>>(scr_l, nch, nbI * nbJ, nbJ, cI, cJ, p, p, G);
                ATE_LAUNCHED(ctx);
            }
        }
        return 0;
    }
    const size_t smem = stages * (stage_bytes + G_ROWS * sizeof(double)) + 2 * stages * sizeof(uint64_t);
    const int ngroups = (nblk + G_CWARPS - 1) / G_CWARPS;
    int nchunks = ctx->num_sms / ngroups;
    if (nchunks < 1) nchunks = 1;
    long long max_chunks = (nrows + 4 * G_ROWS - 1) / (4 * G_ROWS);
    if (max_chunks < 1) max_chunks = 1;
    if (nchunks > max_chunks) nchunks = (int)max_chunks;
    long long rpc = (nrows + nchunks - 1) / nchunks;
    rpc = (rpc + G_ROWS - 1) / G_ROWS * G_ROWS;
    if (rpc < G_ROWS) rpc = G_ROWS;
    nchunks = (int)((nrows + rpc - 1) / rpc);
    if (nchunks < 1) nchunks = 1;
    void *scr = nullptr;
    ATE_TRY(ate_gram_scratch(ctx, sizeof(double) * 1024 * (size_t)nblk * nchunks, &scr));
    ATE_TRY(gram_allow_big_smem(ctx));
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ctx->profile & 2) {
        ATE_CUDA(ctx, cudaEventCreate(&e0));
        ATE_CUDA(ctx, cudaEventCreate(&e1));
        ATE_CUDA(ctx, cudaEventRecord(e0, ctx->stream));
    }
    gram_dmma_kernel<<<dim3(ngroups, nchunks), G_THREADS, smem, ctx->stream>>>(
        X, ld, ncols, rows, row_begin, nrows, hole_begin, hole_len, rpc, v, nblk, stages, colidx, (double *)scr, 0, 0);
    ATE_LAUNCHED(ctx);
    if (ctx->profile & 2) {
        ATE_CUDA(ctx, cudaEventRecord(e1, ctx->stream));
        ctx->prof_events.emplace_back(e0, e1);
        ctx->prof_shape.emplace_back(-p, nrows);
        ctx->prof_launches++;
        ctx->prof_flops += (double)nrows * p * (p + 1.0);
        ctx->prof_bytes += (double)nrows * (ncols + 1.0) * 8.0;  // whole staged rows + the weight
    }
    gram_fold_kernel<<<dim3(nblk, 4), 256, 0, ctx->stream>>>((const double *)scr, nchunks, nblk, p, ldg, G);
    ATE_LAUNCHED(ctx);
    return 0;
}

extern "C" int ate_weighted_gram(ate_ctx *ctx, long long n, int p, const double *x, long long ld,
                                 const double *v, const int32_t *rows, long long nrows, double *G,
                                 int on_device) {
    if (!ctx) return -1;
    if (!x || !G || n <= 0 || p <= 0) return ate_fail(ctx, -1, "ate_weighted_gram: bad arguments");
    ATE_CUDA(ctx, cudaSetDevice(ctx->device));
    const long long m = rows ? nrows : n;
    if (on_device) {
        ATE_TRY(ate_i_gram(ctx, x, ld, p, rows, 0, m, v, G));
        ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        return 0;
    }
    const long long ldd = (p + 31) / 32 * 32;
    DevBuf<double> dx, dX, dv, dG;
    DevBuf<int32_t> dr;
    ATE_CUDA(ctx, dx.alloc((size_t)n * p));
    ATE_CUDA(ctx, dX.alloc((size_t)n * ldd));
    ATE_CUDA(ctx, dG.alloc((size_t)p * p));
    ATE_CUDA(ctx, cudaMemcpyAsync(dx.p, x, sizeof(double) * (size_t)n * p, cudaMemcpyHostToDevice, ctx->stream));
    ATE_CUDA(ctx, cudaMemsetAsync(dX.p, 0, sizeof(double) * (size_t)n * ldd, ctx->stream));
    ATE_TRY(ate_i_transpose_to_rowmajor(ctx, dx.p, n, p, dX.p, ldd, 0));
    if (v) {
        ATE_CUDA(ctx, dv.alloc((size_t)m));
        ATE_CUDA(ctx, cudaMemcpyAsync(dv.p, v, sizeof(double) * (size_t)m, cudaMemcpyHostToDevice, ctx->stream));
    }
    if (rows) {
        for (long long i = 0; i < nrows; ++i)
            if (rows[i] < 0 || rows[i] >= n) return ate_fail(ctx, -1, "ate_weighted_gram: row %d out of range", rows[i]);
        ATE_CUDA(ctx, dr.alloc((size_t)nrows));
        ATE_CUDA(ctx, cudaMemcpyAsync(dr.p, rows, sizeof(int32_t) * (size_t)nrows, cudaMemcpyHostToDevice, ctx->stream));
    }
    ATE_TRY(ate_i_gram(ctx, dX.p, ldd, p, rows ? dr.p : nullptr, 0, m, v ? dv.p : nullptr, dG.p));
    ATE_CUDA(ctx, cudaMemcpyAsync(G, dG.p, sizeof(double) * (size_t)p * p, cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}
