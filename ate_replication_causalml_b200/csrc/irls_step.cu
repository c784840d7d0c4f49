// irls_step.cu -- one IRLS step of the L1-logistic fit (glmnet's lognet, ate_functions.R:139-141)
// as ONE pass over the observations: the design rows are read from HBM once and serve both
//   (1) the working-response refresh   eta = a0 + x beta, q = plogis(eta), v = w q (1-q),
//       r = w (y - q), the deviance, X^T r and X^T v, and
//   (2) the weighted Gram  H = X_G^T V X_G  over a column list G (the strong set plus the
//       columns predicted to join it) on the fp64 DMMA pipe.
// glmnet spends these O(n) pieces inside its coordinate-descent loop; here they are one
// HBM-bound kernel (algorithmic bytes: nrows * (ncols + 1) * 8) and the coordinate descent runs
// on the p x p result (path_binom.cu).
//
// CTA = 8 MMA warps + 2 helper warps + 2 producer warps + 4 evaluation warps, one CTA per SM.
// The working-response refresh is a long dependent chain (dot -> shuffle reduce -> exp -> divide
// -> log), so it is latency-bound per warp; slots are therefore small (16 rows, up to 6 in flight)
// and handed to the evaluation warps round-robin: each warp works on every 4th slot and has four
// slot periods to finish it.
//   producer : stages 16 rows per slot with 16 one-dimensional TMA bulk copies (one per lane)
//              into an mbarrier ring, plus y and a validity flag per row.
//   eval     : lanes hold a column slice of beta; the 16 partial dot products of a slot are
//              reduce-scattered across the warp (5 exchange steps), so lane l ends up with row
//              l/2 and the exp/log work of 16 rows runs in parallel lanes.  v and r go to shared
//              memory (second mbarrier) for the MMA and helper warps.
//   helpers  : (lead CTA of a row chunk) accumulate X^T r, X^T v on the FP64 FMA pipe from the
//              same staged rows -- off the eval warps' critical path and off the tensor pipe.
//   MMA      : as gram.cu -- one 32x32 block of the lower triangle per warp, operands gathered from
//              the staged rows through the column list.
// Row chunks are folded in fixed order afterwards: bit-reproducible.
#include <math.h>
#include <stdlib.h>

#include "irls_step.cuh"
#include "path.cuh"
#include "tma_dmma.cuh"

namespace {

using namespace atk;

constexpr int F_ROWS = 16;    // rows per slot: small slots, many of them in flight
constexpr int F_CWARPS = 8;   // MMA warps;  8 + 2 + 2 + 4 = 16 warps: 128 registers per thread
constexpr int F_HWARPS = 2;   // helper warps (X^T r, X^T v); slot `it` belongs to helper it % 2
constexpr int F_EWARPS = 4;   // evaluation warps; slot `it` belongs to eval warp it % 4
constexpr int F_PWARPS = 2;   // producer warps; slot `it` belongs to producer it % 2.  One warp issues a
                              // 2 KB bulk copy every ~140 cycles (the copies of a warp's lanes are
                              // serialised on the uniform datapath): 4.1 TB/s chip-wide, measured --
                              // two warps lift the issue rate above what HBM delivers
constexpr int F_THREADS = (F_CWARPS + F_HWARPS + F_PWARPS + F_EWARPS) * 32;
constexpr int F_PRODUCER = F_CWARPS + F_HWARPS;  // warp index of the first producer
constexpr int F_PAD = 4;
constexpr int F_MAX_STAGES = 8;
constexpr size_t F_SMEM_LIMIT = 227 * 1024;
static_assert(F_ROWS == 16 && F_HWARPS <= F_EWARPS, "the eval reduce-scatter is written for 16-row slots");

struct StepArgs {
    const double *X;
    long long ld;
    int ncols;
    const int32_t *rows;
    long long nrows, hole_begin, hole_len, rows_per_chunk;
    int ni, iy;
    const double *coef;
    double w;
    const int *colidx;
    int nblk, stages;
    double *gram_partials, *eval_partials, *v_out;
};

template <int ITER, bool INIT>
__global__ void __launch_bounds__(F_THREADS, 1) irls_step_kernel(const StepArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int ncols = a.ncols, pitch = ncols + F_PAD, stages = a.stages;
    const size_t stage_doubles = (size_t)F_ROWS * pitch;
    double *sx = reinterpret_cast<double *>(smem_raw);
    double *sv = sx + stage_doubles * stages;  // [stages][32]: validity from the producer, then v
    double *sy = sv + (size_t)stages * F_ROWS;  // [stages][F_ROWS]: r from the eval warp
    uint64_t *full = reinterpret_cast<uint64_t *>(sy + (size_t)stages * F_ROWS);
    uint64_t *vfull = full + stages, *empty = vfull + stages;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long c_begin = (long long)blockIdx.y * a.rows_per_chunk;
    long long c_rows = a.nrows - c_begin;
    if (c_rows > a.rows_per_chunk) c_rows = a.rows_per_chunk;
    if (c_rows < 0) c_rows = 0;
    const long long c_end = c_begin + c_rows;
    const int nit = (int)((c_rows + F_ROWS - 1) / F_ROWS);
    const uint32_t row_bytes = (uint32_t)ncols * 8u;

    for (int e = threadIdx.x; e < stages * F_ROWS * F_PAD; e += blockDim.x)
        sx[(size_t)(e / F_PAD) * pitch + ncols + (e % F_PAD)] = 0.0;  // the list's "zero column"
    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(full + s, 1);
            mbar_init(vfull + s, 1);
            mbar_init(empty + s, F_CWARPS + 2);  // MMA warps + the slot's eval warp + its helper warp
        }
        mbar_fence_init();
    }
    __syncthreads();

    if (warp >= F_PRODUCER && warp < F_PRODUCER + F_PWARPS) {
        // ------------------------------ producer ------------------------------
        // Row ids are fetched F_PF slots ahead (a register ring), so the global-load latency of the
        // list never sits between "slot free" and "copy issued".  y is not loaded here: the staged
        // width covers the response column and the eval warps read it from shared memory.
        const int pw = warp - F_PRODUCER;
        if (a.rows == nullptr && a.ld == (long long)pitch) {
            // Contiguous rows whose pitch in HBM equals the slot pitch: the whole slot is ONE bulk
            // copy (two when it straddles the held-out hole).  The TMA unit serves small copies at
            // ~140 cycles each, so per-row copies cap at ~4.2 TB/s chip-wide; slot-sized ones do not.
            for (int it = pw; it < nit; it += F_PWARPS) {
                const int s = it % stages;
                const uint32_t ph = (uint32_t)(it / stages) & 1u;
                const long long t0 = c_begin + (long long)it * F_ROWS;
                const int cnt = (int)((c_end - t0) < F_ROWS ? (c_end - t0) : F_ROWS);
                mbar_wait(empty + s, ph ^ 1u);
                double *dst0 = sx + stage_doubles * s;
                if (lane < F_ROWS) {
                    sv[s * F_ROWS + lane] = lane < cnt ? 1.0 : 0.0;
                    if (lane >= cnt)
                        for (int c = 0; c < ncols; ++c) dst0[(size_t)lane * pitch + c] = 0.0;
                }
                __syncwarp();
                if (lane == 0) {
                    // rows t0 .. t0+cnt-1; positions >= hole_begin are shifted by hole_len
                    long long n1 = a.hole_begin - t0;
                    if (n1 < 0) n1 = 0;
                    if (n1 > cnt) n1 = cnt;
                    const long long n2 = cnt - n1;
                    const uint32_t rowb = (uint32_t)pitch * 8u;
                    mbar_arrive_expect_tx(full + s, rowb * (uint32_t)cnt);
                    if (n1 > 0) bulk_g2s(dst0, a.X + (size_t)t0 * (size_t)a.ld, rowb * (uint32_t)n1, full + s);
                    if (n2 > 0)
                        bulk_g2s(dst0 + (size_t)n1 * pitch, a.X + (size_t)(t0 + n1 + a.hole_len) * (size_t)a.ld,
                                 rowb * (uint32_t)n2, full + s);
                }
                __syncwarp();
            }
        } else {
        constexpr int F_PF = 4;
        auto fetch = [&](int it, long long &r) -> bool {
            const long long t = c_begin + (long long)it * F_ROWS + lane;
            if (lane >= F_ROWS || it >= nit || t >= c_end) return false;
            const long long tp = t + (t >= a.hole_begin ? a.hole_len : 0);
            r = a.rows ? (long long)__ldg(a.rows + tp) : tp;
            return true;
        };
        long long ring_r[F_PF];
        bool ring_ok[F_PF];
#pragma unroll
        for (int d = 0; d < F_PF; ++d) {
            ring_r[d] = 0;
            ring_ok[d] = fetch(pw + d * F_PWARPS, ring_r[d]);
        }
        for (int it0 = pw; it0 < nit; it0 += F_PF * F_PWARPS) {
#pragma unroll
            for (int d = 0; d < F_PF; ++d) {
                const int it = it0 + d * F_PWARPS;
                if (it >= nit) break;
                const long long cur_r = ring_r[d];
                const bool cur_ok = ring_ok[d];
                ring_ok[d] = fetch(it + F_PF * F_PWARPS, ring_r[d]);
                const int s = it % stages;
                const uint32_t ph = (uint32_t)(it / stages) & 1u;
                mbar_wait(empty + s, ph ^ 1u);
                double *dst = sx + stage_doubles * s + (size_t)lane * pitch;
                if (lane < F_ROWS) {
                    sv[s * F_ROWS + lane] = cur_ok ? 1.0 : 0.0;
                    if (!cur_ok)
                        for (int c = 0; c < ncols; ++c) dst[c] = 0.0;
                }
                const unsigned vmask = __ballot_sync(0xffffffffu, cur_ok);
                __syncwarp();
                if (lane == 0) mbar_arrive_expect_tx(full + s, row_bytes * (uint32_t)__popc(vmask));
                __syncwarp();
                if (cur_ok) bulk_g2s(dst, a.X + (size_t)cur_r * (size_t)a.ld, row_bytes, full + s);
            }
        }
        }
    } else if (warp >= F_PRODUCER + F_PWARPS) {
        // ------------------------------ evaluation ------------------------------
        const int e = warp - (F_PRODUCER + F_PWARPS);
        const bool lead = (blockIdx.x == 0);
        const int ni = a.ni;
        double beta[2 * ITER];
#pragma unroll
        for (int i = 0; i < ITER; ++i) {
            const int c = 2 * lane + 64 * i;
            beta[2 * i] = c < ni ? a.coef[c] : 0.0;
            beta[2 * i + 1] = c + 1 < ni ? a.coef[c + 1] : 0.0;
        }
        const double a0 = a.coef[ni], w = a.w;
        const double pmin = GLM_PMIN, fmax_ = log(1.0 / pmin - 1.0), fmin_ = -fmax_;
        double s_xmz = 0.0, s_sr = 0.0, s_dev = 0.0;
        const int rr = lane >> 1;  // the row of the slot this lane finishes
        for (int it = e; it < nit; it += F_EWARPS) {
            const int s = it % stages;
            const uint32_t ph = (uint32_t)(it / stages) & 1u;
            mbar_wait(full + s, ph);
            const double *base = sx + stage_doubles * s;
            double part[F_ROWS];
#pragma unroll
            for (int q = 0; q < F_ROWS; ++q) {
                const double *rowp = base + (size_t)q * pitch;
                double dot = 0.0;
#pragma unroll
                for (int i = 0; i < ITER; ++i) {
                    const int c = 2 * lane + 64 * i;
                    if (c < ncols) {
                        const double2 t2 = *reinterpret_cast<const double2 *>(rowp + c);
                        dot += t2.x * beta[2 * i] + t2.y * beta[2 * i + 1];
                    }
                }
                part[q] = dot;
            }
            // reduce-scatter the 16 partial dots over the 32 lanes: lane l ends with row l >> 1
#pragma unroll
            for (int half = 8, off = 16; half >= 1; half >>= 1, off >>= 1) {
                const bool up = (lane & off) != 0;
#pragma unroll
                for (int k = 0; k < half; ++k) {
                    const double send = up ? part[k] : part[k + half], keep = up ? part[k + half] : part[k];
                    part[k] = keep + __shfl_xor_sync(0xffffffffu, send, off);
                }
            }
            part[0] += __shfl_xor_sync(0xffffffffu, part[0], 1);
            const double valid = sv[s * F_ROWS + rr], yi = base[(size_t)rr * pitch + a.iy];
            const double fi = a0 + part[0];
            double q;
            if (fi < fmin_)
                q = 0.0;
            else if (fi > fmax_)
                q = 1.0;
            else
                q = 1.0 / (1.0 + exp(-fi));
            const double vi = w * q * (1.0 - q) * valid, ri = w * (yi - q) * valid;
            __syncwarp();  // every lane has read the validity flag before it is overwritten
            if ((lane & 1) == 0) {
                sv[s * F_ROWS + rr] = vi;
                sy[s * F_ROWS + rr] = ri;
            }
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(vfull + s);  // MMA and helper warps may start on this slot
                mbar_arrive(empty + s);  // ... and this warp is done with its staged rows
            }
            if (lead && (lane & 1) == 0 && valid != 0.0) {
                a.v_out[c_begin + (long long)it * F_ROWS + rr] = vi;
                const double qc = fmin(fmax(q, pmin), 1.0 - pmin);
                s_xmz += vi;
                s_sr += ri;
                // y is 0/1 on this path, so one of the two log terms vanishes
                const double ll = (yi == 1.0) ? log(qc) : (yi == 0.0) ? log(1.0 - qc)
                                                        : yi * log(qc) + (1.0 - yi) * log(1.0 - qc);
                s_dev -= w * ll;
            }
        }
        if (lead) {
            const int plen = 3 * ni + 3;
            double *ep = a.eval_partials + ((size_t)blockIdx.y * F_EWARPS + e) * plen;
            if (e >= F_HWARPS)  // no helper writes this slot's vector part
                for (int c = lane; c < 3 * ni; c += 32) ep[c] = 0.0;
            s_xmz = warp_sum(s_xmz);
            s_sr = warp_sum(s_sr);
            s_dev = warp_sum(s_dev);
            if (lane == 0) {
                ep[3 * ni] = s_xmz;
                ep[3 * ni + 1] = s_sr;
                ep[3 * ni + 2] = s_dev;
            }
        }
    } else if (warp >= F_CWARPS) {
        // ------------------------------ helpers: X^T r, X^T v (and sum v x^2 on the first step) -----
        const int h = warp - F_CWARPS;
        const bool lead = (blockIdx.x == 0);
        const int ni = a.ni;
        double gr[2 * ITER], hx[2 * ITER], hxx[INIT ? 2 * ITER : 1];
#pragma unroll
        for (int i = 0; i < ITER; ++i) {
            gr[2 * i] = gr[2 * i + 1] = hx[2 * i] = hx[2 * i + 1] = 0.0;
            if (INIT) hxx[2 * i] = hxx[2 * i + 1] = 0.0;
        }
        for (int it = h; it < nit; it += F_HWARPS) {
            const int s = it % stages;
            const uint32_t ph = (uint32_t)(it / stages) & 1u;
            mbar_wait(full + s, ph);
            if (lead) {
                mbar_wait(vfull + s, ph);
                const double *base = sx + stage_doubles * s;
#pragma unroll 4
                for (int q2 = 0; q2 < F_ROWS; ++q2) {
                    const double vj = sv[s * F_ROWS + q2];
                    const double rj = sy[s * F_ROWS + q2];
                    const double *rowp = base + (size_t)q2 * pitch;
#pragma unroll
                    for (int i = 0; i < ITER; ++i) {
                        const int c = 2 * lane + 64 * i;
                        if (c < ncols) {
                            const double2 t2 = *reinterpret_cast<const double2 *>(rowp + c);
                            gr[2 * i] = fma(rj, t2.x, gr[2 * i]);
                            gr[2 * i + 1] = fma(rj, t2.y, gr[2 * i + 1]);
                            hx[2 * i] = fma(vj, t2.x, hx[2 * i]);
                            hx[2 * i + 1] = fma(vj, t2.y, hx[2 * i + 1]);
                            if (INIT) {
                                hxx[2 * i] = fma(vj * t2.x, t2.x, hxx[2 * i]);
                                hxx[2 * i + 1] = fma(vj * t2.y, t2.y, hxx[2 * i + 1]);
                            }
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty + s);
        }
        if (lead) {
            const int plen = 3 * ni + 3;
            double *ep = a.eval_partials + ((size_t)blockIdx.y * F_EWARPS + h) * plen;
#pragma unroll
            for (int i = 0; i < ITER; ++i)
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    const int c = 2 * lane + 64 * i + hh;
                    if (c < ni) {
                        ep[c] = gr[2 * i + hh];
                        ep[ni + c] = hx[2 * i + hh];
                        ep[2 * ni + c] = INIT ? hxx[INIT ? 2 * i + hh : 0] : 0.0;
                    }
                }
        }
    } else {
        // ------------------------------ MMA ------------------------------
        const int q = blockIdx.x * F_CWARPS + warp;
        const bool active = q < a.nblk;
        int bi = 0, bj = 0;
        if (active) tri_block(q, bi, bj);
        const bool diag = (bi == bj);
        double acc[4][4][2];
#pragma unroll
        for (int m = 0; m < 4; ++m)
#pragma unroll
            for (int n = 0; n < 4; ++n) acc[m][n][0] = acc[m][n][1] = 0.0;
        const int g = lane >> 2, k4 = lane & 3;
        int offA[4], offB[4];
#pragma unroll
        for (int m = 0; m < 4; ++m) {
            const int ia = bi * 32 + m * 8 + g, ib = bj * 32 + m * 8 + g;
            offA[m] = active ? __ldg(a.colidx + ia) : 0;
            offB[m] = active ? __ldg(a.colidx + ib) : 0;
        }
        for (int it = 0; it < nit; ++it) {
            const int s = it % stages;
            const uint32_t ph = (uint32_t)(it / stages) & 1u;
            mbar_wait(full + s, ph);
            mbar_wait(vfull + s, ph);
            if (active) {
                const double *base = sx + stage_doubles * s;
                const double *vs = sv + s * F_ROWS;
#pragma unroll
                for (int kk = 0; kk < F_ROWS / 4; ++kk) {
                    const int r = kk * 4 + k4;
                    const double *rowp = base + (size_t)r * pitch;
                    const double vv = vs[r];
                    double fa[4], fb[4];
#pragma unroll
                    for (int m = 0; m < 4; ++m) fa[m] = rowp[offA[m]];
#pragma unroll
                    for (int m = 0; m < 4; ++m) fb[m] = rowp[offB[m]] * vv;
#pragma unroll
                    for (int m = 0; m < 4; ++m)
#pragma unroll
                        for (int n = 0; n < 4; ++n)
                            if (!diag || n <= m)  // a diagonal block only needs its lower 8x8 tiles
                                dmma884(acc[m][n][0], acc[m][n][1], fa[m], fb[n]);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty + s);
        }
        if (active) {
            double *out = a.gram_partials + ((size_t)blockIdx.y * a.nblk + q) * 1024;
#pragma unroll
            for (int m = 0; m < 4; ++m)
#pragma unroll
                for (int n = 0; n < 4; ++n) {
                    const int r = m * 8 + g, c = n * 8 + 2 * k4;
                    *reinterpret_cast<double2 *>(out + r * 32 + c) = make_double2(acc[m][n][0], acc[m][n][1]);
                }
        }
    }
}

// one warp per output element; partials folded lane-strided then by butterfly: a fixed order
__global__ void __launch_bounds__(256)
fold_eval_kernel(const double *__restrict__ partials, int nparts, int len, double *__restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int e = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (e >= len) return;
    double s = 0.0;
    for (int q = lane; q < nparts; q += 32) s += partials[(size_t)q * len + e];
    s = warp_sum(s);
    if (lane == 0) out[e] = s;
}

template <int ITER>
int launch_step(ate_ctx *ctx, const StepArgs &a, dim3 grid, size_t smem, bool init) {
    if (init) {
        auto k = irls_step_kernel<ITER, true>;
        if (cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)F_SMEM_LIMIT) != cudaSuccess) return -1;
        k<<<grid, F_THREADS, smem, ctx->stream>>>(a);
    } else {
        auto k = irls_step_kernel<ITER, false>;
        if (cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)F_SMEM_LIMIT) != cudaSuccess) return -1;
        k<<<grid, F_THREADS, smem, ctx->stream>>>(a);
    }
    return 0;
}

}  // namespace

int ate_i_irls_step(ate_ctx *ctx, const IrlsStep &in) {
    if (in.nrows <= 0 || in.ni <= 0) return ate_fail(ctx, -1, "irls_step: bad shape");
    const int ncols = in.ncols;
    if (ncols % 8 != 0 || ncols <= in.ni || in.iy >= ncols || ncols > in.ld)
        return ate_fail(ctx, -1, "irls_step: staged width %d must be a multiple of 8 that covers the response column", ncols);
    if ((in.ld & 1) != 0) return ate_fail(ctx, -1, "irls_step: ld must be even");
    const int nG = in.nG;
    const int nb = (nG + 31) / 32, nblk = nb * (nb + 1) / 2;
    const size_t stage_bytes = sizeof(double) * ((size_t)F_ROWS * (ncols + F_PAD) + 2 * F_ROWS);
    int stages = (int)((F_SMEM_LIMIT - 1024) / stage_bytes);
    if (stages > F_MAX_STAGES) stages = F_MAX_STAGES;
    if (stages < 2) return ate_fail(ctx, -1, "irls_step: %d staged columns do not fit the full-row kernel", ncols);
    const size_t smem = stages * stage_bytes + 3 * stages * sizeof(uint64_t);
    const int ngroups = nblk > 0 ? (nblk + F_CWARPS - 1) / F_CWARPS : 1;
    int nchunks = ctx->num_sms / ngroups;
    if (nchunks < 1) nchunks = 1;
    long long max_chunks = (in.nrows + 4 * F_ROWS - 1) / (4 * F_ROWS);
    if (max_chunks < 1) max_chunks = 1;
    if (nchunks > max_chunks) nchunks = (int)max_chunks;
    long long rpc = (in.nrows + nchunks - 1) / nchunks;
    rpc = (rpc + F_ROWS - 1) / F_ROWS * F_ROWS;
    nchunks = (int)((in.nrows + rpc - 1) / rpc);
    const int plen = 3 * in.ni + 3;
    const size_t gram_doubles = (size_t)1024 * nblk * nchunks;
    const size_t eval_doubles = (size_t)nchunks * F_EWARPS * plen;
    void *scr = nullptr;
    ATE_TRY(ate_gram_scratch(ctx, sizeof(double) * (gram_doubles + eval_doubles), &scr));
    StepArgs a;
    a.X = in.X; a.ld = in.ld; a.ncols = ncols; a.rows = in.rows;
    a.nrows = in.nrows;
    a.hole_begin = in.hole_len > 0 ? in.hole_begin : in.nrows;
    a.hole_len = in.hole_len > 0 ? in.hole_len : 0;
    a.rows_per_chunk = rpc;
    a.ni = in.ni; a.iy = in.iy; a.coef = in.coef; a.w = in.w;
    a.colidx = in.colidx; a.nblk = nblk; a.stages = stages;
    a.gram_partials = (double *)scr;
    a.eval_partials = (double *)scr + gram_doubles;
    a.v_out = in.v_out;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ctx->profile) {
        ATE_CUDA(ctx, cudaEventCreate(&e0));
        ATE_CUDA(ctx, cudaEventCreate(&e1));
        ATE_CUDA(ctx, cudaEventRecord(e0, ctx->stream));
    }
    const int it = (ncols + 63) / 64;
    int rc;
    const dim3 grid(ngroups, nchunks);
    if (it <= 1) rc = launch_step<1>(ctx, a, grid, smem, in.init != 0);
    else if (it <= 2) rc = launch_step<2>(ctx, a, grid, smem, in.init != 0);
    else if (it <= 3) rc = launch_step<3>(ctx, a, grid, smem, in.init != 0);
    else if (it <= 4) rc = launch_step<4>(ctx, a, grid, smem, in.init != 0);
    else if (it <= 5) rc = launch_step<5>(ctx, a, grid, smem, in.init != 0);
    else if (it <= 6) rc = launch_step<6>(ctx, a, grid, smem, in.init != 0);
    else if (it <= 8) rc = launch_step<8>(ctx, a, grid, smem, in.init != 0);
    else rc = -1;
    if (rc < 0) return ate_fail(ctx, -1, "irls_step: cannot launch for %d staged columns", ncols);
    ATE_LAUNCHED(ctx);
    if (ctx->profile) {
        ATE_CUDA(ctx, cudaEventRecord(e1, ctx->stream));
        ctx->prof_events.emplace_back(e0, e1);
        ctx->prof_shape.emplace_back(nG, in.nrows);
        ctx->prof_launches++;
        ctx->prof_flops += (double)in.nrows * ((double)nG * (nG + 1.0) + 6.0 * in.ni);
        ctx->prof_bytes += (double)in.nrows * (ncols + 2.0) * 8.0;  // staged rows read + v written
    }
    fold_eval_kernel<<<(plen + 7) / 8, 256, 0, ctx->stream>>>(a.eval_partials, nchunks * F_EWARPS, plen, in.sums);
    ATE_LAUNCHED(ctx);
    if (nG > 0) ATE_TRY(ate_i_gram_fold(ctx, a.gram_partials, nchunks, nblk, nG, in.ldh, in.H));
    return 0;
}
