// irls_round.cu -- one IRLS step of EVERY live L1-logistic fit of a cv.glmnet batch in one launch
// (glmnet's lognet behind ate_functions.R:139-141; the O(n) pieces of its outer loop).
//
// What a step needs from the rows of a fit: eta = a0 + x beta, q = plogis(eta) with glmnet's
// clamping, v = w q (1 - q), r = w (y - q), the deviance, X^T r and X^T v over all columns (KKT and
// strong-rule checks) and H = X_L^T V X_L over a column list L that covers the strong set (fp64 DMMA).
// The eleven fits of one cv.glmnet problem read the SAME rows (a cv-fold fit gives its held-out
// tenth weight 0), so the rows are staged once per CTA for up to two fits, and the CTAs that serve
// the other fits of the problem walk the same rows at the same time, which the 126 MB L2 absorbs.
//
// CTA = 8 MMA warps + 2 producer warps + 2 helper warps + 4 evaluation warps around a ring of
// 16- or 32-row slots in shared memory; one CTA per SM; a host-built descriptor per CTA says which
// rows and which fits it serves and which Gram tiles each MMA warp owns.
//   producers : a slot of a private contiguous row copy is ONE 1-D TMA bulk copy (the TMA unit
//               serves ~140 cycles per copy whatever its size); through a row list, one copy per row
//               issued by two warps.
//   eval      : lanes hold a column slice of beta for every fit of the CTA; a staged row is read
//               once for all fits.  Eight partial dots per fit are reduce-scattered over the warp so
//               that lane l ends with row l / 4; the four lanes of a row then each finish ONE fit
//               (exp, divide, log run once per lane whatever the number of fits).
//   helpers   : X^T r, X^T v (and sum v x^2 on the first step) on the FP64 FMA pipe, the columns split
//               between the two warps, every row read once for all fits.
//   MMA       : the lower triangle of each fit's H in 8x8 tiles; a warp owns up to four tile rows of
//               one 32-column segment of one fit (<= 16 tiles, 64 accumulator registers).
// Row-chunk partials are folded in a fixed order by round_fold_*: bit-reproducible for a given
// layout of the round.  Every barrier wait is bounded: a wait that lasts seconds traps instead of
// hanging the GPU.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <type_traits>
#include <vector>

#include "irls_round.cuh"
#include "path.cuh"
#include "tma_dmma.cuh"

namespace {

using namespace atk;

constexpr int R_CWARPS = 8, R_PWARPS = 2, R_HWARPS = 2, R_EWARPS = 4;
constexpr int R_THREADS = (R_CWARPS + R_PWARPS + R_HWARPS + R_EWARPS) * 32;
constexpr int R_PRODUCER = R_CWARPS, R_HELPER = R_PRODUCER + R_PWARPS, R_EVAL = R_HELPER + R_HWARPS;
constexpr int R_PAD = 4;
constexpr int R_MAX_STAGES = 8;
constexpr int R_FMAX = 4;  // fits served by one CTA
constexpr size_t R_SMEM_LIMIT = 227 * 1024;

struct DProb {
    const double *X;
    long long ld;
    const int32_t *rows;
    long long nrows;
    int padded, pad_;
};
struct DFit {
    const double *coef;
    const int *colidx;
    long long hole_begin, hole_len;
    double w;
    double *gram_part;  // [chunks][nblk][1024]
    double *eval_part;  // [chunks][3 ni + 3 R_EWARPS]
    int nblk, ones;     // ones >= 0: X^T v comes out of the Gram (RoundFit::ones)
};
struct DCta {
    int prob, nfit, chunk, pad;
    long long row_begin, row_end;
    int fit[R_FMAX], lead[R_FMAX];
    uint2 mma[R_CWARPS];  // per MMA warp two halves (x, y): fit slot | segment << 2 | count << 8 | row0 << 10 | row1 << 18
};
struct DFold {
    const double *gram_part, *eval_part;
    double *sums, *H;
    const int *colidx;
    int nchunks, nblk, nG, ldh, ones, ni;
};
struct RArgs {
    const DProb *probs;
    const DFit *fits;
    const DCta *ctas;
    int ncols, ni, iy, stages;
    int rolemap;    // 1: warps of a sub-partition share a role (see the kernel)
    int wait_hint;  // ns of suspend-time hint for the barrier waits (0: plain try_wait)
    int skip;  // measurement only (ate_round_kernel_bench): 1 MMA idle, 2 helpers idle, 4 eval without exp/log, 8 eval without dot
};

__device__ __forceinline__ unsigned long long gtimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// mbarrier wait that cannot hang the GPU: after ~4 s (a launch takes milliseconds) it traps.  hint_ns > 0
// passes try_wait a suspend-time hint (the warp sleeps in hardware until the phase completes or the time
// is up); 0 uses the plain form, which re-issues the test after a short hardware-chosen wait.
__device__ __forceinline__ void wait_g(uint64_t *bar, uint32_t parity, uint32_t hint_ns) {
    unsigned long long t0 = 0;
    for (unsigned spins = 1;; ++spins) {
        uint32_t done;
        if (hint_ns) {
            asm volatile(
                "{\n\t.reg .pred p;\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
                "selp.u32 %0, 1, 0, p;\n\t}"
                : "=r"(done)
                : "r"(smem_u32(bar)), "r"(parity), "r"(hint_ns)
                : "memory");
        } else {
            asm volatile(
                "{\n\t.reg .pred p;\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                "selp.u32 %0, 1, 0, p;\n\t}"
                : "=r"(done)
                : "r"(smem_u32(bar)), "r"(parity)
                : "memory");
        }
        if (done) return;
        if (hint_ns || (spins & 0xfffu) == 0) {
            const unsigned long long t = gtimer_ns();
            if (t0 == 0)
                t0 = t;
            else if (t - t0 > 4000000000ull)
                __trap();
        }
    }
}


// -DATE_ROUND_TRACE: CTA 0 stamps clock64() per slot and role into a device buffer which
// ate_round_kernel_bench writes to $ATE_ROUND_TRACE_OUT (tools/round_trace.py prints the timeline).
#ifdef ATE_ROUND_TRACE
constexpr int TR_SLOTS = 160, TR_EVENTS = 16;
__device__ long long g_rtrace[TR_SLOTS * TR_EVENTS + 8];
#define RTRACE(ev, it)                                                                          \
    do {                                                                                        \
        if (blockIdx.x == 0 && (it) < TR_SLOTS && (threadIdx.x & 31) == 0) g_rtrace[8 + (it) * TR_EVENTS + (ev)] = clock64(); \
    } while (0)
#else
#define RTRACE(ev, it) do { } while (0)
#endif

// One slot of one HALF of an MMA warp: tile rows a (NA tiles) and b (NB <= NA tiles) of one 32-column
// segment, all k-steps of the slot.  NA / NB are compile-time so that the loop body is straight-line
// DMMAs: a run-time tile count inside the loop costs a jump table and a WARPSYNC per tile row and
// k-step (measured: more than half of the MMA warps' issue slots).  sbase: the slot's rows (shared
// memory, bytes); the operand offsets are per-lane byte offsets into a row, row k4 of each k-step.
#ifndef ATE_MMA_UNROLL
#define ATE_MMA_UNROLL 2
#endif
// one-hot code of a half's tile counts (na >= nb)
__host__ __device__ constexpr unsigned pat_bit(int na, int nb) { return 1u << (na * 5 + nb); }
constexpr int MMA_UNROLL = ATE_MMA_UNROLL;  // k-steps per trip of the loop below (A/B builds: tools/build_variant_lib.sh)
template <int NA, int NB, int ROWS>
__device__ __forceinline__ void mma_half_slot(double (&accA)[4][2], double (&accB)[4][2], const char *sbase, uint32_t kstep_bytes,
                                              const double *vs, uint32_t offA0, uint32_t offA1, const uint32_t (&offB)[4]) {
#pragma unroll(MMA_UNROLL)
    for (int kk = 0; kk < ROWS / 4; ++kk) {
        const char *rowp = sbase + (size_t)kk * kstep_bytes;  // (warp-uniform)
#ifdef ATE_MMA_NOLOAD  // measurement only: the DMMA stream of this loop without its operand loads
        const double vv = 1.0 + kk;
        double fb[NA];
#pragma unroll
        for (int n = 0; n < NA; ++n) fb[n] = __longlong_as_double((long long)offB[n] + kk);
        const double fa = __longlong_as_double((long long)offA0 + kk) * vv;
        (void)rowp; (void)vs;
#else
        const double vv = vs[kk * 4];
        // v scales the A operand (one or two products per k-step) rather than the NA B operands
        double fb[NA];
#pragma unroll
        for (int n = 0; n < NA; ++n) fb[n] = *reinterpret_cast<const double *>(rowp + offB[n]);
        const double fa = *reinterpret_cast<const double *>(rowp + offA0) * vv;
#endif
#pragma unroll
        for (int n = 0; n < NA; ++n) dmma884(accA[n][0], accA[n][1], fa, fb[n]);
        if (NB > 0) {
#ifdef ATE_MMA_NOLOAD
            const double fa1 = __longlong_as_double((long long)offA1 + kk) * vv;
#else
            const double fa1 = *reinterpret_cast<const double *>(rowp + offA1) * vv;
#endif
#pragma unroll
            for (int n = 0; n < NB; ++n) dmma884(accB[n][0], accB[n][1], fa1, fb[n]);
        }
    }
}

template <int ROWS, int ITER, int FMAX, bool INIT>
__global__ void __launch_bounds__(R_THREADS, 1) irls_round_kernel(const RArgs a) {
    constexpr int PSHARE = ROWS / R_PWARPS;      // rows of a slot staged by one producer (row-list mode)
    constexpr int ITERH = (ITER + 1) / 2;        // column pairs per lane of an FMA-pipe helper
    constexpr bool HDMMA = !INIT && ITER <= 5;   // helpers on the DMMA pipe (rows up to 320 staged columns, not on the first step)
    constexpr int HT = HDMMA ? (ITER * 8 + 1) / 2 : 1;  // column tiles per helper warp (two helpers: the most any gets)
    constexpr int EWPS = ROWS / 8;               // eval warps per slot: each refreshes eight rows
    constexpr int ETURNS = R_EWARPS / EWPS;      // ... and they take turns on the slots when a slot needs fewer than all four
    static_assert(ROWS == 16 || ROWS == 32, "slot geometry");
    static_assert(FMAX >= 1 && FMAX <= 4 && FMAX <= R_FMAX, "fits per CTA");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ DCta scta;
    __shared__ int s_nnz[R_FMAX];
    const int ncols = a.ncols, pitch = ncols + R_PAD, stages = a.stages;
    const size_t stage_doubles = (size_t)ROWS * pitch;
    double *sx = reinterpret_cast<double *>(smem_raw);
    double *sv = sx + stage_doubles * stages;           // [stages][FMAX][ROWS]: v of every fit
    double *sr = sv + (size_t)stages * FMAX * ROWS;     // [stages][FMAX][ROWS]: r of every fit
    // the non-zero coefficients of every fit, compacted in ascending column order and padded to a multiple
    // of four with (zero column, 0.0): eta is a sparse dot product over them
    double *snzv = sr + (size_t)stages * FMAX * ROWS;                             // [FMAX][ncols] values
    uint16_t *snzi = reinterpret_cast<uint16_t *>(snzv + (size_t)FMAX * ncols);   // [FMAX][ncols] columns
    uint64_t *full = reinterpret_cast<uint64_t *>(snzi + (size_t)FMAX * ncols);  // (ncols % 8 == 0: aligned)
    uint64_t *vfull = full + stages, *empty = vfull + stages;

    // Hardware warp w runs on SM sub-partition w % 4.  With rolemap the roles are dealt so that the warps
    // of a sub-partition run the SAME loop (MMA on sub-partitions 0-1, producer + helpers on 2, eval on
    // 3): the sub-partition's instruction cache then holds one hot loop instead of three.
    const int hw_warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int warp = a.rolemap ? (((hw_warp & 3) << 2) | (hw_warp >> 2)) : hw_warp;
    const uint32_t whint = (uint32_t)a.wait_hint;
    if (threadIdx.x < (int)(sizeof(DCta) / sizeof(int)))
        reinterpret_cast<int *>(&scta)[threadIdx.x] = reinterpret_cast<const int *>(a.ctas + blockIdx.x)[threadIdx.x];
    __syncthreads();
    const DProb prob = a.probs[scta.prob];
    const long long c_begin = scta.row_begin, c_end = scta.row_end;
    const int nit = (int)((c_end - c_begin + ROWS - 1) / ROWS);
    const int nfit = scta.nfit;
    const bool contiguous = (prob.rows == nullptr && prob.ld == (long long)pitch && prob.padded != 0);
    // warp roles: 0-7 MMA, 8 producer, 9 second producer (row-list mode) or third helper, 10-11 helpers, 12-15 eval
    const int nhelp = contiguous ? 3 : 2;
    const bool is_producer = (warp == R_PRODUCER) || (warp == R_PRODUCER + 1 && !contiguous);
    const int hidx = (warp == R_HELPER) ? 0 : (warp == R_HELPER + 1) ? 1 : (warp == R_PRODUCER + 1 && contiguous) ? 2 : -1;

    for (int e = threadIdx.x; e < stages * ROWS * R_PAD; e += blockDim.x)  // the lists' "zero column" and "one column"
        sx[(size_t)(e / R_PAD) * pitch + ncols + (e % R_PAD)] = (e % R_PAD) == 1 ? 1.0 : 0.0;
    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(full + s, contiguous ? 1 : R_PWARPS);
            mbar_init(vfull + s, EWPS);
            mbar_init(empty + s, R_CWARPS + nhelp + R_EWARPS);  // EVERY consumer warp, every slot
        }
        mbar_fence_init();
    }
    __syncthreads();
    if (hw_warp < FMAX) {
        // warp f compacts the non-zero coefficients of fit f (a few dozen of ni on a lasso path)
        const int f = hw_warp;
        int nz = 0;
        if (f < nfit && !(a.skip & 8)) {
            const double *coef = a.fits[scta.fit[f]].coef;
            for (int c0 = 0; c0 < a.ni; c0 += 32) {
                const int c = c0 + lane;
                const double bv = c < a.ni ? __ldg(coef + c) : 0.0;
                const unsigned m = __ballot_sync(0xffffffffu, bv != 0.0);
                if (bv != 0.0) {
                    const int at = nz + __popc(m & ((1u << lane) - 1u));
                    snzv[(size_t)f * ncols + at] = bv;
                    snzi[(size_t)f * ncols + at] = (uint16_t)c;
                }
                nz += __popc(m);
            }
            const int nzp = (nz + 3) & ~3;
            if (lane < nzp - nz) {
                snzv[(size_t)f * ncols + nz + lane] = 0.0;
                snzi[(size_t)f * ncols + nz + lane] = (uint16_t)ncols;  // the staged rows' zero column
            }
            nz = nzp;
        }
        if (lane == 0) s_nnz[f] = nz;
    }
    __syncthreads();
    // Ring protocol: every consumer warp visits every slot in order and arrives on `empty` for it,
    // whether or not it has work there, and every producer waits on `empty` for every slot: no
    // barrier can complete a phase that one of its waiters has not seen the previous phase of.

    if (is_producer) {
        // ------------------------------ producers ------------------------------
        const int pw = warp - R_PRODUCER;
        if (contiguous) {
            const uint32_t rowb = (uint32_t)pitch * 8u;
            for (int it = 0; it < nit; ++it) {
                const int s = it % stages;
                const uint32_t ph = (uint32_t)(it / stages) & 1u;
                const long long t0 = c_begin + (long long)it * ROWS;
                const int cnt = (int)((c_end - t0) < ROWS ? (c_end - t0) : ROWS);
                wait_g(empty + s, ph ^ 1u, whint);
                RTRACE(0, it);  // producer: slot free
                double *dst0 = sx + stage_doubles * s;
                if (lane < ROWS && lane >= cnt)  // a short last slot: the missing rows read as zeros
                    for (int c = 0; c < ncols; ++c) dst0[(size_t)lane * pitch + c] = 0.0;
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive_expect_tx(full + s, rowb * (uint32_t)cnt);
                    bulk_g2s(dst0, prob.X + (size_t)t0 * (size_t)prob.ld, rowb * (uint32_t)cnt, full + s);
                }
                __syncwarp();
            }
        } else {
            // one bulk copy per row; row ids are fetched PF slots ahead so that the latency of the
            // list never sits between "slot free" and "copy issued"
            constexpr int PF = 4;
            const int myrow = PSHARE * pw + lane;  // meaningful for lane < PSHARE
            const uint32_t row_bytes = (uint32_t)ncols * 8u;
            auto fetch = [&](int it, long long &r) -> bool {
                const long long t = c_begin + (long long)it * ROWS + myrow;
                if (lane >= PSHARE || it >= nit || t >= c_end) return false;
                r = prob.rows ? (long long)__ldg(prob.rows + t) : t;
                return true;
            };
            long long ring_r[PF];
            bool ring_ok[PF];
#pragma unroll
            for (int d = 0; d < PF; ++d) {
                ring_r[d] = 0;
                ring_ok[d] = fetch(d, ring_r[d]);
            }
            for (int it0 = 0; it0 < nit; it0 += PF) {
#pragma unroll
                for (int d = 0; d < PF; ++d) {
                    const int it = it0 + d;
                    if (it >= nit) break;
                    const long long cur_r = ring_r[d];
                    const bool cur_ok = ring_ok[d];
                    ring_ok[d] = fetch(it + PF, ring_r[d]);
                    const int s = it % stages;
                    const uint32_t ph = (uint32_t)(it / stages) & 1u;
                    wait_g(empty + s, ph ^ 1u, whint);
                    double *dst = sx + stage_doubles * s + (size_t)myrow * pitch;
                    if (lane < PSHARE && !cur_ok)
                        for (int c = 0; c < ncols; ++c) dst[c] = 0.0;
                    const unsigned vmask = __ballot_sync(0xffffffffu, cur_ok);
                    __syncwarp();
                    if (lane == 0) mbar_arrive_expect_tx(full + s, row_bytes * (uint32_t)__popc(vmask));
                    __syncwarp();
                    if (cur_ok) bulk_g2s(dst, prob.X + (size_t)cur_r * (size_t)prob.ld, row_bytes, full + s);
                }
            }
        }
    } else if (warp >= R_EVAL) {
        // ------------------------------ evaluation ------------------------------
        // eta = a0 + x . beta is a SPARSE dot product over the fit's non-zero coefficients (a few dozen of ni
        // on a lasso path) on the FMA pipe: as an 8 x 8 DMMA accumulation (eight rows x eight fits) it took
        // 21-31 % of the DMMA pipe, which the Gram needs, for at most half-used tiles.  A quad of lanes owns a
        // row; a fit's entries form four interleaved subsets (t mod 4) summed as (s0 + s1) + (s2 + s3)
        // whatever the number of fits, so a fit's eta does not depend on which fits share its CTA: with
        // three or four fits a lane sums all four subsets of its fit, with two fits two lanes share a fit,
        // with one fit the four lanes take one subset each.  Lane (row g, k4) then holds fit k4, and exp /
        // divide / log run once per lane.  v and r are published (and the slot released) BEFORE the log of
        // the deviance is taken.
        const int e = warp - R_EVAL;
        const int ni = a.ni;
        const int grp = e % EWPS, turn = e / EWPS;
        const int g = lane >> 2, k4 = lane & 3;
        const int fsel = k4;
        const bool mine = fsel < nfit && fsel < FMAX;
        const double pmin = GLM_PMIN, fmax_ = log(1.0 / pmin - 1.0), fmin_ = -fmax_;
        double a0 = 0.0, w = 0.0;
        long long hb = 0, he = 0;
        bool lead = false;
        if (mine) {
            const DFit &ft = a.fits[scta.fit[fsel]];
            a0 = ft.coef[ni];
            w = ft.w;
            hb = ft.hole_begin;
            he = ft.hole_begin + ft.hole_len;
            lead = scta.lead[fsel] != 0;
        }
        double s_xmz = 0.0, s_sr = 0.0, s_dev = 0.0, lik = 1.0;
        int nlik = 0;
        const int r0 = 8 * grp, rr = r0 + g;
        // this lane's share of the dot products: fit dfit, subsets [sub0, sub0 + nsub)
        const int dfit = nfit >= 3 ? k4 : nfit == 2 ? (k4 & 1) : 0;
        const int sub0 = nfit >= 3 ? 0 : nfit == 2 ? 2 * (k4 >> 1) : k4;
        const int nnz = (dfit < nfit && dfit < FMAX) ? s_nnz[dfit] : 0;
        const double *nzv = snzv + (size_t)(dfit < FMAX ? dfit : 0) * ncols + sub0;
        const uint16_t *nzi = snzi + (size_t)(dfit < FMAX ? dfit : 0) * ncols + sub0;
        for (int it = 0; it < nit; ++it) {
            const int s = it % stages;
            const uint32_t ph = (uint32_t)(it / stages) & 1u;
            wait_g(full + s, ph, whint);
            if ((it % ETURNS) != turn) {  // not this warp's slot: release it at once
                if (lane == 0) mbar_arrive(empty + s);
                continue;
            }
            RTRACE(1 + (grp & 1), it);  // eval warp of group 0 / 1: rows landed
            const double *slot = sx + stage_doubles * s;
            const double *xrow = slot + (size_t)rr * pitch;
            double dot;
            if (nfit >= 3) {
                double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
#pragma unroll 2
                for (int t = 0; t < nnz; t += 4) {
                    const ushort4 ci = *reinterpret_cast<const ushort4 *>(nzi + t);
                    const double2 b01 = *reinterpret_cast<const double2 *>(nzv + t), b23 = *reinterpret_cast<const double2 *>(nzv + t + 2);
                    s0 = fma(xrow[ci.x], b01.x, s0);
                    s1 = fma(xrow[ci.y], b01.y, s1);
                    s2 = fma(xrow[ci.z], b23.x, s2);
                    s3 = fma(xrow[ci.w], b23.y, s3);
                }
                dot = (s0 + s1) + (s2 + s3);
            } else if (nfit == 2) {
                double sa = 0.0, sb = 0.0;
#pragma unroll 2
                for (int t = 0; t < nnz; t += 4) {
                    const ushort2 ci = *reinterpret_cast<const ushort2 *>(nzi + t);
                    const double2 bb = *reinterpret_cast<const double2 *>(nzv + t);
                    sa = fma(xrow[ci.x], bb.x, sa);
                    sb = fma(xrow[ci.y], bb.y, sb);
                }
                const double part = sa + sb;  // (s0 + s1) on lanes k4 < 2, (s2 + s3) on the others
                const double lo = __shfl_sync(0xffffffffu, part, (lane & ~3) | (k4 & 1));
                const double hi = __shfl_sync(0xffffffffu, part, (lane & ~3) | (k4 & 1) | 2);
                dot = lo + hi;
            } else {
                double sa = 0.0;
#pragma unroll 4
                for (int t = 0; t < nnz; t += 4) sa = fma(xrow[nzi[t]], nzv[t], sa);
                const double s01 = sa + __shfl_xor_sync(0xffffffffu, sa, 1);  // s0 + s1 | s2 + s3 (commutative: same bits on both lanes)
                const double oth = __shfl_xor_sync(0xffffffffu, s01, 2);
                dot = k4 < 2 ? s01 + oth : oth + s01;                         // (s0 + s1) + (s2 + s3)
            }
            const long long t = c_begin + (long long)it * ROWS + rr;
            const bool valid = mine && t < c_end && !(t >= hb && t < he);
            const double yi = slot[(size_t)rr * pitch + a.iy];
            const double fi = a0 + dot;
            double q;
            if (a.skip & 4)
                q = 0.5 + 1e-3 * fi;
            else if (fi < fmin_)
                q = 0.0;
            else if (fi > fmax_)
                q = 1.0;
            else
                q = 1.0 / (1.0 + exp(-fi));
            const double vi = valid ? w * q * (1.0 - q) : 0.0, ri = valid ? w * (yi - q) : 0.0;
            if (mine) {
                sv[((size_t)s * FMAX + fsel) * ROWS + rr] = vi;
                sr[((size_t)s * FMAX + fsel) * ROWS + rr] = ri;
            }
            __syncwarp();
            RTRACE(3 + (grp & 1), it);  // eval: v, r published
            if (lane == 0) {
                mbar_arrive(vfull + s);  // MMA and helper warps may start on this slot
                mbar_arrive(empty + s);  // ... and this warp is done with its staged rows
            }
            if (!(a.skip & 4)) {
                // log-likelihood term, off the slot's critical path.  y is 0/1 on this path, so the term of a row
                // is log(y ? q : 1 - q): the lane multiplies these probabilities up (each >= pmin = 1e-5, so
                // sixteen of them stay above 1e-80) and takes ONE log per sixteen slots -- the log was a fifth of
                // the eval warps' chain.  A non-binary response takes the general form row by row.
                const double qc = fmin(fmax(q, pmin), 1.0 - pmin);
                const bool count = lead && valid;
                const bool binary = (yi == 0.0 || yi == 1.0);
                if (__all_sync(0xffffffffu, binary || !count)) {
                    lik *= count ? (yi == 1.0 ? qc : 1.0 - qc) : 1.0;
                    if (++nlik == 16) {
                        s_dev -= w * log(lik);
                        lik = 1.0;
                        nlik = 0;
                    }
                } else if (count) {
                    s_dev -= w * (yi * log(qc) + (1.0 - yi) * log(1.0 - qc));
                }
                if (count) {
                    s_xmz += vi;
                    s_sr += ri;
                }
            }
        }
        s_dev -= w * log(lik);
        // lanes with the same k4 hold partial sums of the same fit
#pragma unroll
        for (int o = 16; o >= 4; o >>= 1) {
            s_xmz += __shfl_xor_sync(0xffffffffu, s_xmz, o);
            s_sr += __shfl_xor_sync(0xffffffffu, s_sr, o);
            s_dev += __shfl_xor_sync(0xffffffffu, s_dev, o);
        }
        if (lane < 4 && lead) {
            double *ep = a.fits[scta.fit[fsel]].eval_part + (size_t)scta.chunk * (3 * ni + 3 * R_EWARPS) + 3 * ni + 3 * e;
            ep[0] = s_xmz;
            ep[1] = s_sr;
            ep[2] = s_dev;
        }
    } else if (hidx >= 0) {
        // ------------------------------ helpers: X^T r, X^T v (and sum v x^2 on the first step) -----
        const int h = hidx;
        const int ni = a.ni;
        bool any_lead = false;
#pragma unroll
        for (int f = 0; f < FMAX; ++f) any_lead |= (f < nfit && scta.lead[f] != 0);
        if (HDMMA && nfit > 1) {
            // [r_0 v_0 r_1 v_1 ...]^T X on the DMMA pipe: the eight rows of A are (fit, r | v) pairs, a tile
            // of B is four staged rows x eight CONSECUTIVE columns (conflict-free loads), so one DMMA per
            // tile and k-step serves X^T r and X^T v of up to four fits.  The column tiles are split
            // between the helper warps (three when one producer suffices, else two).
            // Every helper takes the same number of tiles, a compile-time constant per helper count, and
            // issues them WITHOUT a guard: a guarded mma.sync costs a re-convergence barrier per instruction
            // (measured: ~25 cycles each, more than the DMMA itself).  Tiles past the last column read
            // whatever follows the row (finite staged data) and their sums are never written out.
            const int ntile = ncols >> 3;
            const int g = lane >> 2, k4 = lane & 3;
            const int af = g >> 1;               // the fit whose r (g even) or v (g odd) this lane feeds into A
            const bool a_ok = af < nfit && af < FMAX;
            double acc[HT][2];
#pragma unroll
            for (int t = 0; t < HT; ++t) acc[t][0] = acc[t][1] = 0.0;
            constexpr int HT3 = (ITER * 8 + 2) / 3;   // tiles per helper when there are three
            const int tpw = nhelp == 3 ? HT3 : HT;
            const int tile0 = h * tpw;
            for (int it = 0; it < nit; ++it) {
                const int s = it % stages;
                const uint32_t ph = (uint32_t)(it / stages) & 1u;
                wait_g(full + s, ph, whint);
                if (any_lead && !(a.skip & 2)) {
                    wait_g(vfull + s, ph, whint);
                    if (h < 2) RTRACE(5 + h, it);  // helper h: start
                    const double *slot = sx + stage_doubles * s;
                    const double *srcA = ((g & 1) ? sv : sr) + ((size_t)s * FMAX + (a_ok ? af : 0)) * ROWS;
                    if (nhelp == 3) {
#pragma unroll 1
                        for (int kk = 0; kk < ROWS / 4; ++kk) {
                            const int r = kk * 4 + k4;
                            const double av = a_ok ? srcA[r] : 0.0;
                            const double *rowp = slot + (size_t)r * pitch + tile0 * 8 + g;
#pragma unroll
                            for (int t = 0; t < HT3; ++t) dmma884(acc[t][0], acc[t][1], av, rowp[t * 8]);
                        }
                    } else {
#pragma unroll 1
                        for (int kk = 0; kk < ROWS / 4; ++kk) {
                            const int r = kk * 4 + k4;
                            const double av = a_ok ? srcA[r] : 0.0;
                            const double *rowp = slot + (size_t)r * pitch + tile0 * 8 + g;
#pragma unroll
                            for (int t = 0; t < HT; ++t) dmma884(acc[t][0], acc[t][1], av, rowp[t * 8]);
                        }
                    }
                }
                __syncwarp();
                if (h < 2) RTRACE(7 + h, it);  // helper h: done
                if (lane == 0) mbar_arrive(empty + s);
            }
            // C[m][n]: lane (g, k4) holds m = g (fit g / 2, r | v), columns 8 (tile0 + t) + 2 k4 + {0, 1}
            if (a_ok && scta.lead[af] != 0) {
                double *ep = a.fits[scta.fit[af]].eval_part + (size_t)scta.chunk * (3 * ni + 3 * R_EWARPS) + ((g & 1) ? ni : 0);
#pragma unroll
                for (int t = 0; t < HT; ++t)
                    if (t < tpw && tile0 + t < ntile) {
                        const int c = (tile0 + t) * 8 + 2 * k4;
                        if (c < ni) ep[c] = acc[t][0];
                        if (c + 1 < ni) ep[c + 1] = acc[t][1];
                    }
            }
        } else {
            // FMA-pipe form, for a CTA that serves ONE fit (first step of a fit, rows wider than 320 columns,
            // lists too wide to share a CTA): an 8-row MMA tile would carry two useful rows, so the plain
            // FMA pipe is cheaper here and leaves the DMMA pipe to the Gram.  Lanes own column pairs, the
            // columns are split between the helper warps.
            const int npairs = ncols >> 1, ppw = (npairs + nhelp - 1) / nhelp;
            const bool xtv_from_gram = nfit > 0 && a.fits[scta.fit[0]].ones >= 0;
            double gr[2 * ITERH], hx[2 * ITERH], hxx[INIT ? 2 * ITERH : 1];
#pragma unroll
            for (int i = 0; i < 2 * ITERH; ++i) {
                gr[i] = hx[i] = 0.0;
                if (INIT) hxx[i] = 0.0;
            }
            int col[ITERH];
#pragma unroll
            for (int i = 0; i < ITERH; ++i) {
                const int idx = lane + 32 * i;
                col[i] = (idx < ppw && h * ppw + idx < npairs) ? 2 * (h * ppw + idx) : -1;
            }
            for (int it = 0; it < nit; ++it) {
                const int s = it % stages;
                const uint32_t ph = (uint32_t)(it / stages) & 1u;
                wait_g(full + s, ph, whint);
                if (any_lead && !(a.skip & 2)) {
                    wait_g(vfull + s, ph, whint);
                    if (h < 2) RTRACE(5 + h, it);  // helper h: start
                    const double *slot = sx + stage_doubles * s;
                    if (!INIT && xtv_from_gram) {
                        // X^T v of the list columns comes out of the Gram (its ONE column): X^T r alone here.
                        // Under the MMA warps' DMMA stream every fp64 FMA of a helper waits its turn on the
                        // shared pipe, and the helpers of a single-fit CTA paced the whole ring (measured).
#pragma unroll 4
                        for (int r = 0; r < ROWS; ++r) {
                            const double *rowp = slot + (size_t)r * pitch;
                            const double rj = sr[(size_t)s * FMAX * ROWS + r];
#pragma unroll
                            for (int i = 0; i < ITERH; ++i) {
                                const double2 x = col[i] >= 0 ? *reinterpret_cast<const double2 *>(rowp + col[i]) : make_double2(0.0, 0.0);
                                gr[2 * i] = fma(rj, x.x, gr[2 * i]);
                                gr[2 * i + 1] = fma(rj, x.y, gr[2 * i + 1]);
                            }
                        }
                    } else {
#pragma unroll 2
                        for (int r = 0; r < ROWS; ++r) {
                            const double *rowp = slot + (size_t)r * pitch;
                            const double vj = sv[(size_t)s * FMAX * ROWS + r];
                            const double rj = sr[(size_t)s * FMAX * ROWS + r];
#pragma unroll
                            for (int i = 0; i < ITERH; ++i) {
                                const double2 x = col[i] >= 0 ? *reinterpret_cast<const double2 *>(rowp + col[i]) : make_double2(0.0, 0.0);
                                gr[2 * i] = fma(rj, x.x, gr[2 * i]);
                                gr[2 * i + 1] = fma(rj, x.y, gr[2 * i + 1]);
                                hx[2 * i] = fma(vj, x.x, hx[2 * i]);
                                hx[2 * i + 1] = fma(vj, x.y, hx[2 * i + 1]);
                                if (INIT) {
                                    hxx[2 * i] = fma(vj * x.x, x.x, hxx[2 * i]);
                                    hxx[2 * i + 1] = fma(vj * x.y, x.y, hxx[2 * i + 1]);
                                }
                            }
                        }
                    }
                }
                __syncwarp();
                if (h < 2) RTRACE(7 + h, it);  // helper h: done
                if (lane == 0) mbar_arrive(empty + s);
            }
            if (nfit > 0 && scta.lead[0] != 0) {
                double *ep = a.fits[scta.fit[0]].eval_part + (size_t)scta.chunk * (3 * ni + 3 * R_EWARPS);
#pragma unroll
                for (int i = 0; i < ITERH; ++i)
#pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
                        const int c = col[i] + hh;
                        if (col[i] >= 0 && c < ni) {
                            ep[c] = gr[2 * i + hh];
                            ep[ni + c] = hx[2 * i + hh];
                            if (INIT) ep[2 * ni + c] = hxx[2 * i + hh];
                        }
                    }
            }
        }
    } else if (warp < R_CWARPS) {
        // ------------------------------ MMA ------------------------------
        // two halves per warp; a half = up to two tile rows of one 32-column segment of one fit
        const uint2 d = scta.mma[warp];
        const int g = lane >> 2, k4 = lane & 3;
        int fsl[2], seg[2], nt[4], tr[4], pat[2];
        uint32_t offA[4], offB[2][4];  // per-lane byte offsets into a staged row (row k4 of a k-step folded in)
        bool any = false;
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
            const unsigned wd = hh ? d.y : d.x;
            fsl[hh] = (int)(wd & 3u);
            seg[hh] = (int)((wd >> 2) & 63u);
            const int cnt = (int)((wd >> 8) & 3u);
            const int *colidx = cnt > 0 ? a.fits[scta.fit[fsl[hh]]].colidx : nullptr;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int u = 2 * hh + j;
                tr[u] = (int)((wd >> (10 + 8 * j)) & 255u);
                const int lim = tr[u] - 4 * seg[hh] + 1;  // tiles left of (and on) the diagonal in this segment
                nt[u] = j < cnt ? (lim < 4 ? lim : 4) : 0;
                offA[u] = (uint32_t)((nt[u] > 0 ? __ldg(colidx + tr[u] * 8 + g) : 0) + k4 * pitch) * 8u;
            }
#pragma unroll
            for (int n = 0; n < 4; ++n)
                offB[hh][n] = (uint32_t)((cnt > 0 ? __ldg(colidx + (4 * seg[hh] + n) * 8 + g) : 0) + k4 * pitch) * 8u;
            any |= cnt > 0;
            // the planner puts the heavier tile row first; (0, 0) is an empty half
            pat[hh] = nt[2 * hh] > 0 ? (int)pat_bit(nt[2 * hh], nt[2 * hh + 1]) : 0;
        }
        const uint32_t kstep_bytes = (uint32_t)pitch * 32u;  // four rows per k-step
        double acc[4][4][2];
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int n = 0; n < 4; ++n) acc[u][n][0] = acc[u][n][1] = 0.0;
        for (int it = 0; it < nit; ++it) {
            const int s = it % stages;
            const uint32_t ph = (uint32_t)(it / stages) & 1u;
            wait_g(full + s, ph, whint);
            if (any && !(a.skip & 1)) {
                wait_g(vfull + s, ph, whint);
                if (warp < 3) RTRACE(9 + 2 * warp, it);  // MMA warp 0..2: start
                const char *sbase = reinterpret_cast<const char *>(sx + stage_doubles * s);
#ifdef ATE_MMA_TWICE  // measurement only: every slot's DMMA work issued twice
                for (int rep = 0; rep < 2; ++rep)
#endif
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    const double *vs = sv + ((size_t)s * FMAX + fsl[hh]) * ROWS + k4;
                    // one (warp-uniform) dispatch per half and slot, then straight-line loops.  The pattern is
                    // a ONE-HOT mask tested bit by bit, commonest first: a switch over dense codes becomes a
                    // jump table (LDC + BRX + WARPSYNC, ~90 cycles per dispatch, measured), a bit test is a
                    // predicated branch.
#define MMA_HALF(NA_, NB_)                                                                                          \
    if (pat[hh] & pat_bit(NA_, NB_)) {                                                                               \
        mma_half_slot<NA_, NB_, ROWS>(acc[2 * hh], acc[2 * hh + 1], sbase, kstep_bytes, vs, offA[2 * hh], offA[2 * hh + 1], \
                                      offB[hh]);                                                                     \
    } else
                    MMA_HALF(4, 4) MMA_HALF(4, 3) MMA_HALF(4, 2) MMA_HALF(4, 1) MMA_HALF(3, 2) MMA_HALF(4, 0)
                    MMA_HALF(3, 1) MMA_HALF(2, 1) MMA_HALF(3, 0) MMA_HALF(2, 0) MMA_HALF(1, 0)
                    if (pat[hh] != 0) __trap();  // not a pair the planner deals (an empty half is 0)
#undef MMA_HALF
                }
            }
            __syncwarp();
            if (warp < 3) RTRACE(10 + 2 * warp, it);  // MMA warp 0..2: done
            if (lane == 0) mbar_arrive(empty + s);
        }
        // tile (tr[u], 4 seg + n) sits in 32x32 block (tr[u] / 4, seg) of the fold kernel's layout
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (nt[u] > 0) {
                const DFit &ft = a.fits[scta.fit[fsl[u >> 1]]];
                const int bi = tr[u] >> 2, m = tr[u] & 3;
                double *out = ft.gram_part + ((size_t)scta.chunk * ft.nblk + (bi * (bi + 1) / 2 + seg[u >> 1])) * 1024;
#pragma unroll
                for (int n = 0; n < 4; ++n)
                    if (n < nt[u]) {
                        const int r = m * 8 + g, c = n * 8 + 2 * k4;
                        *reinterpret_cast<double2 *>(out + r * 32 + c) = make_double2(acc[u][n][0], acc[u][n][1]);
                    }
            }
        }
    }
}

// sums of one fit: one warp per output element; chunk partials folded lane-strided then by butterfly
__global__ void __launch_bounds__(256)
round_fold_eval_kernel(const DFold *__restrict__ folds, int ni, int init) {
    const DFold fd = folds[blockIdx.y];
    const int lane = threadIdx.x & 31;
    const int e = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int plen = 3 * ni + 3, ep = 3 * ni + 3 * R_EWARPS;
    if (e >= plen) return;
    double s = 0.0;
    if (e < 3 * ni) {
        if (fd.ones >= 0 && e >= ni && e < 2 * ni) {  // X^T v: the list columns' entries are written by
            if (lane == 0) fd.sums[e] = 0.0;          // round_fold_gram_kernel (launched after this one) from
            return;                                   // the Gram's ONE column; the solve reads no others
        }
        if (e < 2 * ni || init)
            for (int q = lane; q < fd.nchunks; q += 32) s += fd.eval_part[(size_t)q * ep + e];
    } else {
        const int j = e - 3 * ni;
        for (int q = lane; q < fd.nchunks * R_EWARPS; q += 32)
            s += fd.eval_part[(size_t)(q / R_EWARPS) * ep + 3 * ni + 3 * (q % R_EWARPS) + j];
    }
    s = warp_sum(s);
    if (lane == 0) fd.sums[e] = s;
}

// H of one fit (both triangles, ldh x ldh) from the lower-triangular 8x8 tiles of its chunk partials
__global__ void __launch_bounds__(256)
round_fold_gram_kernel(const DFold *__restrict__ folds) {
    const DFold fd = folds[blockIdx.y];
    const int ldh = fd.ldh;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= ldh * ldh) return;
    const int r = e / ldh, c = e - r * ldh;
    const int npad = (fd.nG + 7) / 8 * 8;  // tile rows the plan covers
    double s = 0.0;
    if (r < npad && c < npad) {
        const int rr = (r >> 3) >= (c >> 3) ? r : c, cc = (r >> 3) >= (c >> 3) ? c : r;
        const int bi = rr >> 5, bj = cc >> 5;
        const double *src = fd.gram_part + (size_t)(bi * (bi + 1) / 2 + bj) * 1024 + (rr & 31) * 32 + (cc & 31);
        const size_t stride = (size_t)fd.nblk * 1024;
        for (int q = 0; q < fd.nchunks; ++q) s += src[q * stride];
    }
    fd.H[e] = s;
    // the ONE column of the list: H[c][ones] = sum_i v_i x_i,c = (X^T v)[colidx[c]]
    if (fd.ones >= 0 && c == fd.ones && r < fd.ones) fd.sums[fd.ni + fd.colidx[r]] = s;
}

// ---- host: who computes which Gram tiles ----------------------------------------------------------
// For a list of ntr = ceil(nG / 8) tile rows: segment sg (columns 32 sg .. 32 sg + 31 of the list)
// holds tile rows r >= 4 sg with min(4, r - 4 sg + 1) tiles each.  The unit of work is a HALF: one or two
// tile rows of ONE segment of ONE fit (they share the four B operands), at most eight tiles.  An MMA warp
// owns two halves (16 tiles, 64 accumulator registers), which may belong to different fits: narrow lists
// of several fits then fill a CTA's eight warps, and the staged rows serve up to four fits.
struct Half {
    int fit, seg, cnt, load, rows[2];
};
// halves of one fit: within a segment the rows are paired heaviest with lightest
void make_halves(int fit, int ntr, std::vector<Half> &out) {
    const int nseg = (ntr + 3) / 4;
    for (int sg = 0; sg < nseg; ++sg) {
        std::vector<int> rows;
        for (int r = 4 * sg; r < ntr; ++r) rows.push_back(r);
        auto tiles = [&](int r) { return r - 4 * sg + 1 < 4 ? r - 4 * sg + 1 : 4; };
        std::stable_sort(rows.begin(), rows.end(), [&](int x, int y) { return tiles(x) > tiles(y); });
        size_t lo = 0, hi = rows.size();
        while (lo < hi) {
            Half h{fit, sg, 0, 0, {0, 0}};
            h.rows[h.cnt++] = rows[lo];
            h.load += tiles(rows[lo]);
            ++lo;
            if (lo < hi) {
                --hi;
                h.rows[h.cnt++] = rows[hi];
                h.load += tiles(rows[hi]);
            }
            out.push_back(h);
        }
    }
}

struct Group {
    int prob = 0;
    int nfit = 0, fit[R_FMAX] = {0, 0, 0, 0}, lead[R_FMAX] = {0, 0, 0, 0};
    std::vector<Half> halves;   // .fit = fit SLOT; <= 2 R_CWARPS, dealt to the warps by deal_halves
    Half warp_half[R_CWARPS][2];
    int tiles = 0;
    double cost = 0.0;
    int chunks = 1;
};

// two halves per warp: heaviest first onto the lightest warp that has room, preferring a warp that
// already holds a half of the same fit and segment (its B operands are then loaded once)
// `handicap`: tiles per k-step an MMA warp of sub-partitions 1-3 starts with.  MMA warp w runs on
// sub-partition w % 4; the three helper warps sit on sub-partitions 1-3 and their X^T r / X^T v tiles use
// the same fp64 datapath (staged columns / 8 tiles per k-step in a CTA of several fits), while
// sub-partition 0 holds only the producer -- its two MMA warps take more of the Gram.
void deal_halves(Group &g, int handicap) {
    for (int w = 0; w < R_CWARPS; ++w) g.warp_half[w][0] = g.warp_half[w][1] = Half{0, 0, 0, 0, {0, 0}};
    std::vector<Half> hs = g.halves;
    std::stable_sort(hs.begin(), hs.end(), [](const Half &x, const Half &y) { return x.load > y.load; });
    int cnt[R_CWARPS] = {0}, load[R_CWARPS] = {0};
    for (int w = 0; w < R_CWARPS; ++w) load[w] = (w & 3) ? handicap : 0;
    for (const Half &h : hs) {
        int best = -1;
        for (int w = 0; w < R_CWARPS; ++w) {
            if (cnt[w] >= 2) continue;
            if (best < 0 || load[w] < load[best]) best = w;
        }
        // a partner of the same fit and segment with (almost) the same load wins
        for (int w = 0; w < R_CWARPS; ++w)
            if (cnt[w] == 1 && g.warp_half[w][0].fit == h.fit && g.warp_half[w][0].seg == h.seg && best >= 0 &&
                load[w] <= load[best] + 2)
                best = w;
        if (best < 0) return;  // cannot happen: the caller admits at most 2 R_CWARPS halves
        g.warp_half[best][cnt[best]++] = h;
        load[best] += h.load;
    }
    // The slot is released when the SLOWEST warp is through its tiles (an MMA warp issues its DMMAs at its
    // own pace, whatever the other warp of its sub-partition does), so the busiest warp paces the ring:
    // while it helps, move one tile row out of a two-row half of the busiest warp to the lightest warp that
    // still has a free half (both then hold single-row halves of that segment).
    auto row_tiles = [](const Half &h, int j) { const int t = h.rows[j] - 4 * h.seg + 1; return t < 4 ? t : 4; };
    for (int pass = 0; pass < 4 * R_CWARPS; ++pass) {
        int hi = 0, lo = -1;
        for (int w = 1; w < R_CWARPS; ++w)
            if (load[w] > load[hi]) hi = w;
        for (int w = 0; w < R_CWARPS; ++w)
            if (w != hi && cnt[w] < 2 && (lo < 0 || load[w] < load[lo])) lo = w;
        if (lo < 0) break;
        int bq = -1, bj = -1, bt = 0;  // the row whose move lowers max(load[hi], load[lo]) most
        for (int q = 0; q < cnt[hi]; ++q) {
            const Half &h = g.warp_half[hi][q];
            if (h.cnt != 2) continue;
            for (int j = 0; j < 2; ++j) {
                const int t = row_tiles(h, j);
                const int worst = std::max(load[hi] - t, load[lo] + t);
                if (worst < load[hi] && (bq < 0 || worst < std::max(load[hi] - bt, load[lo] + bt))) {
                    bq = q; bj = j; bt = t;
                }
            }
        }
        if (bq < 0) break;
        Half &h = g.warp_half[hi][bq];
        Half moved{h.fit, h.seg, 1, bt, {h.rows[bj], 0}};
        const int keep = h.rows[1 - bj];
        h.rows[0] = keep; h.rows[1] = 0; h.cnt = 1; h.load -= bt;
        g.warp_half[lo][cnt[lo]++] = moved;
        load[hi] -= bt;
        load[lo] += bt;
    }
}

// Which fits share a CTA.  Fits of one problem are sorted by list width; a list with more halves than one
// CTA holds is split over several CTAs that stage the same rows (the first leads: it alone evaluates the
// sums); the others are packed first-fit, up to `fmax` fits and 2 R_CWARPS halves per CTA.
// Returns -1 for a list the tile numbering cannot hold.
int build_groups(int P, const std::vector<int> &prob_of, const std::vector<int> &live, const std::vector<int> &nGs,
                 int fmax, int ncs, std::vector<Group> &groups) {
    // helper tiles per k-step and MMA warp of sub-partitions 1-3 (two MMA warps each): DMMA helpers of a
    // CTA with several fits issue ncs / 8 tiles over three warps; the FMA helpers of a single fit cost
    // about a tile's worth of pipe time per warp
    static const int hc_scale = getenv("ATE_ROUND_HANDICAP") ? atoi(getenv("ATE_ROUND_HANDICAP")) : 100;  // percent (A/B measurements)
    const int hc_multi = (ncs / 8 + 5) / 6 * hc_scale / 100, hc_single = 1 * hc_scale / 100;
    const int nfits = (int)prob_of.size();
    const int cap = 2 * R_CWARPS;
    std::vector<Half> hv;
    std::vector<int> order;
    for (int p = 0; p < P; ++p) {
        order.clear();
        for (int f = 0; f < nfits; ++f)
            if (live[f] && prob_of[f] == p) order.push_back(f);
        std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return nGs[x] > nGs[y]; });
        std::vector<Group> open;
        for (int f : order) {
            const int nG = nGs[f];
            if (nG < 0 || nG > 255 * 8) return -1;
            const int ntr = (nG + 7) / 8;
            hv.clear();
            make_halves(0, ntr, hv);
            if ((int)hv.size() > cap) {
                // the CTAs of a split list all stage and evaluate the rows, so they should finish together:
                // the halves are dealt heaviest first, back and forth, to equally many per CTA
                const int ng = ((int)hv.size() + cap - 1) / cap;
                std::vector<Half> sorted = hv;
                std::stable_sort(sorted.begin(), sorted.end(), [](const Half &x, const Half &y) { return x.load > y.load; });
                std::vector<Group> parts(ng);
                for (size_t b = 0; b < sorted.size(); ++b) {
                    const int lap = (int)(b / ng), pos = (int)(b % ng);
                    Group &g = parts[(lap & 1) ? ng - 1 - pos : pos];
                    g.halves.push_back(sorted[b]);
                    g.tiles += sorted[b].load;
                }
                for (int q = 0; q < ng; ++q) {
                    Group &g = parts[q];
                    g.prob = p; g.nfit = 1; g.fit[0] = f; g.lead[0] = (q == 0);
                    deal_halves(g, q == 0 ? hc_single : 0);
                    groups.push_back(g);
                }
                continue;
            }
            bool placed = false;
            for (Group &g : open)
                if (g.nfit < fmax && (int)g.halves.size() + (int)hv.size() <= cap) {
                    const int sl = g.nfit++;
                    g.fit[sl] = f; g.lead[sl] = 1;
                    for (Half h : hv) {
                        h.fit = sl;
                        g.halves.push_back(h);
                        g.tiles += h.load;
                    }
                    placed = true;
                    break;
                }
            if (!placed) {
                Group g;
                g.prob = p; g.nfit = 1; g.fit[0] = f; g.lead[0] = 1;
                for (Half h : hv) {
                    g.halves.push_back(h);
                    g.tiles += h.load;
                }
                open.push_back(g);
            }
        }
        for (Group &g : open) {
            deal_halves(g, g.nfit > 1 && fmax > 1 ? hc_multi : hc_single);
            groups.push_back(g);
        }
    }
    return 0;
}

// descriptor of a half: fit slot | segment << 2 | count << 8 | row0 << 10 | row1 << 18
unsigned half_word(const Half &h) {
    return (unsigned)h.fit | ((unsigned)h.seg << 2) | ((unsigned)h.cnt << 8) | ((unsigned)h.rows[0] << 10) | ((unsigned)h.rows[1] << 18);
}

template <int ROWS, int ITER, int FMAX>
int launch_round(ate_ctx *ctx, const RArgs &a, int nctas, size_t smem, bool init) {
    if (init) {
        auto k = irls_round_kernel<ROWS, ITER, FMAX, true>;
        if (cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return -1;
        k<<<nctas, R_THREADS, smem, ctx->stream>>>(a);
    } else {
        auto k = irls_round_kernel<ROWS, ITER, FMAX, false>;
        if (cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return -1;
        k<<<nctas, R_THREADS, smem, ctx->stream>>>(a);
    }
    return 0;
}
// narrow rows: a fit's first step runs one fit per CTA on the FMA-pipe helpers, every other step four
template <int ROWS, int ITER>
int launch_narrow(ate_ctx *ctx, const RArgs &a, int nctas, size_t smem, bool init) {
    if (init) {
        auto k = irls_round_kernel<ROWS, ITER, 1, true>;
        if (cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return -1;
        k<<<nctas, R_THREADS, smem, ctx->stream>>>(a);
    } else {
        auto k = irls_round_kernel<ROWS, ITER, R_FMAX, false>;
        if (cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return -1;
        k<<<nctas, R_THREADS, smem, ctx->stream>>>(a);
    }
    return 0;
}

}  // namespace

static int round_fmax(int ncs, int init);
// host-only introspection of the planner (include/ate_b200.h)
extern "C" int ate_round_plan_describe(int nfit, const int *list_cols, int staged_cols, int max_groups, unsigned *desc) {
    if (nfit < 1 || !list_cols || !desc || max_groups < 1 || staged_cols < 8) return -1;
    const int fmax = round_fmax(staged_cols, 0);
    std::vector<int> prob_of(nfit, 0), live(nfit, 1), nGs(list_cols, list_cols + nfit);
    std::vector<Group> groups;
    if (build_groups(1, prob_of, live, nGs, fmax, staged_cols, groups) < 0) return -1;
    if ((int)groups.size() > max_groups) return -2;
    for (size_t gi = 0; gi < groups.size(); ++gi) {
        unsigned *d = desc + 24 * gi;
        const Group &g = groups[gi];
        for (int k = 0; k < 24; ++k) d[k] = 0u;
        d[0] = (unsigned)g.nfit;
        for (int sl = 0; sl < g.nfit; ++sl) {
            d[1 + sl] = (unsigned)g.fit[sl];
            d[5] |= (unsigned)(g.lead[sl] ? 1 : 0) << sl;
        }
        for (int w = 0; w < R_CWARPS; ++w) {
            d[8 + 2 * w] = half_word(g.warp_half[w][0]);
            d[9 + 2 * w] = half_word(g.warp_half[w][1]);
        }
    }
    return (int)groups.size();
}

// slot geometry for a staged width: 32-row slots while three of them fit, else 16-row slots
// fixed shared memory beside the ring: the fits' coefficients, the list of non-zero column steps, slack
static size_t round_fixed_smem(int ncs, int fmax) {
    return (sizeof(double) + sizeof(uint16_t)) * (size_t)fmax * ncs + 64 + 2048;  // non-zero lists; barriers and slack
}
static int round_rows_for(int ncs, int fmax) {
    static const int forced = getenv("ATE_ROUND_ROWS") ? atoi(getenv("ATE_ROUND_ROWS")) : 0;  // A/B measurements
    if ((forced == 16 || forced == 32) && (ncs + 63) / 64 == 5) return forced;
    const size_t b32 = sizeof(double) * ((size_t)32 * (ncs + R_PAD) + 2 * (size_t)fmax * 32);
    return (R_SMEM_LIMIT - round_fixed_smem(ncs, fmax)) / b32 >= 3 ? 32 : 16;
}
// fits one CTA serves: four where the helpers run on the DMMA pipe (rows up to 320 staged columns, any
// step but a fit's first); one on first steps and for wider rows (FMA-pipe helpers)
static int round_fmax(int ncs, int init) { return ((ncs + 63) / 64 <= 5 && !init) ? R_FMAX : 1; }

int ate_i_irls_round(ate_ctx *ctx, int ncs, int ni, int iy, int init, int bank, const std::vector<RoundProblem> &probs,
                     const std::vector<RoundFit> &fits, std::vector<RoundSlot> &slots, double **packed,
                     long long *packed_count, RoundStats *stats) {
    // two launches of one round (fits on their first step beside fits under way) must not share result,
    // partial or descriptor buffers: every cached buffer exists once per bank
    const char *const nm_packed = bank ? "round.packed.1" : "round.packed", *const nm_part = bank ? "round.partials.1" : "round.partials",
               *const nm_hdesc = bank ? "round.hdesc.1" : "round.hdesc", *const nm_ddesc = bank ? "round.ddesc.1" : "round.ddesc";
    const int nfits = (int)fits.size(), P = (int)probs.size();
    if (ncs % 8 != 0 || ncs <= ni - 1 || iy >= ncs || ni < 1) return ate_fail(ctx, -1, "irls_round: bad staged width %d", ncs);
    const int it_e = (ncs + 63) / 64;
    const int fmax = round_fmax(ncs, init);
    // the slot geometry follows the widest variant (four fits) so that a fit's first step uses the same one
    const int ROWS = round_rows_for(ncs, it_e <= 5 ? R_FMAX : 1);
    const size_t stage_bytes = sizeof(double) * ((size_t)ROWS * (ncs + R_PAD) + 2 * (size_t)fmax * ROWS);
    int stages = (int)((R_SMEM_LIMIT - round_fixed_smem(ncs, fmax)) / stage_bytes);
    if (stages > R_MAX_STAGES) stages = R_MAX_STAGES;
    if (stages < 2 || it_e > 13) return ate_fail(ctx, -1, "irls_round: %d staged columns do not fit the step kernel", ncs);
    const size_t smem = stages * stage_bytes + round_fixed_smem(ncs, fmax) - 1024;
    for (auto &pr : probs)
        if ((pr.ld & 1) != 0 || pr.ld < ncs) return ate_fail(ctx, -1, "irls_round: bad leading dimension");

    // ---- groups: which fits share a CTA, which tiles each MMA warp owns ----
    std::vector<Group> groups;
    {
        std::vector<int> prob_of(nfits), live(nfits), nGs(nfits);
        for (int f = 0; f < nfits; ++f) {
            prob_of[f] = fits[f].prob;
            live[f] = fits[f].live;
            nGs[f] = fits[f].nG;
        }
        if (build_groups(P, prob_of, live, nGs, fmax, ncs, groups) < 0)
            return ate_fail(ctx, -1, "irls_round: a column list exceeds the tile plan");
    }
    if (groups.empty()) {
        *packed_count = 0;
        return 0;
    }
    // ---- row chunks: every CTA kind of a problem cuts the problem's rows the same way (a fit's partials are
    //      indexed by chunk).  About twelve waves of CTAs are laid out, the work of a problem deciding its
    //      share, and handed out heaviest kind first: CTA kinds differ in cost by up to 2x (fits per CTA,
    //      tiles per fit), so with few waves the tail of the launch idles most SMs (measured at C1:
    //      3 waves 1690 ms per fit, 6 waves 1514, 12 waves 1483, heaviest first 1465; a CTA still runs
    //      for hundreds of slots, its prologue is noise).
    std::vector<double> pcost(P, 0.0);
    std::vector<int> pkinds(P, 0), pchunks(P, 1);
    double total_cost = 0.0;
    for (Group &g : groups) {
        int wmax = 0;
        for (int w = 0; w < R_CWARPS; ++w) wmax = std::max(wmax, g.warp_half[w][0].load + g.warp_half[w][1].load);
        g.cost = (double)probs[g.prob].nrows * (30.0 + 25.0 * g.nfit + 4.0 * wmax + 0.5 * g.tiles);
        pcost[g.prob] += g.cost;
        pkinds[g.prob]++;
        total_cost += g.cost;
    }
    {
        static const int waves = getenv("ATE_ROUND_WAVES") ? std::max(1, atoi(getenv("ATE_ROUND_WAVES"))) : 12;
        const double budget = (double)waves * ctx->num_sms;
        for (int p = 0; p < P; ++p) {
            if (!pkinds[p]) continue;
            long long maxc = (probs[p].nrows + 16LL * ROWS - 1) / (16LL * ROWS);  // at least 16 slots per CTA
            if (maxc < 1) maxc = 1;
            long long c = (long long)(pcost[p] / total_cost * budget / pkinds[p] + 0.5);
            if (c < 1) c = 1;
            if (c > maxc) c = maxc;
            pchunks[p] = (int)c;
        }
    }
    for (Group &g : groups) g.chunks = pchunks[g.prob];

    // ---- per-fit outputs: partial slots and the packed result buffer ----
    slots.assign(nfits, RoundSlot{0, 0, 0});
    std::vector<int> fit_chunks(nfits, 0);
    for (Group &g : groups)
        for (int sl = 0; sl < g.nfit; ++sl)
            if (g.lead[sl]) fit_chunks[g.fit[sl]] = g.chunks;
    const int plen = 3 * ni + 3, eplen = 3 * ni + 3 * R_EWARPS;
    long long pk = 0;
    size_t part_doubles = 0;
    std::vector<size_t> gram_off(nfits, 0), eval_off(nfits, 0);
    int nlive = 0, ldh_max = 0;
    for (int f = 0; f < nfits; ++f) {
        if (!fits[f].live) continue;
        ++nlive;
        const int ldh = (fits[f].nG + 31) / 32 * 32;
        ldh_max = std::max(ldh_max, ldh);
        slots[f].sums_off = pk;
        pk += (plen + 3) / 4 * 4;
        slots[f].H_off = pk;
        slots[f].ldh = ldh;
        pk += (long long)ldh * ldh;
        const int nb = ldh / 32, nblk = nb * (nb + 1) / 2;
        gram_off[f] = part_doubles;
        part_doubles += (size_t)1024 * nblk * fit_chunks[f];
        eval_off[f] = part_doubles;
        part_doubles += ((size_t)eplen * fit_chunks[f] + 3) / 4 * 4;  // the next fit's tiles are stored 16 bytes at a time
    }
    void *dpacked = nullptr, *dpart = nullptr;
    ATE_TRY(ate_cached(ctx, nm_packed, sizeof(double) * (size_t)std::max(pk, 1LL), &dpacked));
    ATE_TRY(ate_cached(ctx, nm_part, sizeof(double) * std::max(part_doubles, (size_t)1), &dpart));
    *packed = (double *)dpacked;
    *packed_count = pk;

    // ---- descriptors: one pinned staging block, one copy ----
    int nctas = 0;
    for (Group &g : groups) nctas += g.chunks;
    const size_t b_probs = sizeof(DProb) * P, b_fits = sizeof(DFit) * nfits, b_ctas = sizeof(DCta) * nctas,
                 b_folds = sizeof(DFold) * nlive;
    auto up16 = [](size_t v) { return (v + 15) / 16 * 16; };
    const size_t o_fits = up16(b_probs), o_ctas = o_fits + up16(b_fits), o_folds = o_ctas + up16(b_ctas),
                 total = o_folds + up16(b_folds);
    void *hst = nullptr, *ddesc = nullptr;
    ATE_TRY(ate_cached_pinned(ctx, nm_hdesc, total, &hst));
    ATE_TRY(ate_cached(ctx, nm_ddesc, total, &ddesc));
    memset(hst, 0, total);
    DProb *hp = (DProb *)hst;
    DFit *hf = (DFit *)((char *)hst + o_fits);
    DCta *hc = (DCta *)((char *)hst + o_ctas);
    DFold *hd = (DFold *)((char *)hst + o_folds);
    for (int p = 0; p < P; ++p) hp[p] = DProb{probs[p].X, probs[p].ld, probs[p].rows, probs[p].nrows, probs[p].padded, 0};
    int li = 0;
    for (int f = 0; f < nfits; ++f) {
        if (!fits[f].live) continue;
        const int nb = slots[f].ldh / 32, nblk = nb * (nb + 1) / 2;
        hf[f].coef = fits[f].coef;
        hf[f].colidx = fits[f].colidx;
        hf[f].hole_begin = fits[f].hole_len > 0 ? fits[f].hole_begin : 0;
        hf[f].hole_len = fits[f].hole_len > 0 ? fits[f].hole_len : 0;
        hf[f].w = fits[f].w;
        hf[f].gram_part = (double *)dpart + gram_off[f];
        hf[f].eval_part = (double *)dpart + eval_off[f];
        hf[f].nblk = nblk;
        hd[li].gram_part = hf[f].gram_part;
        hd[li].eval_part = hf[f].eval_part;
        hd[li].sums = (double *)dpacked + slots[f].sums_off;
        hd[li].H = (double *)dpacked + slots[f].H_off;
        hd[li].nchunks = fit_chunks[f];
        hd[li].nblk = nblk;
        hd[li].nG = fits[f].nG;
        hd[li].ldh = slots[f].ldh;
        hd[li].colidx = fits[f].colidx;
        hd[li].ones = init ? -1 : fits[f].ones;
        hd[li].ni = ni;
        hf[f].ones = init ? -1 : fits[f].ones;
        ++li;
    }
    RoundStats st;
    int ci = 0;
    std::vector<double> cta_cost;
    for (int p = 0; p < P; ++p) {
        if (!pkinds[p]) continue;
        const long long nrows = probs[p].nrows;
        long long rpc = (nrows + pchunks[p] - 1) / pchunks[p];
        rpc = (rpc + ROWS - 1) / ROWS * ROWS;
        for (int c = 0; c < pchunks[p]; ++c)          // chunk-major: the kinds of a chunk sit side by side
            for (Group &g : groups) {
                if (g.prob != p) continue;
                DCta &d = hc[ci++];
                d.prob = g.prob;
                d.nfit = g.nfit;
                d.chunk = c;
                d.row_begin = std::min(nrows, (long long)c * rpc);
                d.row_end = std::min(nrows, (long long)(c + 1) * rpc);
                for (int sl = 0; sl < g.nfit; ++sl) {
                    d.fit[sl] = g.fit[sl];
                    d.lead[sl] = g.lead[sl];
                }
                for (int wq = 0; wq < R_CWARPS; ++wq)
                    d.mma[wq] = make_uint2(half_word(g.warp_half[wq][0]), half_word(g.warp_half[wq][1]));
                st.rows_staged += d.row_end - d.row_begin;
                cta_cost.push_back(g.cost / std::max(1, g.chunks));
            }
    }
    {
        // The hardware hands out CTAs in index order.  Heaviest kinds first (longest-processing-time order):
        // the light CTAs then fill the tail of the launch.  ATE_ROUND_ORDER=0 keeps the chunk-major order
        // (the kinds of a chunk side by side, walking the same rows at the same time).
        static const int order = getenv("ATE_ROUND_ORDER") ? atoi(getenv("ATE_ROUND_ORDER")) : 1;
        if (order == 1) {
            std::vector<int> perm(nctas);
            for (int i = 0; i < nctas; ++i) perm[i] = i;
            std::stable_sort(perm.begin(), perm.end(), [&](int x, int y) { return cta_cost[x] > cta_cost[y]; });
            std::vector<DCta> tmp(hc, hc + nctas);
            for (int i = 0; i < nctas; ++i) hc[i] = tmp[perm[i]];
        }
    }
    for (Group &g : groups) {
        const long long nrows = probs[g.prob].nrows;
        for (int sl = 0; sl < g.nfit; ++sl)
            if (g.lead[sl]) {
                const double nG = fits[g.fit[sl]].nG;
                st.row_fits += nrows;
                st.flops += (double)nrows * (nG * (nG + 1.0) + 6.0 * ni);
                st.alg_bytes += (double)nrows * ((double)ni * 8.0 + 24.0);  // n (p 8 + 24) per fit and step
            }
    }
    st.ctas = nctas;
    if (getenv("ATE_ROUND_DEBUG")) {  // the layout of a few rounds, for tuning the cost model
        static int calls = 0;
        if (calls < 4 || calls % 50 == 0) {
            fprintf(stderr, "[ate round %d] %d live fits, %zu CTA kinds, %d CTAs, ROWS %d stages %d fmax %d init %d\n", calls, nlive,
                    groups.size(), nctas, ROWS, stages, fmax, init);
            for (size_t gi = 0; gi < groups.size() && gi < 12; ++gi) {
                const Group &g = groups[gi];
                int wmax = 0;
                for (int w = 0; w < R_CWARPS; ++w) wmax = std::max(wmax, g.warp_half[w][0].load + g.warp_half[w][1].load);
                fprintf(stderr, "   kind %zu: prob %d fits %d (nG %d, %d) lead %d halves %zu tiles %d busiest %d chunks %d rows/chunk %lld\n", gi,
                        g.prob, g.nfit, fits[g.fit[0]].nG, g.nfit > 1 ? fits[g.fit[1]].nG : -1, g.lead[0], g.halves.size(), g.tiles, wmax,
                        g.chunks, (long long)((probs[g.prob].nrows + g.chunks - 1) / g.chunks));
            }
        }
        ++calls;
    }
    ATE_CUDA(ctx, cudaMemcpyAsync(ddesc, hst, total, cudaMemcpyHostToDevice, ctx->stream));
    RArgs a;
    a.probs = (const DProb *)ddesc;
    a.fits = (const DFit *)((char *)ddesc + o_fits);
    a.ctas = (const DCta *)((char *)ddesc + o_ctas);
    a.ncols = ncs; a.ni = ni; a.iy = iy; a.stages = stages;
    a.skip = ctx->debug_skip;
    {
        static const int hint = getenv("ATE_WAIT_HINT_NS") ? atoi(getenv("ATE_WAIT_HINT_NS")) : 0;
        static const int rolemap = getenv("ATE_ROUND_ROLEMAP") ? atoi(getenv("ATE_ROUND_ROLEMAP")) : 0;
        a.wait_hint = hint;
        a.rolemap = rolemap;
    }
    const DFold *dfolds = (const DFold *)((char *)ddesc + o_folds);

    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ctx->profile & 1) {
        ATE_CUDA(ctx, cudaEventCreate(&e0));
        ATE_CUDA(ctx, cudaEventCreate(&e1));
        ATE_CUDA(ctx, cudaEventRecord(e0, ctx->stream));
    }
    int rc = -1;
    const bool in = init != 0;
    if (it_e <= 5) {  // rows up to 320 staged columns: four fits per CTA (one on a fit's first step)
        if (ROWS == 32) {
            if (it_e <= 1) rc = launch_narrow<32, 1>(ctx, a, nctas, smem, in);
            else if (it_e <= 2) rc = launch_narrow<32, 2>(ctx, a, nctas, smem, in);
            else if (it_e <= 3) rc = launch_narrow<32, 3>(ctx, a, nctas, smem, in);
            else if (it_e <= 4) rc = launch_narrow<32, 4>(ctx, a, nctas, smem, in);
            else rc = launch_narrow<32, 5>(ctx, a, nctas, smem, in);
        } else {
            rc = launch_narrow<16, 5>(ctx, a, nctas, smem, in);
        }
    } else if (it_e <= 7) rc = launch_round<16, 7, 1>(ctx, a, nctas, smem, in);
    else if (it_e <= 9) rc = launch_round<16, 9, 1>(ctx, a, nctas, smem, in);
    else rc = launch_round<16, 13, 1>(ctx, a, nctas, smem, in);
    if (rc < 0) return ate_fail(ctx, -1, "irls_round: cannot launch for %d staged columns", ncs);
    ATE_LAUNCHED(ctx);
    if (ctx->profile & 1) {
        ATE_CUDA(ctx, cudaEventRecord(e1, ctx->stream));
        ctx->prof_events.emplace_back(e0, e1);
        ctx->prof_shape.emplace_back(nlive, st.row_fits);
        ctx->prof_launches++;
        ctx->prof_flops += st.flops;
        ctx->prof_bytes += st.alg_bytes;
    }
    round_fold_eval_kernel<<<dim3((plen + 7) / 8, nlive), 256, 0, ctx->stream>>>(dfolds, ni, init);
    ATE_LAUNCHED(ctx);
    if (ldh_max > 0) {
        round_fold_gram_kernel<<<dim3((ldh_max * ldh_max + 255) / 256, nlive), 256, 0, ctx->stream>>>(dfolds);
        ATE_LAUNCHED(ctx);
    }
    if (stats) *stats = st;
    return 0;
}

// Measurement entry (include/ate_b200.h): `iters` launches of the step kernel on synthetic rows.
extern "C" int ate_round_kernel_bench(ate_ctx *ctx, long long nrows, int ni, int nprob, int fits_per_prob, int list_cols,
                                      int private_copy, int iters, int skip_mask, double *ms_per_launch, double *flops,
                                      double *alg_bytes, int *ctas) {
    if (!ctx) return -1;
    if (nrows < 64 || ni < 1 || ni > 1024 || nprob < 1 || fits_per_prob < 1 || list_cols < 0 || list_cols > ni || iters < 1)
        return ate_fail(ctx, -1, "ate_round_kernel_bench: bad arguments");
    ATE_CUDA(ctx, cudaSetDevice(ctx->device));
    const int iy = ni, ncs = (ni + 1 + 7) / 8 * 8;
    const long long ld = private_copy ? ncs + 4 : (ni + 3 + 31) / 32 * 32;
    const int nfits = nprob * fits_per_prob, ldh = (ni + 31) / 32 * 32;
    void *blk = nullptr;
    const size_t xb = sizeof(double) * (size_t)nrows * (size_t)ld * nprob;
    const size_t cb = sizeof(double) * (size_t)(ni + 1) * nfits, lb = sizeof(int) * (size_t)ldh * nfits;
    ATE_TRY(ate_cached(ctx, "roundbench.state", xb + cb + lb + 1024, &blk));
    double *dX = (double *)blk;
    double *dcoef = (double *)((char *)blk + xb);
    int *dlist = (int *)((char *)blk + xb + cb);
    {   // rows: small deterministic values, response column 0/1, pad zero
        std::vector<double> row((size_t)ld * 4096);
        for (long long r0 = 0; r0 < nrows * nprob; r0 += 4096) {
            const long long cnt = std::min<long long>(4096, nrows * nprob - r0);
            for (long long r = 0; r < cnt; ++r)
                for (long long c = 0; c < ld; ++c) {
                    const unsigned h = (unsigned)((r0 + r) * 2654435761u + c * 40503u);
                    row[r * ld + c] = c < ni ? ((h >> 8) & 1023) / 512.0 - 1.0 : (c == iy ? (double)((h >> 5) & 1) : ((c == ni + 2 && !private_copy) || (c == ncs + 1 && private_copy) ? 1.0 : 0.0));
                }
            ATE_CUDA(ctx, cudaMemcpyAsync(dX + (size_t)r0 * ld, row.data(), sizeof(double) * (size_t)cnt * ld, cudaMemcpyHostToDevice, ctx->stream));
            ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        }
        std::vector<double> hc((size_t)(ni + 1) * nfits, 0.0);
        std::vector<int> hl((size_t)ldh * nfits, ncs);
        for (int f = 0; f < nfits; ++f) {
            for (int j = 0; j < list_cols; ++j) {
                hl[(size_t)f * ldh + j] = (j * 7 + f) % ni;  // scattered columns, as a strong set is
                hc[(size_t)f * (ni + 1) + hl[(size_t)f * ldh + j]] = 0.01 * ((j % 5) - 2);
            }
            hc[(size_t)f * (ni + 1) + ni] = -0.3;
            if (list_cols >= R_ONES_MIN) hl[(size_t)f * ldh + list_cols - 1] = ncs + 1;  // as the driver does: the ONE column closes a wide list
        }
        ATE_CUDA(ctx, cudaMemcpyAsync(dcoef, hc.data(), cb, cudaMemcpyHostToDevice, ctx->stream));
        ATE_CUDA(ctx, cudaMemcpyAsync(dlist, hl.data(), lb, cudaMemcpyHostToDevice, ctx->stream));
        ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    std::vector<RoundProblem> probs(nprob);
    std::vector<RoundFit> fits(nfits);
    for (int p = 0; p < nprob; ++p) probs[p] = RoundProblem{dX + (size_t)p * nrows * ld, ld, nullptr, nrows, private_copy ? 1 : 0};
    for (int f = 0; f < nfits; ++f) {
        RoundFit &r = fits[f];
        r.prob = f / fits_per_prob;
        const int k = f % fits_per_prob;
        r.hole_begin = fits_per_prob > 1 ? nrows / fits_per_prob * k : 0;
        r.hole_len = fits_per_prob > 1 ? nrows / fits_per_prob : 0;
        r.w = 1.0 / (double)nrows;
        r.coef = dcoef + (size_t)f * (ni + 1);
        r.colidx = dlist + (size_t)f * ldh;
        r.nG = list_cols;
        r.ones = list_cols >= R_ONES_MIN ? list_cols - 1 : -1;
        r.live = 1;
    }
    std::vector<RoundSlot> slots;
    double *packed = nullptr;
    long long count = 0;
    RoundStats st;
    ctx->debug_skip = skip_mask;
    int rc = ate_i_irls_round(ctx, ncs, ni, iy, 0, 0, probs, fits, slots, &packed, &count, &st);  // warm-up
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    float ms = 0.f;
    if (rc == 0 && cudaEventCreate(&e0) == cudaSuccess && cudaEventCreate(&e1) == cudaSuccess) {
        cudaStreamSynchronize(ctx->stream);
        cudaEventRecord(e0, ctx->stream);
        for (int it = 0; it < iters && rc == 0; ++it) rc = ate_i_irls_round(ctx, ncs, ni, iy, 0, 0, probs, fits, slots, &packed, &count, &st);
        cudaEventRecord(e1, ctx->stream);
        if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) rc = ate_fail(ctx, -100, "ate_round_kernel_bench: %s", cudaGetErrorString(cudaGetLastError()));
        cudaEventElapsedTime(&ms, e0, e1);
    }
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    ctx->debug_skip = 0;
#ifdef ATE_ROUND_TRACE
    if (const char *path = getenv("ATE_ROUND_TRACE_OUT")) {
        static long long hbuf[TR_SLOTS * TR_EVENTS + 8];
        if (cudaMemcpyFromSymbol(hbuf, g_rtrace, sizeof(hbuf)) == cudaSuccess) {
            if (FILE *f = fopen(path, "w")) {
                fprintf(f, "slot,prod_free,e0_start,e1_start,e0_pub,e1_pub,h0_start,h1_start,h0_done,h1_done,m0_start,m0_done,m1_start,m1_done,m2_start,m2_done\n");
                long long base = hbuf[8];
                for (int it = 0; it < TR_SLOTS; ++it) {
                    fprintf(f, "%d", it);
                    for (int e = 0; e < 15; ++e) fprintf(f, ",%lld", hbuf[8 + it * TR_EVENTS + e] ? hbuf[8 + it * TR_EVENTS + e] - base : -1);
                    fprintf(f, "\n");
                }
                fclose(f);
            }
        }
    }
#endif
    if (rc < 0) return rc;
    if (ms_per_launch) *ms_per_launch = ms / iters;
    if (flops) *flops = st.flops;
    if (alg_bytes) *alg_bytes = st.alg_bytes;
    if (ctas) *ctas = st.ctas;
    return 0;
}
