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
// CTA = 8 MMA warps + 2 producer warps + 2 helper warps + 4 evaluation warps, one CTA per SM, around
// a ring of 32-row slots in shared memory (three at the C1 width of 264 staged columns).  No pipe
// bounds this kernel (ncu: shared-memory pipe 50-66 %, DMMA 34-65 %, issue slots 37 %): the per-slot
// trace (-DATE_STEP_TRACE, tools/step_trace.py) shows the serial life of a slot -- fill, evaluate,
// consume -- pacing the ring, at ~2500 cycles per 16 rows where HBM speed would be 1500.  16-row
// slots with six in the ring measure the same; see DESIGN.md section 5 for what was tried.
//   producers: producer pw stages rows [16 pw, 16 pw + 16) of every slot with one-dimensional TMA
//              bulk copies (one per lane; ONE copy for the whole slot when the rows are contiguous
//              at the slot's pitch) into an mbarrier ring, plus a validity flag per row.
//   eval     : the working-response refresh is a long dependent chain (dot -> shuffle reduce -> exp
//              -> divide -> log), latency-bound per warp, so a slot is refreshed by two warps (16
//              rows each) and consecutive slots by different pairs.  Lanes hold a column slice of
//              beta; the 16 partial dots are reduce-scattered across the warp (5 exchange steps),
//              so lane l ends up with row l/2 and the exp/log work of 16 rows runs in parallel
//              lanes.  v and r go to shared memory (second mbarrier) for the MMA and helper warps.
//   helpers  : (lead CTA of a row chunk) accumulate X^T r, X^T v on the FP64 FMA pipe from the
//              same staged rows, 16 rows of every slot each -- off the eval warps' critical path and
//              off the tensor pipe.
//   MMA      : the lower triangle of H in 8x8 tiles.  A warp owns up to four tile ROWS of one
//              32-column segment (<= 16 tiles, 64 accumulator registers), dealt by a host-built plan
//              for every list width.  Operands are gathered from the staged rows through the
//              column list.
// Row chunks are folded in fixed order afterwards: bit-reproducible.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "irls_step.cuh"
#include "path.cuh"
#include "tma_dmma.cuh"

namespace {

using namespace atk;

#ifndef ATE_STEP_ROWS
#define ATE_STEP_ROWS 32
#endif
constexpr int F_ROWS = ATE_STEP_ROWS;  // rows per slot (16, 24 or 32)
constexpr int F_HALF = 16;             // an eval warp refreshes 16 rows at a time (its reduce-scatter is 16 wide)
constexpr int F_HALVES = (F_ROWS + F_HALF - 1) / F_HALF;  // the last one may be short (24-row slots: 16 + 8)
constexpr int F_CWARPS = 8;   // MMA warps;  8 + 2 + 2 + 4 = 16 warps: 128 registers per thread
constexpr int F_HWARPS = 2;   // helper warps (X^T r, X^T v); each owns an equal share of the rows of every slot
constexpr int F_EWARPS = 4;   // evaluation warps, taking turns on the slots (see F_TURNS)
constexpr int F_PWARPS = 2;   // producer warps; each stages an equal share of the rows of every slot.  One warp issues a
                              // 2 KB bulk copy every ~140 cycles (the copies of a warp's lanes are
                              // serialised on the uniform datapath): 4.1 TB/s chip-wide, measured --
                              // two warps lift the issue rate above what HBM delivers
constexpr int F_THREADS = (F_CWARPS + F_HWARPS + F_PWARPS + F_EWARPS) * 32;
// Warp w runs on SM sub-partition w % 4.  MMA warps 0-7 (the plan fills 0-3 first), producers 8-9
// (sub-partitions 0-1, next to the two busiest MMA warps: they hardly compute), helpers 10-11
// (sub-partitions 2-3), eval warps 12-15 (one per sub-partition).
constexpr int F_PRODUCER = F_CWARPS;            // warp index of the first producer
constexpr int F_HELPER = F_CWARPS + F_PWARPS;   // ... of the first helper
constexpr int F_EVAL = F_HELPER + F_HWARPS;     // ... of the first eval warp
constexpr int F_PAD = 4;
constexpr int F_MAX_STAGES = 8;
constexpr size_t F_SMEM_LIMIT = 227 * 1024;
constexpr int F_PSHARE = F_ROWS / F_PWARPS;  // rows of every slot staged by one producer (one per lane)
constexpr int F_HSHARE = F_ROWS / F_HWARPS;  // rows of every slot accumulated by one helper
constexpr int F_TURNS = F_EWARPS / F_HALVES; // eval warp e refreshes half e % F_HALVES of the slots with it % F_TURNS == e / F_HALVES
static_assert(F_HALF == 16 && (F_ROWS == 16 || F_ROWS == 24 || F_ROWS == 32) && F_EWARPS % F_HALVES == 0 && F_ROWS % F_PWARPS == 0 &&
                  F_ROWS % F_HWARPS == 0,
              "slot geometry");

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
    const uint2 *plan;  // this launch's tile-row plan: one descriptor per MMA warp of the grid row
    int nblk, stages;
    double *gram_partials, *eval_partials, *v_out;
    int trace;                // -DATE_STEP_TRACE builds: stamp this launch (value = list width)
    unsigned long long *dbg;  // ATE_STEP_WATCHDOG=1: host-mapped log of barrier waits that never end
};

// mbar_wait with a deadlock detector: when a log buffer is given, a wait that lasts longer than two
// seconds (a launch takes under a millisecond) writes who waits on what, and one second later traps.
__device__ __forceinline__ unsigned long long gtimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
__device__ __noinline__ void wait_report(unsigned long long *dbg, int tag, int it, int s, uint32_t parity,
                                        unsigned long long info) {
    if ((threadIdx.x & 31) == 0) {
        const unsigned long long k = atomicAdd(dbg, 1ull);
        if (k < 120) {
            dbg[4 + 4 * k] = (unsigned long long)tag | ((unsigned long long)(threadIdx.x >> 5) << 8) |
                             ((unsigned long long)blockIdx.x << 16) | ((unsigned long long)blockIdx.y << 32);
            dbg[5 + 4 * k] = (unsigned long long)(unsigned)it;
            dbg[6 + 4 * k] = (unsigned long long)(unsigned)s | ((unsigned long long)parity << 32);
            dbg[7 + 4 * k] = info;
        }
        __threadfence_system();
    }
}
__device__ __forceinline__ void wait_w(uint64_t *bar, uint32_t parity, const StepArgs &a, int tag, int it, int s) {
    unsigned long long *dbg = a.dbg;
    if (dbg == nullptr) {
        mbar_wait(bar, parity);
        return;
    }
    unsigned long long t0 = 0;
    bool reported = false;
    for (unsigned spins = 0;; ++spins) {
        uint32_t done;
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
        if (done) return;
        if ((spins & 255u) == 255u) {
            const unsigned long long t = gtimer_ns();
            if (t0 == 0) t0 = t;
            if (!reported && t - t0 > 2000000000ull) {
                wait_report(dbg, tag, it, s, parity,
                            (unsigned long long)(unsigned)a.ncols | ((unsigned long long)(unsigned)a.stages << 16) |
                                ((unsigned long long)(unsigned)a.nblk << 24) |
                                ((unsigned long long)(a.rows == nullptr && a.ld == (long long)(a.ncols + F_PAD)) << 40) |
                                ((unsigned long long)(a.rows_per_chunk / F_ROWS) << 41));
                reported = true;
            }
            if (t - t0 > 3000000000ull) __trap();
        }
    }
}

// -DATE_STEP_TRACE: CTA (0,0) of launches with a narrow list stamps clock64() per slot and role into
// a device buffer (tools/step_trace.py prints the timeline).  Compiled out otherwise.
#ifdef ATE_STEP_TRACE
constexpr int TR_SLOTS = 96, TR_EVENTS = 12;
__device__ long long g_trace[TR_SLOTS * TR_EVENTS + 8];
#define TRACE(ev, it)                                                                         \
    do {                                                                                      \
        if (trace_on && (it) < TR_SLOTS && (threadIdx.x & 31) == 0) g_trace[8 + (it) * TR_EVENTS + (ev)] = clock64(); \
    } while (0)
#else
#define TRACE(ev, it) do { } while (0)
#endif

template <int ITER, bool INIT>
__global__ void __launch_bounds__(F_THREADS, 1) irls_step_kernel(const StepArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int ncols = a.ncols, pitch = ncols + F_PAD, stages = a.stages;
    const size_t stage_doubles = (size_t)F_ROWS * pitch;
    double *sx = reinterpret_cast<double *>(smem_raw);
    double *sv = sx + stage_doubles * stages;   // [stages][F_ROWS]: validity from the producers, then v
    double *sy = sv + (size_t)stages * F_ROWS;  // [stages][F_ROWS]: r from the eval warps
    uint64_t *full = reinterpret_cast<uint64_t *>(sy + (size_t)stages * F_ROWS);
    uint64_t *vfull = full + stages, *empty = vfull + stages;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#ifdef ATE_STEP_TRACE
    const bool trace_on = blockIdx.x == 0 && blockIdx.y == 0 && a.trace != 0;
    if (trace_on && threadIdx.x == 0) {
        g_trace[0] = clock64(); g_trace[1] = a.ncols; g_trace[2] = a.stages; g_trace[3] = a.nblk;
        g_trace[4] = a.rows_per_chunk / F_ROWS; g_trace[5] = a.trace;
    }
#endif
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
            mbar_init(full + s, F_PWARPS);   // both producers stage half of every slot
            mbar_init(vfull + s, F_HALVES);  // the two eval warps of the slot
            mbar_init(empty + s, F_CWARPS + F_HWARPS + F_EWARPS);  // EVERY consumer warp, every slot
        }
        mbar_fence_init();
    }
    __syncthreads();
    // Ring protocol.  Every warp of a role visits every slot in order and arrives on `empty` for
    // every slot, whether or not it has work in it, and both producers wait on `empty` for every
    // slot.  A parity wait is only sound while the waiter is less than one phase away from the
    // barrier; with this rule no barrier can complete a phase that one of its waiters has not
    // observed the previous phase of, for any stage count.

    if (warp >= F_PRODUCER && warp < F_PRODUCER + F_PWARPS) {
        // ------------------------------ producers ------------------------------
        // Producer pw stages rows [F_PSHARE pw, F_PSHARE (pw + 1)) of every slot, one row per lane.
        const int pw = warp - F_PRODUCER;
        const int myrow = F_PSHARE * pw + lane;  // meaningful for lane < F_PSHARE
        if (a.rows == nullptr && a.ld == (long long)pitch) {
            // Contiguous rows whose pitch in HBM equals the slot pitch: the whole slot is ONE bulk
            // copy (two when it straddles the held-out hole), issued by producer 0.  The TMA unit
            // serves small copies at ~140 cycles each, so per-row copies cap at ~4.2 TB/s chip-wide;
            // slot-sized ones do not.
            for (int it = 0; it < nit; ++it) {
                const int s = it % stages;
                const uint32_t ph = (uint32_t)(it / stages) & 1u;
                const long long t0 = c_begin + (long long)it * F_ROWS;
                const int cnt = (int)((c_end - t0) < F_ROWS ? (c_end - t0) : F_ROWS);
                wait_w(empty + s, ph ^ 1u, a, 1, it, s);
                if (pw == 0) TRACE(0, it);  // producer: slot free
                double *dst0 = sx + stage_doubles * s;
                if (lane < F_PSHARE) {
                    sv[s * F_ROWS + myrow] = myrow < cnt ? 1.0 : 0.0;
                    if (myrow >= cnt)
                        for (int c = 0; c < ncols; ++c) dst0[(size_t)myrow * pitch + c] = 0.0;
                }
                __syncwarp();
                if (lane == 0) {
                    if (pw == 0) {
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
                    } else {
                        mbar_arrive(full + s);  // this producer's validity flags are in place
                    }
                }
                __syncwarp();
            }
        } else {
            // Row ids are fetched F_PF slots ahead (a register ring), so the global-load latency of
            // the list never sits between "slot free" and "copy issued".
            constexpr int F_PF = 4;
            auto fetch = [&](int it, long long &r) -> bool {
                const long long t = c_begin + (long long)it * F_ROWS + myrow;
                if (lane >= F_PSHARE || it >= nit || t >= c_end) return false;
                const long long tp = t + (t >= a.hole_begin ? a.hole_len : 0);
                r = a.rows ? (long long)__ldg(a.rows + tp) : tp;
                return true;
            };
            long long ring_r[F_PF];
            bool ring_ok[F_PF];
#pragma unroll
            for (int d = 0; d < F_PF; ++d) {
                ring_r[d] = 0;
                ring_ok[d] = fetch(d, ring_r[d]);
            }
            for (int it0 = 0; it0 < nit; it0 += F_PF) {
#pragma unroll
                for (int d = 0; d < F_PF; ++d) {
                    const int it = it0 + d;
                    if (it >= nit) break;
                    const long long cur_r = ring_r[d];
                    const bool cur_ok = ring_ok[d];
                    ring_ok[d] = fetch(it + F_PF, ring_r[d]);
                    const int s = it % stages;
                    const uint32_t ph = (uint32_t)(it / stages) & 1u;
                    wait_w(empty + s, ph ^ 1u, a, 2, it, s);
                    if (pw == 0) TRACE(0, it);
                    double *dst = sx + stage_doubles * s + (size_t)myrow * pitch;
                    if (lane < F_PSHARE) {
                        sv[s * F_ROWS + myrow] = cur_ok ? 1.0 : 0.0;
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
    } else if (warp >= F_EVAL) {
        // ------------------------------ evaluation ------------------------------
        // Eval warp e refreshes rows [16 half, 16 half + 16) of the slots whose turn it is.
        const int e = warp - F_EVAL;
        const int half = e % F_HALVES, turn = e / F_HALVES;
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
        const int rr = F_HALF * half + (lane >> 1);  // the row of the slot this lane finishes
        for (int it = 0; it < nit; ++it) {
            const int s = it % stages;
            const uint32_t ph = (uint32_t)(it / stages) & 1u;
            wait_w(full + s, ph, a, 3, it, s);
            if ((it % F_TURNS) != turn) {  // not this warp's slot: release it at once
                if (lane == 0) mbar_arrive(empty + s);
                continue;
            }
            if (half == 0) TRACE(1, it);  // eval: rows landed
            const double *base = sx + stage_doubles * s + (size_t)(F_HALF * half) * pitch;
            double part[F_HALF];
#pragma unroll
            for (int q = 0; q < F_HALF; ++q) {
                // a short last half re-reads its last row for the missing ones; those lanes are masked below
                const int qq = (F_ROWS % F_HALF != 0 && F_HALF * half + q >= F_ROWS) ? F_ROWS - 1 - F_HALF * half : q;
                const double *rowp = base + (size_t)qq * pitch;
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
            for (int hf = 8, off = 16; hf >= 1; hf >>= 1, off >>= 1) {
                const bool up = (lane & off) != 0;
#pragma unroll
                for (int k = 0; k < hf; ++k) {
                    const double send = up ? part[k] : part[k + hf], keep = up ? part[k + hf] : part[k];
                    part[k] = keep + __shfl_xor_sync(0xffffffffu, send, off);
                }
            }
            part[0] += __shfl_xor_sync(0xffffffffu, part[0], 1);
            const bool inslot = rr < F_ROWS;
            const int rc = inslot ? rr : F_ROWS - 1;
            const double valid = inslot ? sv[s * F_ROWS + rc] : 0.0, yi = sx[stage_doubles * s + (size_t)rc * pitch + a.iy];
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
            if ((lane & 1) == 0 && inslot) {
                sv[s * F_ROWS + rr] = vi;
                sy[s * F_ROWS + rr] = ri;
            }
            __syncwarp();
            if (half == 0) TRACE(2, it);  // eval: v, r ready
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
    } else if (warp >= F_HELPER) {
        // ------------------------------ helpers: X^T r, X^T v (and sum v x^2 on the first step) -----
        // Helper h owns rows [F_HSHARE h, F_HSHARE (h + 1)) of every slot.
        const int h = warp - F_HELPER;
        const bool lead = (blockIdx.x == 0);
        const int ni = a.ni;
        double gr[2 * ITER], hx[2 * ITER], hxx[INIT ? 2 * ITER : 1];
#pragma unroll
        for (int i = 0; i < ITER; ++i) {
            gr[2 * i] = gr[2 * i + 1] = hx[2 * i] = hx[2 * i + 1] = 0.0;
            if (INIT) hxx[2 * i] = hxx[2 * i + 1] = 0.0;
        }
        for (int it = 0; it < nit; ++it) {
            const int s = it % stages;
            const uint32_t ph = (uint32_t)(it / stages) & 1u;
            wait_w(full + s, ph, a, 4, it, s);
            if (lead) {
                wait_w(vfull + s, ph, a, 5, it, s);
                if (h == 0) TRACE(3, it);  // helper: start
                const double *base = sx + stage_doubles * s + (size_t)(F_HSHARE * h) * pitch;
#pragma unroll 4
                for (int q2 = 0; q2 < F_HSHARE; ++q2) {
                    const double vj = sv[s * F_ROWS + F_HSHARE * h + q2];
                    const double rj = sy[s * F_ROWS + F_HSHARE * h + q2];
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
            if (h == 0) TRACE(4, it);  // helper: done
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
        // descriptor: x = segment | count << 8, y = four tile-row numbers (bytes)
        const uint2 d = __ldg(a.plan + blockIdx.x * F_CWARPS + warp);
        const int seg = (int)(d.x & 0xffu), cnt = (int)((d.x >> 8) & 0xffu);
        const int g = lane >> 2, k4 = lane & 3;
        int tr[4], nt[4], offA[4], offB[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            tr[u] = (int)((d.y >> (8 * u)) & 0xffu);
            const int lim = tr[u] - 4 * seg + 1;  // tiles left of (and on) the diagonal in this segment
            nt[u] = u < cnt ? (lim < 4 ? lim : 4) : 0;
            offA[u] = nt[u] > 0 ? __ldg(a.colidx + tr[u] * 8 + g) : 0;
            offB[u] = cnt > 0 ? __ldg(a.colidx + (4 * seg + u) * 8 + g) : 0;
        }
        double acc[4][4][2];
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int n = 0; n < 4; ++n) acc[u][n][0] = acc[u][n][1] = 0.0;
        for (int it = 0; it < nit; ++it) {
            const int s = it % stages;
            const uint32_t ph = (uint32_t)(it / stages) & 1u;
            wait_w(full + s, ph, a, 6, it, s);
            wait_w(vfull + s, ph, a, 7, it, s);
            if (warp < 3) TRACE(5 + 2 * warp, it);  // MMA warp 0..2: start
            if (cnt > 0) {
                const double *base = sx + stage_doubles * s;
                const double *vs = sv + s * F_ROWS;
#pragma unroll
                for (int kk = 0; kk < F_ROWS / 4; ++kk) {
                    const int r = kk * 4 + k4;
                    const double *rowp = base + (size_t)r * pitch;
                    const double vv = vs[r];
                    double fa[4], fb[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) fa[u] = nt[u] > 0 ? rowp[offA[u]] : 0.0;
#pragma unroll
                    for (int n = 0; n < 4; ++n) fb[n] = rowp[offB[n]] * vv;
#pragma unroll
                    for (int u = 0; u < 4; ++u)
#pragma unroll
                        for (int n = 0; n < 4; ++n)
                            if (n < nt[u]) dmma884(acc[u][n][0], acc[u][n][1], fa[u], fb[n]);
                }
            }
            __syncwarp();
            if (warp < 3) TRACE(6 + 2 * warp, it);  // MMA warp 0..2: done
            if (lane == 0) mbar_arrive(empty + s);
        }
        // tile (tr[u], 4 seg + n) sits in 32x32 block (tr[u] / 4, seg) of the fold kernel's layout
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (nt[u] > 0) {
                const int bi = tr[u] >> 2, m = tr[u] & 3;
                double *out = a.gram_partials + ((size_t)blockIdx.y * a.nblk + (bi * (bi + 1) / 2 + seg)) * 1024;
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

// ---- tile-row plan -------------------------------------------------------------------------
// For a list of ntr = ceil(nG / 8) tile rows: segment s (columns 32 s .. 32 s + 31 of the list) holds
// tile rows r >= 4 s with min(4, r - 4 s + 1) tiles each.  A warp takes <= 4 tile rows of ONE segment
// (they share the four B operands).  Rows are packed first-fit-decreasing under a load bound T; the
// smallest feasible T over the 8 * ngroups warps wins, and ngroups grows only when 16 tiles per warp
// are not enough.  Built once per process, copied once per ctx.
constexpr int F_MAX_NTR = 52;   // 416 list columns, the widest row the staging ring holds
constexpr int F_PLAN_W = 128;   // descriptor slots per width

struct StepPlan {
    int ngroups[F_MAX_NTR + 1];
    int max_tiles[F_MAX_NTR + 1];
    std::vector<uint2> tab;
    int min_tiles = 1;   // ATE_PLAN_MINT: lower bound on the busiest warp's tile count (experiments)
    StepPlan() : tab((size_t)(F_MAX_NTR + 1) * F_PLAN_W, make_uint2(0u, 0u)) {
        if (const char *e = getenv("ATE_PLAN_MINT")) min_tiles = atoi(e);
        ngroups[0] = 1;
        max_tiles[0] = 0;
        for (int ntr = 1; ntr <= F_MAX_NTR; ++ntr) build(ntr);
    }
    static int pack(int ntr, int T, std::vector<uint2> *out) {
        int bins = 0;
        const int nseg = (ntr + 3) / 4;
        for (int sgm = 0; sgm < nseg; ++sgm) {
            struct Bin { int load, cnt, rows[4]; };
            std::vector<Bin> b;
            // tile rows of this segment by decreasing cost: the full ones (4 tiles), then 4 s + 3 .. 4 s
            std::vector<int> order;
            for (int r = 4 * sgm + 3; r < ntr; ++r) order.push_back(r);
            for (int r = (4 * sgm + 2 < ntr - 1 ? 4 * sgm + 2 : ntr - 1); r >= 4 * sgm; --r) order.push_back(r);
            for (int r : order) {
                const int c = r - 4 * sgm + 1 < 4 ? r - 4 * sgm + 1 : 4;
                size_t k = 0;
                for (; k < b.size(); ++k)
                    if (b[k].cnt < 4 && b[k].load + c <= T) break;
                if (k == b.size()) b.push_back(Bin{0, 0, {0, 0, 0, 0}});
                b[k].rows[b[k].cnt++] = r;
                b[k].load += c;
            }
            for (const Bin &x : b) {
                if (out)
                    out->push_back(make_uint2((unsigned)sgm | ((unsigned)x.cnt << 8),
                                              (unsigned)x.rows[0] | ((unsigned)x.rows[1] << 8) |
                                                  ((unsigned)x.rows[2] << 16) | ((unsigned)x.rows[3] << 24)));
                ++bins;
            }
        }
        return bins;
    }
    void build(int ntr) {
        const int total = ntr * (ntr + 1) / 2;
        // g groups of 8 warp slots.  First try to get by with FOUR warps per group -- warps 0-3 sit on
        // the four SM sub-partitions, and one warp per sub-partition already saturates its DMMA pipe
        // (16.7 cycles per DMMA.8x8x4 alone, 32 with two warps: tools/microbench/dmma_shapes.cu) --
        // then with all eight.
        for (int pass = 0; pass < 2 * (F_PLAN_W / 8); ++pass) {
            const int g = pass / 2 + 1, per = (pass % 2 == 0) ? 4 : 8;
            const int W = per * g;
            int T0 = (total + W - 1) / W;
            if (T0 < min_tiles) T0 = min_tiles;
            for (int T = T0; T <= 16; ++T) {
                if (T < 4 && ntr >= 4) continue;  // a full tile row is 4 tiles
                if (pack(ntr, T, nullptr) <= W) {
                    std::vector<uint2> d;
                    pack(ntr, T, &d);
                    for (size_t k = 0; k < d.size(); ++k)
                        tab[(size_t)ntr * F_PLAN_W + 8 * (k / per) + (k % per)] = d[k];
                    ngroups[ntr] = g;
                    max_tiles[ntr] = 0;  // the busiest warp's real load (a row wider than T sits alone in a bin)
                    for (const uint2 &e : d) {
                        const int sgm = (int)(e.x & 0xffu), cnt = (int)((e.x >> 8) & 0xffu);
                        int load = 0;
                        for (int u = 0; u < cnt; ++u) {
                            const int r = (int)((e.y >> (8 * u)) & 0xffu);
                            load += r - 4 * sgm + 1 < 4 ? r - 4 * sgm + 1 : 4;
                        }
                        if (load > max_tiles[ntr]) max_tiles[ntr] = load;
                    }
                    return;
                }
            }
        }
        ngroups[ntr] = -1;
        max_tiles[ntr] = -1;
    }
};

// ATE_STEP_WATCHDOG=1: a host-mapped log the kernel's barrier waits report into (see wait_w); printed
// at process exit when anything was reported.  Off by default: the kernel then uses the plain wait.
unsigned long long *g_wd_host = nullptr, *g_wd_dev = nullptr;
void wd_print() {
    if (!g_wd_host || g_wd_host[0] == 0) return;
    static const char *names[] = {"?", "producer(slot copy)/empty", "producer(row copies)/empty", "eval/full", "helper/full",
                                  "helper/vfull", "mma/full", "mma/vfull"};
    fprintf(stderr, "[ate] irls_step_kernel: %llu barrier waits never completed:\n", g_wd_host[0]);
    const unsigned long long n = g_wd_host[0] < 120 ? g_wd_host[0] : 120;
    for (unsigned long long k = 0; k < n; ++k) {
        const unsigned long long w = g_wd_host[4 + 4 * k], it = g_wd_host[5 + 4 * k], sp = g_wd_host[6 + 4 * k];
        const unsigned long long in = g_wd_host[7 + 4 * k];
        fprintf(stderr, "  cta(%llu,%llu) warp %2llu %-28s slot %llu stage %llu parity %llu | ncols %llu stages %llu nblk %llu "
                        "contiguous %llu slots/chunk %llu\n",
                (w >> 16) & 0xffff, w >> 32, (w >> 8) & 0xff, names[w & 0xff], it, sp & 0xffffffffu, sp >> 32, in & 0xffff,
                (in >> 16) & 0xff, (in >> 24) & 0xffff, (in >> 40) & 1, in >> 41);
    }
}
unsigned long long *watchdog_buffer() {
    static bool tried = false;
    if (!tried) {
        tried = true;
        const char *e = getenv("ATE_STEP_WATCHDOG");
        if (e && atoi(e) != 0) {
            void *h = nullptr, *d = nullptr;
            if (cudaHostAlloc(&h, 4096, cudaHostAllocMapped) == cudaSuccess && cudaHostGetDevicePointer(&d, h, 0) == cudaSuccess) {
                memset(h, 0, 4096);
                g_wd_host = (unsigned long long *)h;
                g_wd_dev = (unsigned long long *)d;
                atexit(wd_print);
            }
        }
    }
    return g_wd_dev;
}

const StepPlan &step_plan() {
    static const StepPlan plan;
    return plan;
}

}  // namespace

extern "C" int ate_step_plan_describe(int list_cols, int *ngroups, int *max_tiles, unsigned *desc) {
    if (list_cols < 0) return -1;
    const int ntr = (list_cols + 7) / 8;
    if (ntr > F_MAX_NTR) return -1;
    const StepPlan &plan = step_plan();
    if (plan.ngroups[ntr] < 1) return -1;
    if (ngroups) *ngroups = plan.ngroups[ntr];
    if (max_tiles) *max_tiles = plan.max_tiles[ntr];
    if (desc)
        for (int k = 0; k < F_PLAN_W; ++k) {
            desc[2 * k] = plan.tab[(size_t)ntr * F_PLAN_W + k].x;
            desc[2 * k + 1] = plan.tab[(size_t)ntr * F_PLAN_W + k].y;
        }
    return 0;
}

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
    const int ntr = (nG + 7) / 8;
    if (ntr > F_MAX_NTR) return ate_fail(ctx, -1, "irls_step: a %d-column list exceeds the tile plan", nG);
    const StepPlan &plan = step_plan();
    const int ngroups = plan.ngroups[ntr];
    if (ngroups < 1) return ate_fail(ctx, -1, "irls_step: no tile plan for %d list columns", nG);
    const bool plan_fresh = ctx->cache.find("irls.plan") == ctx->cache.end();
    void *dplan = nullptr;
    ATE_TRY(ate_cached(ctx, "irls.plan", sizeof(uint2) * plan.tab.size(), &dplan));
    if (plan_fresh)
        ATE_CUDA(ctx, cudaMemcpyAsync(dplan, plan.tab.data(), sizeof(uint2) * plan.tab.size(), cudaMemcpyHostToDevice,
                                      ctx->stream));
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
    a.plan = (const uint2 *)dplan + (size_t)ntr * F_PLAN_W;
    a.gram_partials = (double *)scr;
    a.eval_partials = (double *)scr + gram_doubles;
    a.v_out = in.v_out;
    a.dbg = watchdog_buffer();
    a.trace = 0;
#ifdef ATE_STEP_TRACE
    {
        static const char *want = getenv("ATE_STEP_TRACE_NG");  // "lo:hi": stamp launches with lo <= nG <= hi
        int lo = 33, hi = 64;
        if (want) sscanf(want, "%d:%d", &lo, &hi);
        if (nG >= lo && nG <= hi && in.nrows > 500000 && ngroups == 1) a.trace = nG;
        static bool reg = false;
        if (!reg) {
            reg = true;
            atexit([] {
                static long long h[TR_SLOTS * TR_EVENTS + 8];
                if (cudaMemcpyFromSymbol(h, g_trace, sizeof(h)) != cudaSuccess) return;
                const char *path = getenv("ATE_STEP_TRACE_OUT");
                FILE *f = fopen(path ? path : "step_trace.csv", "w");
                if (!f) return;
                fprintf(f, "# ncols %lld stages %lld nblk %lld slots %lld nG %lld\n", h[1], h[2], h[3], h[4], h[5]);
                fprintf(f, "slot,prod_free,eval_start,eval_done,help_start,help_done,m0_start,m0_done,m1_start,m1_done,m2_start,m2_done\n");
                for (int it = 0; it < TR_SLOTS; ++it) {
                    fprintf(f, "%d", it);
                    for (int e = 0; e < 11; ++e) fprintf(f, ",%lld", h[8 + it * TR_EVENTS + e] ? h[8 + it * TR_EVENTS + e] - h[0] : -1);
                    fprintf(f, "\n");
                }
                fclose(f);
            });
        }
    }
#endif
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
