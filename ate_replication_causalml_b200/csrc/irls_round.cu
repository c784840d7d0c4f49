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
constexpr int R_FMAX = 2;  // fits served by one CTA
constexpr size_t R_SMEM_LIMIT = 227 * 1024;

struct DProb {
    const double *X;
    long long ld;
    const int32_t *rows;
    long long nrows;
};
struct DFit {
    const double *coef;
    const int *colidx;
    long long hole_begin, hole_len;
    double w;
    double *gram_part;  // [chunks][nblk][1024]
    double *eval_part;  // [chunks][3 ni + 3 R_EWARPS]
    int nblk, pad;
};
struct DCta {
    int prob, nfit, chunk, pad;
    long long row_begin, row_end;
    int fit[R_FMAX], lead[R_FMAX];
    uint2 mma[R_CWARPS];  // x = fit slot | segment << 8 | count << 16, y = four tile-row numbers (bytes)
};
struct DFold {
    const double *gram_part, *eval_part;
    double *sums, *H;
    int nchunks, nblk, nG, ldh;
};
struct RArgs {
    const DProb *probs;
    const DFit *fits;
    const DCta *ctas;
    int ncols, ni, iy, stages;
};

__device__ __forceinline__ unsigned long long gtimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// mbarrier wait that cannot hang the GPU: after ~4 s (a launch takes milliseconds) it traps
__device__ __forceinline__ void wait_g(uint64_t *bar, uint32_t parity) {
    unsigned long long t0 = 0;
    for (unsigned spins = 1;; ++spins) {
        uint32_t done;
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
        if (done) return;
        if ((spins & 0xfffu) == 0) {
            const unsigned long long t = gtimer_ns();
            if (t0 == 0)
                t0 = t;
            else if (t - t0 > 4000000000ull)
                __trap();
        }
    }
}

template <int ROWS, int ITER, int FMAX, bool INIT>
__global__ void __launch_bounds__(R_THREADS, 1) irls_round_kernel(const RArgs a) {
    constexpr int HALVES = ROWS / 16;            // eval warps refresh 16 rows of a slot each
    constexpr int TURNS = R_EWARPS / HALVES;     // ... and take turns on the slots
    constexpr int PSHARE = ROWS / R_PWARPS;      // rows of a slot staged by one producer (row-list mode)
    constexpr int ITERH = (ITER + R_HWARPS - 1) / R_HWARPS;  // column pairs per helper lane
    static_assert(ROWS == 16 || ROWS == 32, "slot geometry");
    static_assert(FMAX >= 1 && FMAX <= 4 && FMAX <= R_FMAX, "fits per CTA");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ DCta scta;
    const int ncols = a.ncols, pitch = ncols + R_PAD, stages = a.stages;
    const size_t stage_doubles = (size_t)ROWS * pitch;
    double *sx = reinterpret_cast<double *>(smem_raw);
    double *sv = sx + stage_doubles * stages;           // [stages][FMAX][ROWS]: v of every fit
    double *sr = sv + (size_t)stages * FMAX * ROWS;     // [stages][FMAX][ROWS]: r of every fit
    uint64_t *full = reinterpret_cast<uint64_t *>(sr + (size_t)stages * FMAX * ROWS);
    uint64_t *vfull = full + stages, *empty = vfull + stages;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x < (int)(sizeof(DCta) / sizeof(int)))
        reinterpret_cast<int *>(&scta)[threadIdx.x] = reinterpret_cast<const int *>(a.ctas + blockIdx.x)[threadIdx.x];
    __syncthreads();
    const DProb prob = a.probs[scta.prob];
    const long long c_begin = scta.row_begin, c_end = scta.row_end;
    const int nit = (int)((c_end - c_begin + ROWS - 1) / ROWS);
    const int nfit = scta.nfit;
    const bool contiguous = (prob.rows == nullptr && prob.ld == (long long)pitch);

    for (int e = threadIdx.x; e < stages * ROWS * R_PAD; e += blockDim.x)
        sx[(size_t)(e / R_PAD) * pitch + ncols + (e % R_PAD)] = 0.0;  // the list's "zero column"
    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(full + s, contiguous ? 1 : R_PWARPS);
            mbar_init(vfull + s, HALVES);
            mbar_init(empty + s, R_CWARPS + R_HWARPS + R_EWARPS);  // EVERY consumer warp, every slot
        }
        mbar_fence_init();
    }
    __syncthreads();
    // Ring protocol: every consumer warp visits every slot in order and arrives on `empty` for it,
    // whether or not it has work there, and every producer waits on `empty` for every slot: no
    // barrier can complete a phase that one of its waiters has not seen the previous phase of.

    if (warp >= R_PRODUCER && warp < R_PRODUCER + R_PWARPS) {
        // ------------------------------ producers ------------------------------
        const int pw = warp - R_PRODUCER;
        if (contiguous) {
            if (pw != 0) return;
            const uint32_t rowb = (uint32_t)pitch * 8u;
            for (int it = 0; it < nit; ++it) {
                const int s = it % stages;
                const uint32_t ph = (uint32_t)(it / stages) & 1u;
                const long long t0 = c_begin + (long long)it * ROWS;
                const int cnt = (int)((c_end - t0) < ROWS ? (c_end - t0) : ROWS);
                wait_g(empty + s, ph ^ 1u);
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
                    wait_g(empty + s, ph ^ 1u);
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
        const int e = warp - R_EVAL;
        const int half = e % HALVES, turn = e / HALVES;
        const int ni = a.ni;
        double beta[FMAX][2 * ITER];
#pragma unroll
        for (int f = 0; f < FMAX; ++f) {
            const double *cf = f < nfit ? a.fits[scta.fit[f]].coef : nullptr;
#pragma unroll
            for (int i = 0; i < ITER; ++i) {
                const int c = 2 * lane + 64 * i;
                beta[f][2 * i] = (cf && c < ni) ? cf[c] : 0.0;
                beta[f][2 * i + 1] = (cf && c + 1 < ni) ? cf[c + 1] : 0.0;
            }
        }
        // the four lanes of a row group finish one fit each
        const int fsel = lane & 3;
        const bool mine = fsel < nfit && fsel < FMAX;
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
        const double pmin = GLM_PMIN, fmax_ = log(1.0 / pmin - 1.0), fmin_ = -fmax_;
        double s_xmz = 0.0, s_sr = 0.0, s_dev = 0.0;
        for (int it = 0; it < nit; ++it) {
            const int s = it % stages;
            const uint32_t ph = (uint32_t)(it / stages) & 1u;
            wait_g(full + s, ph);
            if ((it % TURNS) != turn) {  // not this warp's slot: release it at once
                if (lane == 0) mbar_arrive(empty + s);
                continue;
            }
            const double *slot = sx + stage_doubles * s;
#pragma unroll 1
            for (int grp = 0; grp < 2; ++grp) {
                const int r0 = 16 * half + 8 * grp;
                double part[FMAX][8];
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const double *rowp = slot + (size_t)(r0 + q) * pitch;
#pragma unroll
                    for (int f = 0; f < FMAX; ++f) part[f][q] = 0.0;
#pragma unroll
                    for (int i = 0; i < ITER; ++i) {
                        const int c = 2 * lane + 64 * i;
                        if (c < ncols) {
                            const double2 t2 = *reinterpret_cast<const double2 *>(rowp + c);
#pragma unroll
                            for (int f = 0; f < FMAX; ++f)
                                if (f < nfit)
                                    part[f][q] = fma(t2.x, beta[f][2 * i], fma(t2.y, beta[f][2 * i + 1], part[f][q]));
                        }
                    }
                }
                // reduce-scatter the 8 partial dots of every fit: lane l ends with row l >> 2
#pragma unroll
                for (int f = 0; f < FMAX; ++f) {
#pragma unroll
                    for (int hf = 4, off = 16; hf >= 1; hf >>= 1, off >>= 1) {
                        const bool up = (lane & off) != 0;
#pragma unroll
                        for (int k = 0; k < hf; ++k) {
                            const double send = up ? part[f][k] : part[f][k + hf], keep = up ? part[f][k + hf] : part[f][k];
                            part[f][k] = keep + __shfl_xor_sync(0xffffffffu, send, off);
                        }
                    }
                    part[f][0] += __shfl_xor_sync(0xffffffffu, part[f][0], 2);
                    part[f][0] += __shfl_xor_sync(0xffffffffu, part[f][0], 1);
                }
                double dot = part[0][0];
#pragma unroll
                for (int f = 1; f < FMAX; ++f)
                    if (fsel == f) dot = part[f][0];
                const int rr = r0 + (lane >> 2);
                const long long t = c_begin + (long long)it * ROWS + rr;
                const bool valid = mine && t < c_end && !(t >= hb && t < he);
                const double yi = slot[(size_t)rr * pitch + a.iy];
                const double fi = a0 + dot;
                double q;
                if (fi < fmin_)
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
                if (lead && valid) {
                    const double qc = fmin(fmax(q, pmin), 1.0 - pmin);
                    s_xmz += vi;
                    s_sr += ri;
                    // y is 0/1 on this path, so one of the two log terms vanishes
                    const double ll = (yi == 1.0) ? log(qc) : (yi == 0.0) ? log(1.0 - qc)
                                                            : yi * log(qc) + (1.0 - yi) * log(1.0 - qc);
                    s_dev -= w * ll;
                }
            }
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(vfull + s);  // MMA and helper warps may start on this slot
                mbar_arrive(empty + s);  // ... and this warp is done with its staged rows
            }
        }
        // lanes with the same fsel hold partial sums of the same fit
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
    } else if (warp >= R_HELPER) {
        // ------------------------------ helpers: X^T r, X^T v (and sum v x^2 on the first step) -----
        const int h = warp - R_HELPER;
        const int ni = a.ni;
        const int npairs = ncols >> 1, ppw = (npairs + R_HWARPS - 1) / R_HWARPS;
        bool any_lead = false;
#pragma unroll
        for (int f = 0; f < FMAX; ++f) any_lead |= (f < nfit && scta.lead[f] != 0);
        double gr[FMAX][2 * ITERH], hx[FMAX][2 * ITERH], hxx[INIT ? FMAX : 1][INIT ? 2 * ITERH : 1];
#pragma unroll
        for (int f = 0; f < FMAX; ++f)
#pragma unroll
            for (int i = 0; i < 2 * ITERH; ++i) {
                gr[f][i] = hx[f][i] = 0.0;
                if (INIT) hxx[f][i] = 0.0;
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
            wait_g(full + s, ph);
            if (any_lead) {
                wait_g(vfull + s, ph);
                const double *slot = sx + stage_doubles * s;
#pragma unroll 2
                for (int r = 0; r < ROWS; ++r) {
                    const double *rowp = slot + (size_t)r * pitch;
                    double2 x[ITERH];
#pragma unroll
                    for (int i = 0; i < ITERH; ++i)
                        x[i] = col[i] >= 0 ? *reinterpret_cast<const double2 *>(rowp + col[i]) : make_double2(0.0, 0.0);
#pragma unroll
                    for (int f = 0; f < FMAX; ++f) {
                        if (f < nfit) {
                            const double vj = sv[((size_t)s * FMAX + f) * ROWS + r];
                            const double rj = sr[((size_t)s * FMAX + f) * ROWS + r];
#pragma unroll
                            for (int i = 0; i < ITERH; ++i) {
                                gr[f][2 * i] = fma(rj, x[i].x, gr[f][2 * i]);
                                gr[f][2 * i + 1] = fma(rj, x[i].y, gr[f][2 * i + 1]);
                                hx[f][2 * i] = fma(vj, x[i].x, hx[f][2 * i]);
                                hx[f][2 * i + 1] = fma(vj, x[i].y, hx[f][2 * i + 1]);
                                if (INIT) {
                                    hxx[f][2 * i] = fma(vj * x[i].x, x[i].x, hxx[f][2 * i]);
                                    hxx[f][2 * i + 1] = fma(vj * x[i].y, x[i].y, hxx[f][2 * i + 1]);
                                }
                            }
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty + s);
        }
#pragma unroll
        for (int f = 0; f < FMAX; ++f) {
            if (f < nfit && scta.lead[f] != 0) {
                double *ep = a.fits[scta.fit[f]].eval_part + (size_t)scta.chunk * (3 * ni + 3 * R_EWARPS);
#pragma unroll
                for (int i = 0; i < ITERH; ++i)
#pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
                        const int c = col[i] + hh;
                        if (col[i] >= 0 && c < ni) {
                            ep[c] = gr[f][2 * i + hh];
                            ep[ni + c] = hx[f][2 * i + hh];
                            if (INIT) ep[2 * ni + c] = hxx[f][2 * i + hh];
                        }
                    }
            }
        }
    } else {
        // ------------------------------ MMA ------------------------------
        const uint2 d = scta.mma[warp];
        const int fs = (int)(d.x & 0xffu), seg = (int)((d.x >> 8) & 0xffu), cnt = (int)((d.x >> 16) & 0xffu);
        const int g = lane >> 2, k4 = lane & 3;
        int nt[4], offA[4], offB[4], tr[4];
        const int *colidx = cnt > 0 ? a.fits[scta.fit[fs]].colidx : nullptr;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            tr[u] = (int)((d.y >> (8 * u)) & 0xffu);
            const int lim = tr[u] - 4 * seg + 1;  // tiles left of (and on) the diagonal in this segment
            nt[u] = u < cnt ? (lim < 4 ? lim : 4) : 0;
            offA[u] = nt[u] > 0 ? __ldg(colidx + tr[u] * 8 + g) : 0;
            offB[u] = cnt > 0 ? __ldg(colidx + (4 * seg + u) * 8 + g) : 0;
        }
        double acc[4][4][2];
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int n = 0; n < 4; ++n) acc[u][n][0] = acc[u][n][1] = 0.0;
        for (int it = 0; it < nit; ++it) {
            const int s = it % stages;
            const uint32_t ph = (uint32_t)(it / stages) & 1u;
            wait_g(full + s, ph);
            if (cnt > 0) {
                wait_g(vfull + s, ph);
                const double *base = sx + stage_doubles * s;
                const double *vs = sv + ((size_t)s * FMAX + fs) * ROWS;
#pragma unroll
                for (int kk = 0; kk < ROWS / 4; ++kk) {
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
            if (lane == 0) mbar_arrive(empty + s);
        }
        // tile (tr[u], 4 seg + n) sits in 32x32 block (tr[u] / 4, seg) of the fold kernel's layout
        if (cnt > 0) {
            const DFit &ft = a.fits[scta.fit[fs]];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (nt[u] > 0) {
                    const int bi = tr[u] >> 2, m = tr[u] & 3;
                    double *out = ft.gram_part + ((size_t)scta.chunk * ft.nblk + (bi * (bi + 1) / 2 + seg)) * 1024;
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
}

// ---- host: tile-row bins ---------------------------------------------------------------------
// For a list of ntr = ceil(nG / 8) tile rows: segment sg (columns 32 sg .. 32 sg + 31 of the list)
// holds tile rows r >= 4 sg with min(4, r - 4 sg + 1) tiles each.  A bin is what one MMA warp takes:
// <= 4 tile rows of ONE segment (they share the four B operands), packed first-fit-decreasing under
// a load bound T.
struct Bin {
    int seg, cnt, load, rows[4];
};
void pack_bins(int ntr, int T, std::vector<Bin> &out) {
    out.clear();
    const int nseg = (ntr + 3) / 4;
    for (int sg = 0; sg < nseg; ++sg) {
        const size_t first = out.size();
        std::vector<int> order;
        for (int r = 4 * sg + 3; r < ntr; ++r) order.push_back(r);
        for (int r = (4 * sg + 2 < ntr - 1 ? 4 * sg + 2 : ntr - 1); r >= 4 * sg; --r) order.push_back(r);
        for (int r : order) {
            const int c = r - 4 * sg + 1 < 4 ? r - 4 * sg + 1 : 4;
            size_t k = first;
            for (; k < out.size(); ++k)
                if (out[k].cnt < 4 && out[k].load + c <= T) break;
            if (k == out.size()) out.push_back(Bin{sg, 0, 0, {0, 0, 0, 0}});
            out[k].rows[out[k].cnt++] = r;
            out[k].load += c;
        }
    }
}

struct Group {
    int prob = 0;
    int nfit = 0, fit[R_FMAX] = {0, 0}, lead[R_FMAX] = {0, 0};
    std::vector<std::pair<int, Bin>> bins;  // (fit slot, bin), <= R_CWARPS
    int tiles = 0;
    double cost = 0.0;
    int chunks = 1;
};


// Which fits share a CTA and which Gram tiles each of its MMA warps owns.  Fits of one problem are
// sorted by list width; a list whose bins exceed one CTA's eight MMA warps is split over several CTAs
// that stage the same rows (the first leads: it alone evaluates the sums); the others are packed
// first-fit, up to `fmax` fits per CTA; finally a group's tiles are re-dealt with the smallest load bound
// that still fits eight warps.  Returns -1 for a list the tile numbering cannot hold.
int build_groups(int P, const std::vector<int> &prob_of, const std::vector<int> &live, const std::vector<int> &nGs,
                 int fmax, std::vector<Group> &groups) {
    const int nfits = (int)prob_of.size();
    std::vector<Bin> bins;
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
            const int ntr = (nG + 7) / 8, tiles = ntr * (ntr + 1) / 2;
            pack_bins(ntr, 16, bins);
            if ((int)bins.size() > R_CWARPS) {
                for (size_t b0 = 0; b0 < bins.size(); b0 += R_CWARPS) {
                    Group g;
                    g.prob = p; g.nfit = 1; g.fit[0] = f; g.lead[0] = (b0 == 0);
                    for (size_t b = b0; b < bins.size() && b < b0 + R_CWARPS; ++b) {
                        g.bins.emplace_back(0, bins[b]);
                        g.tiles += bins[b].load;
                    }
                    groups.push_back(g);
                }
                continue;
            }
            bool placed = false;
            for (Group &g : open)
                if (g.nfit < fmax && (int)g.bins.size() + (int)bins.size() <= R_CWARPS) {
                    const int sl = g.nfit++;
                    g.fit[sl] = f; g.lead[sl] = 1;
                    for (auto &b : bins) g.bins.emplace_back(sl, b);
                    g.tiles += tiles;
                    placed = true;
                    break;
                }
            if (!placed) {
                Group g;
                g.prob = p; g.nfit = 1; g.fit[0] = f; g.lead[0] = 1;
                for (auto &b : bins) g.bins.emplace_back(0, b);
                g.tiles = tiles;
                open.push_back(g);
            }
        }
        for (Group &g : open) {
            if (g.tiles > 0 && (int)g.bins.size() < R_CWARPS) {
                for (int T = std::max(4, (g.tiles + R_CWARPS - 1) / R_CWARPS); T < 16; ++T) {
                    std::vector<std::pair<int, Bin>> nb;
                    for (int sl = 0; sl < g.nfit; ++sl) {
                        pack_bins((nGs[g.fit[sl]] + 7) / 8, T, bins);
                        for (auto &b : bins) nb.emplace_back(sl, b);
                    }
                    if ((int)nb.size() <= R_CWARPS) {
                        g.bins.swap(nb);
                        break;
                    }
                }
            }
            groups.push_back(g);
        }
    }
    return 0;
}

unsigned bin_x(int fitslot, const Bin &bn) { return (unsigned)fitslot | ((unsigned)bn.seg << 8) | ((unsigned)bn.cnt << 16); }
unsigned bin_y(const Bin &bn) {
    return (unsigned)bn.rows[0] | ((unsigned)bn.rows[1] << 8) | ((unsigned)bn.rows[2] << 16) | ((unsigned)bn.rows[3] << 24);
}

template <int ROWS, int ITER, int FMAX>
int launch_round(ate_ctx *ctx, const RArgs &a, int nctas, size_t smem, bool init) {
    if (init) {
        auto k = irls_round_kernel<ROWS, ITER, FMAX, true>;
        if (cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)R_SMEM_LIMIT) != cudaSuccess) return -1;
        k<<<nctas, R_THREADS, smem, ctx->stream>>>(a);
    } else {
        auto k = irls_round_kernel<ROWS, ITER, FMAX, false>;
        if (cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)R_SMEM_LIMIT) != cudaSuccess) return -1;
        k<<<nctas, R_THREADS, smem, ctx->stream>>>(a);
    }
    return 0;
}

}  // namespace

// host-only introspection of the planner (include/ate_b200.h)
extern "C" int ate_round_plan_describe(int nfit, const int *list_cols, int staged_cols, int max_groups, unsigned *desc) {
    if (nfit < 1 || !list_cols || !desc || max_groups < 1 || staged_cols < 8) return -1;
    const int fmax = (staged_cols + 63) / 64 <= 5 ? R_FMAX : 1;
    std::vector<int> prob_of(nfit, 0), live(nfit, 1), nGs(list_cols, list_cols + nfit);
    std::vector<Group> groups;
    if (build_groups(1, prob_of, live, nGs, fmax, groups) < 0) return -1;
    if ((int)groups.size() > max_groups) return -2;
    for (size_t gi = 0; gi < groups.size(); ++gi) {
        unsigned *d = desc + 24 * gi;
        const Group &g = groups[gi];
        for (int k = 0; k < 24; ++k) d[k] = 0u;
        d[0] = (unsigned)g.nfit;
        for (int sl = 0; sl < g.nfit; ++sl) {
            d[1 + sl] = (unsigned)g.fit[sl];
            d[3 + sl] = (unsigned)g.lead[sl];
        }
        for (size_t b = 0; b < g.bins.size(); ++b) {
            d[8 + 2 * b] = bin_x(g.bins[b].first, g.bins[b].second);
            d[9 + 2 * b] = bin_y(g.bins[b].second);
        }
    }
    return (int)groups.size();
}

// slot geometry for a staged width: 32-row slots while three of them fit, else 16-row slots
static int round_rows_for(int ncs) {
    const size_t b32 = sizeof(double) * ((size_t)32 * (ncs + R_PAD) + 2 * R_FMAX * 32);
    return (R_SMEM_LIMIT - 2048) / b32 >= 3 ? 32 : 16;
}

int ate_i_irls_round(ate_ctx *ctx, int ncs, int ni, int iy, int init, const std::vector<RoundProblem> &probs,
                     const std::vector<RoundFit> &fits, std::vector<RoundSlot> &slots, double **packed,
                     long long *packed_count, RoundStats *stats) {
    const int nfits = (int)fits.size(), P = (int)probs.size();
    if (ncs % 8 != 0 || ncs <= ni - 1 || iy >= ncs || ni < 1) return ate_fail(ctx, -1, "irls_round: bad staged width %d", ncs);
    const int it_e = (ncs + 63) / 64;
    const int ROWS = round_rows_for(ncs);
    const int fmax = it_e <= 5 ? R_FMAX : 1;  // the eval warps hold a beta slice per fit in registers
    const size_t stage_bytes = sizeof(double) * ((size_t)ROWS * (ncs + R_PAD) + 2 * (size_t)fmax * ROWS);
    int stages = (int)((R_SMEM_LIMIT - 2048) / stage_bytes);
    if (stages > R_MAX_STAGES) stages = R_MAX_STAGES;
    if (stages < 2 || it_e > 13) return ate_fail(ctx, -1, "irls_round: %d staged columns do not fit the step kernel", ncs);
    const size_t smem = stages * stage_bytes + 3 * stages * sizeof(uint64_t);
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
        if (build_groups(P, prob_of, live, nGs, fmax, groups) < 0)
            return ate_fail(ctx, -1, "irls_round: a column list exceeds the tile plan");
    }
    if (groups.empty()) {
        *packed_count = 0;
        return 0;
    }
    // ---- row chunks per group: the SMs are dealt in proportion to the estimated work ----
    double total_cost = 0.0;
    for (Group &g : groups) {
        int wmax = 0;
        for (auto &b : g.bins) wmax = std::max(wmax, b.second.load);
        g.cost = (double)probs[g.prob].nrows * (30.0 + 25.0 * g.nfit + 4.0 * wmax + 0.5 * g.tiles);
        total_cost += g.cost;
    }
    const int budget = std::max(ctx->num_sms, (int)groups.size());
    int used = 0;
    for (Group &g : groups) {
        long long maxc = (probs[g.prob].nrows + 4LL * ROWS - 1) / (4LL * ROWS);
        if (maxc < 1) maxc = 1;
        int c = (int)(g.cost / total_cost * budget);
        if (c < 1) c = 1;
        if (c > maxc) c = (int)maxc;
        g.chunks = c;
        used += c;
    }
    for (bool grew = true; used < budget && grew;) {  // hand out the remainder to the most loaded groups
        grew = false;
        Group *best = nullptr;
        for (Group &g : groups) {
            const long long maxc = std::max(1LL, (probs[g.prob].nrows + 4LL * ROWS - 1) / (4LL * ROWS));
            if (g.chunks >= maxc) continue;
            if (!best || g.cost / g.chunks > best->cost / best->chunks) best = &g;
        }
        if (best) {
            best->chunks++;
            used++;
            grew = true;
        }
    }
    // all CTAs of one fit must cut the rows the same way (they write the same partial slots)
    for (size_t i = 0; i < groups.size(); ++i)
        if (!groups[i].lead[0] && i > 0) groups[i].chunks = groups[i - 1].chunks;

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
        part_doubles += (size_t)eplen * fit_chunks[f];
    }
    void *dpacked = nullptr, *dpart = nullptr;
    ATE_TRY(ate_cached(ctx, "round.packed", sizeof(double) * (size_t)std::max(pk, 1LL), &dpacked));
    ATE_TRY(ate_cached(ctx, "round.partials", sizeof(double) * std::max(part_doubles, (size_t)1), &dpart));
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
    ATE_TRY(ate_cached_pinned(ctx, "round.hdesc", total, &hst));
    ATE_TRY(ate_cached(ctx, "round.ddesc", total, &ddesc));
    memset(hst, 0, total);
    DProb *hp = (DProb *)hst;
    DFit *hf = (DFit *)((char *)hst + o_fits);
    DCta *hc = (DCta *)((char *)hst + o_ctas);
    DFold *hd = (DFold *)((char *)hst + o_folds);
    for (int p = 0; p < P; ++p) hp[p] = DProb{probs[p].X, probs[p].ld, probs[p].rows, probs[p].nrows};
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
        ++li;
    }
    RoundStats st;
    int ci = 0;
    for (Group &g : groups) {
        const long long nrows = probs[g.prob].nrows;
        long long rpc = (nrows + g.chunks - 1) / g.chunks;
        rpc = (rpc + ROWS - 1) / ROWS * ROWS;
        for (int c = 0; c < g.chunks; ++c, ++ci) {
            DCta &d = hc[ci];
            d.prob = g.prob;
            d.nfit = g.nfit;
            d.chunk = c;
            d.row_begin = std::min(nrows, (long long)c * rpc);
            d.row_end = std::min(nrows, (long long)(c + 1) * rpc);
            for (int sl = 0; sl < g.nfit; ++sl) {
                d.fit[sl] = g.fit[sl];
                d.lead[sl] = g.lead[sl];
            }
            for (size_t b = 0; b < g.bins.size() && b < (size_t)R_CWARPS; ++b) {
                d.mma[b] = make_uint2(bin_x(g.bins[b].first, g.bins[b].second), bin_y(g.bins[b].second));
            }
            st.rows_staged += d.row_end - d.row_begin;
        }
        for (int sl = 0; sl < g.nfit; ++sl)
            if (g.lead[sl]) {
                const double nG = fits[g.fit[sl]].nG;
                st.row_fits += nrows;
                st.flops += (double)nrows * (nG * (nG + 1.0) + 6.0 * ni);
                st.alg_bytes += (double)nrows * ((double)ni * 8.0 + 24.0);  // n (p 8 + 24) per fit and step
            }
    }
    st.ctas = nctas;
    ATE_CUDA(ctx, cudaMemcpyAsync(ddesc, hst, total, cudaMemcpyHostToDevice, ctx->stream));
    RArgs a;
    a.probs = (const DProb *)ddesc;
    a.fits = (const DFit *)((char *)ddesc + o_fits);
    a.ctas = (const DCta *)((char *)ddesc + o_ctas);
    a.ncols = ncs; a.ni = ni; a.iy = iy; a.stages = stages;
    const DFold *dfolds = (const DFold *)((char *)ddesc + o_folds);

    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ctx->profile & 1) {
        ATE_CUDA(ctx, cudaEventCreate(&e0));
        ATE_CUDA(ctx, cudaEventCreate(&e1));
        ATE_CUDA(ctx, cudaEventRecord(e0, ctx->stream));
    }
    int rc = -1;
    const bool in = init != 0;
    if (ROWS == 32) {  // staged widths up to ~286 columns: at most five 64-column slices, two fits per CTA
        if (it_e <= 1) rc = launch_round<32, 1, 2>(ctx, a, nctas, smem, in);
        else if (it_e <= 2) rc = launch_round<32, 2, 2>(ctx, a, nctas, smem, in);
        else if (it_e <= 3) rc = launch_round<32, 3, 2>(ctx, a, nctas, smem, in);
        else if (it_e <= 4) rc = launch_round<32, 4, 2>(ctx, a, nctas, smem, in);
        else rc = launch_round<32, 5, 2>(ctx, a, nctas, smem, in);
    } else {
        if (fmax == 2) rc = launch_round<16, 5, 2>(ctx, a, nctas, smem, in);
        else if (it_e <= 7) rc = launch_round<16, 7, 1>(ctx, a, nctas, smem, in);
        else if (it_e <= 9) rc = launch_round<16, 9, 1>(ctx, a, nctas, smem, in);
        else rc = launch_round<16, 13, 1>(ctx, a, nctas, smem, in);
    }
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
