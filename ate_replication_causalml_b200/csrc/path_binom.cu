// path_binom.cu -- a4: L1-logistic path + cv.glmnet (ate_functions.R:139-141 -> glmnet's
// lognet, two-class Newton) as lock-step IRLS rounds in Gram space.
//
// glmnet's lognet2n touches the n observations in three places per IRLS step: the curvature
// xv_j = sum v x_j^2, the coordinate gradients g_k = <r, x_k> inside the coordinate-descent
// sweeps (O(n) each), and the refresh of (q, v, r).  Here every O(n) piece of one IRLS step
// is ONE pass over the rows (irls_step.cu): eta, q (glmnet's clamping), v, r, sum v, sum r,
// deviance, X^T r, X^T v, and H = X_L^T V X_L on the DMMA pipe over a column list L that
// covers the strong set.  The coordinate descent then runs on the p x p quadratic:
//   lognet_pre_kernel  : one CTA per fit: convergence test of the previous quadratic, KKT check,
//                        recording / advancing lambda, strong-rule update.  Only the strong
//                        set's columns enter the Gram (on average ~1/6 of p at C1, i.e. ~4 % of
//                        the flops).  If a column joined that the step's list missed (rare: the
//                        list carries a 10 % margin), the solve is put off by one round: the list
//                        is rebuilt around the exact strong set and the step repeats at the same
//                        coefficients -- no host decision is involved.  Last, the column list of
//                        the NEXT step is built (strong set + columns within 10 % of the next
//                        threshold): the host reads its length while the solve still runs.
//   lognet_solve_kernel: one CTA per fit.  Standardises the compact H in place (lstandard1's
//                        centring and 1/N scaling are affine in H, X^T v, sum v), then follows
//                        lognet2n's loop order exactly -- strong-set sweep, intercept step,
//                        active-set loop -- with g_k = <r, x_k> kept current by covariance
//                        updates g -= d * H[:,k].  In exact arithmetic the iterates equal glmnet's.
// All fits of a batch (the full-data fits of every outer fold, then all their CV folds) advance
// one IRLS step per round, ONE launch of the step kernel for all of them (irls_round.cu); finished
// fits drop out.  The host reads one small flag vector per round (who is done, next list lengths).
// When several GPUs share the fit each holds a slice of the rows: the packed per-fit sums and Grams
// of a round are all-reduced on the stream before the bookkeeping, after which every rank runs the
// same p x p work on the same numbers and takes the same decisions.
#include <math.h>
#include <stdlib.h>

#include <chrono>
#include <type_traits>

#include "irls_round.cuh"
#include "tma_dmma.cuh"
#include "path.cuh"

namespace {

enum { PH_INIT = 0, PH_REFRESH = 1 };
enum { SD_B0 = 0, SD_BS0, SD_AL, SD_AL0, SD_DEV0, SD_DEV1, SD_SHR, SD_ALF, SD_Q0, SD_BZ, SD_SUMS, SD_SUMS2, SD_SUMA, SD_N };
enum { SI_NIN = 0, SI_ILM, SI_PHASE, SI_DONE, SI_NLP, SI_JERR, SI_LMU, SI_ROUNDS, SI_NS, SI_SOLVE, SI_NG, SI_RETRY, SI_MISS, SI_ONES, SI_NEWLIST, SI_N };
// sums layout: gr[ni] = X^T r, hx[ni] = X^T v, hxx[ni] = sum v x^2, then XMZ, SR, DEV
enum { SU_XMZ = 0, SU_SR, SU_DEV, SU_N };

struct LognetFit {
    const double *X;      // the design this fit reads: the shared one (with a row list) or a
    long long ld;         //   private contiguous copy of its problem's rows (list == NULL)
    const int32_t *list;  // row ids (NULL = identity)
    long long hole_begin, hole_len;  // this rank's rows of the held-out inner fold (positions in the problem's rows)
    double nt;            // observations of the fit over ALL ranks
    int prob, seg;        // problem index; held-out segment (-1: the full-data fit)
    int ni, iy, ldh, ncs;  // ncs: columns staged per row by the step kernel = its zero slot
    const double *vp;
    double flmin;
    const double *ulam;
    const int *nlam_ptr;  // user lambdas: how many; a cv-fold fit: its full-data fit's lmu so far
    const int *lead_done; // a cv-fold fit: its full-data fit's SI_DONE (nlam_ptr is final only then); else NULL
    double *seq;          // a full-data fit: the lambda sequence its cv-fold fits follow (their ulam)
    int nlam;
    double thr;
    int maxit;
    double *coef;     // ni + 1 (a0 last): raw-scale coefficients for the next eval
    double *st_d;     // b, bs, xv, ga, xm, xs, vq [7 ni] + SD_N
    int *st_i;        // ixx, mm, mlist, ju, spos, gpos [6 ni] + slist [ldh] + glist [ldh] + SI_N + next glist [ldh] + next gpos [ni]
    double *alm, *a0, *dev, *beta;  // outputs
};

constexpr int EV_THREADS = 256;
constexpr int CD_THREADS = 256;

// where a fit's sums and Gram of the current round sit (the packed buffer of irls_round.cu)
struct RoundDev {
    const double *sums;  // X^T r [ni], X^T v [ni], sum v x^2 [ni], then XMZ, SR, DEV
    double *H;           // ldh x ldh over the column list of the round
    int ldh, live;       // live = 0: the fit takes no part in this round
};
// lambdas a fit may run: all of its own sequence, or, for a cv-fold fit, those its full-data fit reached
// (known once that fit is done; until then the fold fit simply continues)
__device__ __forceinline__ int fit_nlam(const LognetFit &f) {
    if (f.flmin < 1.0) return f.nlam;
    if (f.lead_done && !*f.lead_done) return f.nlam;
    return min(f.nlam, *f.nlam_ptr);
}
__device__ __forceinline__ int *fit_slist(const LognetFit &f) { return f.st_i + 6 * (size_t)f.ni; }
__device__ __forceinline__ int *fit_glist(const LognetFit &f) { return f.st_i + 6 * (size_t)f.ni + f.ldh; }
__device__ __forceinline__ int *fit_scal_i(const LognetFit &f) { return f.st_i + 6 * (size_t)f.ni + 2 * (size_t)f.ldh; }
// the list the NEXT step will use (built by lognet_pre_kernel while this step's H is still in use; copied over
// glist / gpos by lognet_solve_kernel when it is through with H)
__device__ __forceinline__ int *fit_glist_next(const LognetFit &f) { return f.st_i + 6 * (size_t)f.ni + 2 * (size_t)f.ldh + SI_N; }
__device__ __forceinline__ int *fit_gpos_next(const LognetFit &f) { return fit_glist_next(f) + f.ldh; }
__device__ __forceinline__ void fit_write_flags(const LognetFit &f, int *flags_out) {
    const int *si = f.st_i + 6 * (size_t)f.ni + 2 * (size_t)f.ldh;
    flags_out[4 * blockIdx.x] = si[SI_DONE] ? 1 : 0;
    flags_out[4 * blockIdx.x + 1] = si[SI_DONE] ? 0 : si[SI_NG];
    flags_out[4 * blockIdx.x + 2] = si[SI_LMU];
    flags_out[4 * blockIdx.x + 3] = si[SI_ONES];
}
__device__ __forceinline__ double *fit_scal_d(const LognetFit &f) { return f.st_d + 7 * (size_t)f.ni; }

// one warp per output element, partials folded in a fixed (lane-strided, then butterfly) order
__global__ void __launch_bounds__(256)
fold_partials_kernel(const double *__restrict__ partials, int nparts, int len, double *__restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int e = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (e >= len) return;
    double s = 0.0;
    for (int q = lane; q < nparts; q += 32) s += partials[(size_t)q * len + e];
    s = warp_sum(s);
    if (lane == 0) out[e] = s;
}

// sum of the response column over inner-fold segment blockIdx.y of problem blockIdx.x (this rank's rows)
struct SegDesc {
    const double *X;
    long long ld;
    const int32_t *list;
    long long begin, end;
    int iy, pad;
};
__global__ void __launch_bounds__(256)
lognet_ysum_kernel(const SegDesc *__restrict__ segs, double *__restrict__ out) {
    const SegDesc sgm = segs[blockIdx.x];
    __shared__ double red[32];
    double acc[1] = {0.0};
    for (long long t = sgm.begin + threadIdx.x; t < sgm.end; t += blockDim.x) {
        const long long row = sgm.list ? (long long)__ldg(sgm.list + t) : t;
        acc[0] += __ldg(sgm.X + (size_t)row * (size_t)sgm.ld + sgm.iy);
    }
    block_sum<1>(acc, red);
    if (threadIdx.x == 0) out[blockIdx.x] = acc[0];
}

// q0 = mean of the response over the fit's rows (the problem's rows minus the held-out segment),
// bz = logit(q0); arms the first evaluation pass.  ysum[prob * nsegp + s], nsegp = max(nseg, 1).
__global__ void __launch_bounds__(256)
lognet_init_kernel(const LognetFit *__restrict__ fits, const double *__restrict__ ysum, int nsegp) {
    const LognetFit f = fits[blockIdx.x];
    for (int j = threadIdx.x; j < f.ni; j += blockDim.x) f.coef[j] = 0.0;
    if (threadIdx.x == 0) {
        double tot = 0.0;
        for (int sg = 0; sg < nsegp; ++sg)
            if (sg != f.seg) tot += ysum[f.prob * nsegp + sg];
        const double q0 = tot / f.nt;
        const double bz = log(q0 / (1.0 - q0));
        f.coef[f.ni] = bz;
        double *sd = fit_scal_d(f);
        int *si = fit_scal_i(f);
        for (int k = 0; k < SD_N; ++k) sd[k] = 0.0;
        for (int k = 0; k < SI_N; ++k) si[k] = 0;
        sd[SD_Q0] = q0;
        sd[SD_BZ] = bz;
        sd[SD_B0] = bz;
        si[SI_PHASE] = PH_INIT;
        si[SI_ONES] = -1;
        if (q0 <= GLM_PMIN) {
            si[SI_JERR] = 8001;
            si[SI_DONE] = 1;
        } else if (q0 >= 1.0 - GLM_PMIN) {
            si[SI_JERR] = 9001;
            si[SI_DONE] = 1;
        }
    }
}

__device__ __forceinline__ double bsum_all(double v, double *red) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) red[wid] = v;
    __syncthreads();
    double t = 0.0;
    for (int i = 0; i < nw; ++i) t += red[i];
    __syncthreads();
    return t;
}
__device__ __forceinline__ double bmax_all(double v, double *red) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();
    if (lane == 0) red[wid] = v;
    __syncthreads();
    double t = red[0];
    for (int i = 1; i < nw; ++i) t = fmax(t, red[i]);
    __syncthreads();
    return t;
}

// ---- IRLS round, part 1: bookkeeping that needs only the evaluation pass --------------------
// Convergence of the previous quadratic, KKT re-check, recording lambda, advancing it, and the
// strong-set update.  The step computed H over last round's prediction of the strong set; when a
// column joined that the prediction missed, SI_RETRY is raised: the solve kernel then only rebuilds
// the list (around the exact strong set), the step repeats at the same coefficients, and the next
// visit here goes straight to the solve.  flags_out[4 fit + {0, 1, 2, 3}] = {done, length of the next list,
// lambdas recorded, position of the ONE column in that list or -1}: what the host needs to plan the next launch.
__global__ void __launch_bounds__(CD_THREADS)
lognet_pre_kernel(const LognetFit *__restrict__ fits, const RoundDev *__restrict__ rdev, int *__restrict__ flags_out) {
    const LognetFit f = fits[blockIdx.x];
    const RoundDev rd = rdev[blockIdx.x];
    const int ni = f.ni, tid = threadIdx.x, T = blockDim.x;
    int *si = fit_scal_i(f);
    double *sd = fit_scal_d(f);
    // every thread reads the state these early exits depend on BEFORE thread 0 changes any of it (a warp that
    // read SI_RETRY after thread 0 had cleared it would run the bookkeeping below on its own)
    const bool st_done = si[SI_DONE] != 0, st_retry = si[SI_RETRY] != 0;
    const bool st_lead_stop = f.lead_done && *f.lead_done && si[SI_ILM] >= *f.nlam_ptr && si[SI_PHASE] != PH_INIT;
    __syncthreads();
    if (st_done || !rd.live) {
        if (tid == 0) {
            si[SI_SOLVE] = 0;
            fit_write_flags(f, flags_out);
        }
        return;
    }
    if (st_lead_stop) {
        // a cv-fold fit that ran ahead of its full-data fit, which has now stopped short of this lambda
        if (tid == 0) {
            si[SI_DONE] = 1;
            si[SI_SOLVE] = 0;
            fit_write_flags(f, flags_out);
        }
        return;
    }
    if (st_retry) {  // the bookkeeping of this step was done last round; H now covers the strong set
        if (tid == 0) {
            si[SI_RETRY] = 0;
            fit_write_flags(f, flags_out);  // the next step keeps the list (the strong set has not changed)
        }
        return;
    }
    extern __shared__ double sm[];
    double *g = sm, *red = g + ni;  // red[32]
    int *cand = reinterpret_cast<int *>(red + 32), *srt = cand + f.ldh;  // scratch of the list builder
    double *Gb = f.st_d, *Gbs = Gb + ni, *Gxv = Gbs + ni, *Gga = Gxv + ni, *Gxm = Gga + ni,
           *Gxs = Gxm + ni, *Gvq = Gxs + ni;
    int *Gixx = f.st_i, *Gmm = Gixx + ni, *Gml = Gmm + ni, *Gju = Gml + ni, *Gspos = Gju + ni, *Ggpos = Gspos + ni;
    int *slist = fit_slist(f);
    const double *gr = rd.sums, *hx = rd.sums + ni, *hxx = rd.sums + 2 * ni;
    const double xmz = rd.sums[3 * ni + SU_XMZ], sr_in = rd.sums[3 * ni + SU_SR];
    const double devsum = rd.sums[3 * ni + SU_DEV];
    const double pmin = GLM_PMIN, vmin = (1.0 + pmin) * pmin * (1.0 - pmin);

    int ilm = si[SI_ILM], nS = si[SI_NS];
    const int nin = si[SI_NIN];
    const int phase = si[SI_PHASE];
    const double b0 = sd[SD_B0];
    double al = sd[SD_AL], al0 = sd[SD_AL0];
    const int nlam = fit_nlam(f);
    const int mnl = min(GLM_MNLAM, nlam);
    double dev0 = sd[SD_DEV0], dev1 = sd[SD_DEV1], shr = sd[SD_SHR], alf = sd[SD_ALF];

    if (phase == PH_INIT) {
        // lstandard1 from the first pass (v = w q0 (1 - q0) is constant there)
        double vs = 0.0;
        for (int j = tid; j < ni; j += T) {
            const double m = hx[j] / xmz;
            const double sq = hxx[j] / xmz;
            const double var = sq - m * m;
            const bool ok = var > 1.0e-14 * (sq > 1.0e-300 ? sq : 1.0e-300);
            Gxm[j] = m;
            Gxs[j] = ok ? sqrt(var) : 1.0;
            Gju[j] = ok ? 1 : 0;
            const double t = f.vp ? fmax(f.vp[j], 0.0) : 1.0;
            Gvq[j] = t;
            vs += t;
            Gb[j] = 0.0;
            Gbs[j] = 0.0;
            Gxv[j] = 0.0;
            Gixx[j] = 0;
            Gmm[j] = 0;
            Gml[j] = 0;
            Gspos[j] = -1;
            Ggpos[j] = -1;
        }
        for (int c = tid; c < f.ldh; c += T) {
            slist[c] = f.ncs;  // sentinel: the staged rows' zero slot
            fit_glist(f)[c] = f.ncs;
        }
        vs = bsum_all(vs, red);
        for (int j = tid; j < ni; j += T) Gvq[j] = Gvq[j] * (double)ni / vs;
        __syncthreads();
    }
    // gradient of the refreshed quadratic on standardised columns: g_j = <r, x_j>
    for (int j = tid; j < ni; j += T) g[j] = Gju[j] ? (gr[j] - Gxm[j] * sr_in) / Gxs[j] : 0.0;
    __syncthreads();
    bool begin_lambda = false, finished = false;

    if (phase == PH_INIT) {
        const double q0 = sd[SD_Q0], bz = sd[SD_BZ];
        dev1 = -(bz * q0 + log(1.0 - q0));
        dev0 = dev1;  // y is 0/1: the y log y terms vanish
        shr = f.thr * dev0;
        alf = (f.flmin < 1.0) ? pow(fmax(GLM_EPS, f.flmin), 1.0 / (double)(nlam - 1)) : 1.0;
        for (int j = tid; j < ni; j += T) Gga[j] = Gju[j] ? fabs(g[j]) : 0.0;
        __syncthreads();
        ilm = 0;
        al = 0.0;
        begin_lambda = true;
    } else {
        // ---- the quadratic was just refreshed at the new coefficients: converged? ----
        const bool stop_path = (xmz <= vmin);
        bool again = false;
        if (!stop_path) {
            const double d0 = b0 - sd[SD_BS0];
            int moved = (xmz * d0 * d0 >= shr) ? 1 : 0;
            for (int l = tid; l < nin; l += T) {
                const int k = Gml[l];
                const double d = Gb[k] - Gbs[k];
                if (Gxv[k] * d * d >= shr) moved = 1;
            }
            moved = __syncthreads_or(moved);
            if (moved)
                again = true;
            else {
                int add = 0;
                for (int k = tid; k < ni; k += T) {
                    if (Gixx[k] || !Gju[k]) continue;
                    const double ga = fabs(g[k]);
                    Gga[k] = ga;
                    if (ga > al * Gvq[k]) {
                        Gixx[k] = 2;  // newly strong: appended to the list below
                        add = 1;
                    }
                }
                if (__syncthreads_or(add)) again = true;
            }
        }
        if (!again) {
            // ---- lambda ilm is done: record it on the caller's scale ----
            double part = 0.0;
            int nz = 0;
            for (int j = tid; j < ni; j += T) {
                const double bj = (Gb[j] != 0.0) ? Gb[j] / Gxs[j] : 0.0;
                f.beta[(size_t)ilm * ni + j] = bj;
                part += bj * Gxm[j];
                nz += (Gb[j] != 0.0);
            }
            const double dotbm = bsum_all(part, red);
            const int me = (int)bsum_all((double)nz, red);
            const double devr = (dev1 - devsum) / dev0;
            const double dprev = ilm > 0 ? f.dev[ilm - 1] : 0.0;
            if (tid == 0) {
                f.a0[ilm] = b0 - dotbm;
                f.alm[ilm] = al;
                f.dev[ilm] = devr;
                si[SI_LMU] = ilm + 1;
                if (f.seq && f.flmin < 1.0 && ilm == 2) {
                    // Three lambdas are recorded: the whole sequence is now determined (lambda_k =
                    // lambda_{k-1} alf, the products this fit itself will form; lambda_1 by R's fix.lam
                    // extrapolation), so the cv-fold fits may start while this fit is still running.
                    const double l1 = f.alm[1];
                    f.seq[0] = exp(2.0 * log(l1) - log(al));
                    f.seq[1] = l1;
                    double a = al;
                    for (int k = 2; k < f.nlam; ++k) {
                        f.seq[k] = a;
                        a *= alf;
                    }
                }
            }
            bool stop = stop_path;
            if (!stop && ilm + 1 >= mnl && f.flmin < 1.0) {
                if (me > ni + 1) stop = true;
                if (devr > GLM_DEVMAX) stop = true;
                if (devr - dprev < GLM_FDEV) stop = true;  // [EXT] absolute, as in glmnet 2.0-x
            }
            ++ilm;
            if (stop || ilm >= nlam)
                finished = true;
            else
                begin_lambda = true;
        }
    }
    if (begin_lambda) {
        al0 = al;
        if (f.flmin >= 1.0)
            al = f.ulam[ilm];
        else if (ilm > 1)
            al *= alf;
        else if (ilm == 0)
            al = GLM_BIG;
        else {
            double t = 0.0;
            for (int j = tid; j < ni; j += T)
                if (Gju[j] && Gvq[j] > 0.0) t = fmax(t, Gga[j] / Gvq[j]);
            al0 = bmax_all(t, red);
            al = alf * al0;
        }
        const double tlam = 2.0 * al - al0;
        for (int k = tid; k < ni; k += T)
            if (!Gixx[k] && Gju[k] && Gga[k] > tlam * Gvq[k]) Gixx[k] = 2;
    }
    __syncthreads();
    if (tid == 0) {
        // newly strong columns join the Gram's column list in ascending order (deterministic)
        int miss = 0;
        for (int k = 0; k < ni; ++k)
            if (Gixx[k] == 2) {
                Gixx[k] = 1;
                Gspos[k] = nS;
                slist[nS++] = k;
                if (Ggpos[k] < 0) miss = 1;  // not in the list the step built H over
            }
        if (finished) miss = 0;
        si[SI_RETRY] = miss;  // 1: the step repeats over a list rebuilt around the strong set
        si[SI_MISS] += miss;
        sd[SD_AL] = al;
        sd[SD_AL0] = al0;
        sd[SD_DEV0] = dev0;
        sd[SD_DEV1] = dev1;
        sd[SD_SHR] = shr;
        sd[SD_ALF] = alf;
        si[SI_ILM] = ilm;
        si[SI_NS] = nS;
        si[SI_PHASE] = PH_REFRESH;
        si[SI_SOLVE] = finished ? 0 : 1;
        if (finished) {
            si[SI_DONE] = 1;
            // fix.lam: lambda[1] was `big`; R rewrites it by log-linear extrapolation before the
            // sequence is handed to the CV folds
            if (f.flmin < 1.0 && si[SI_LMU] > 2) f.alm[0] = exp(2.0 * log(f.alm[1]) - log(f.alm[2]));
            if (f.seq)
                for (int k = 0; k < si[SI_LMU]; ++k) f.seq[k] = f.alm[k];
        }
    }
    __syncthreads();
    if (finished) {
        if (tid == 0) fit_write_flags(f, flags_out);
        return;
    }
    // ---- column list of the NEXT fused step: the strong set, then the likely joiners -- every column whose
    //      gradient is within 10 % of the next strong-rule / KKT threshold.  It depends on what this kernel has
    //      just decided (strong set, lambda) and on the gradient the round started from, not on the solve:
    //      built here, its length reaches the host while the solve still runs, and the host plans the next
    //      launch in that time.  The list goes to the `next` arrays: this round's H is indexed by the current
    //      ones until lognet_solve_kernel is through with it. ----
    double tnext;
    {
        const int nlam_n = fit_nlam(f);
        double aln;
        if (f.flmin >= 1.0)
            aln = (ilm + 1 < nlam_n) ? f.ulam[ilm + 1] : al;
        else if (ilm == 0) {
            double t = 0.0;
            for (int j = tid; j < ni; j += T)
                if (Gju[j] && Gvq[j] > 0.0) t = fmax(t, fabs(g[j]) / Gvq[j]);
            aln = alf * bmax_all(t, red);
        } else
            aln = al * alf;
        const double cur = (ilm == 0 && f.flmin < 1.0) ? aln / alf : al;  // lambda_1 = "big" has no meaning here
        tnext = 0.9 * fmin(2.0 * aln - cur, cur);
        if (!(tnext > 0.0)) tnext = 0.0;
    }
    int *glist = fit_glist_next(f), *gpos = fit_gpos_next(f);
    for (int j = tid; j < ni; j += T) gpos[j] = -1;
    nS = si[SI_NS];  // thread 0 appended the newly strong columns above (its copy alone was current)
    __syncthreads();
    if (tid < 32) {
        // one warp builds the list (a single thread took ~50 us per launch, all of it on the round's
        // critical path): ordered compaction by ballot, then every element computes its own position
        const int lane = tid;
        const unsigned lt = (1u << lane) - 1u;
        int nG = nS < f.ldh ? nS : f.ldh;
        for (int c = lane; c < nG; c += 32) cand[c] = slist[c];
        for (int k0 = 0; k0 < ni; k0 += 32) {  // the joiners, in ascending column order
            const int k = k0 + lane;
            const bool take = k < ni && !Gixx[k] && Gju[k] && fabs(g[k]) > tnext * Gvq[k];
            const unsigned m = __ballot_sync(0xffffffffu, take);
            const int at = nG + __popc(m & lt);
            if (take && at < f.ldh) cand[at] = k;
            nG = min(f.ldh, nG + __popc(m));
        }
        __syncwarp();
        // Order of the list.  An MMA warp loads, per 8x8 tile, 8 list columns x 4 staged rows; consecutive
        // rows sit 12 (mod 16) doubles apart, so four columns x four rows hit the 16 eight-byte banks once
        // each exactly when the four columns have four different residues mod 4 (a 64-bit shared load is
        // served half a warp at a time: class order 0 1 2 3 0 1 2 3, not 0 0 1 1 2 2 3 3, which cost 40 % extra
        // wavefronts in the C1 capture).  The classes take turns while they last: the r-th column of class q
        // (in the order above) goes to position sum_q' min(cnt[q'], r) + #{q' < q : cnt[q'] > r}.
        int run[4] = {0, 0, 0, 0};
        for (int c0 = 0; c0 < nG; c0 += 32) {
            const int c = c0 + lane;
            const int q = c < nG ? (cand[c] & 3) : -1;
            int r = 0;
#pragma unroll
            for (int qq = 0; qq < 4; ++qq) {
                const unsigned m = __ballot_sync(0xffffffffu, q == qq);
                if (q == qq) r = run[qq] + __popc(m & lt);
                run[qq] += __popc(m);
            }
            if (c < nG) srt[c] = r;
        }
        __syncwarp();
        for (int c = lane; c < nG; c += 32) {
            const int k = cand[c], q = k & 3, r = srt[c];
            int at = 0;
#pragma unroll
            for (int qq = 0; qq < 4; ++qq) at += min(run[qq], r) + ((qq < q && run[qq] > r) ? 1 : 0);
            glist[at] = k;
            gpos[k] = at;
        }
        // A wide list gets a CTA of its own in the step kernel, whose helpers are its slowest warps: such a
        // list ends with the staged rows' ONE column, the Gram's row of that column IS X^T v of the list
        // columns (all the solve reads of it), and the helpers compute X^T r only.
        int ones = -1;
        if (nG + 1 >= R_ONES_MIN && nG < f.ldh) {
            ones = nG;
            if (lane == 0) glist[nG] = f.ncs + 1;
            ++nG;
        }
        const int npad = (nG + 31) / 32 * 32;
        for (int c = nG + lane; c < npad && c < f.ldh; c += 32) glist[c] = f.ncs;
        __syncwarp();
        if (lane != 0) return;
        si[SI_NG] = nG;
        si[SI_ONES] = ones;
        si[SI_NEWLIST] = 1;
        fit_write_flags(f, flags_out);
    }
}

// ---- IRLS round, part 2: solve the penalised quadratic on the strong set -------------------
// H holds X_L^T V X_L over the column list L the step used (positions gpos).  At the end the list
// lognet_pre_kernel built for the next step replaces L.
__global__ void __launch_bounds__(CD_THREADS)
lognet_solve_kernel(const LognetFit *__restrict__ fits, const RoundDev *__restrict__ rdev, int hs_cap) {
    const LognetFit f = fits[blockIdx.x];
    const RoundDev rd = rdev[blockIdx.x];
    const int ni = f.ni, tid = threadIdx.x, T = blockDim.x, ldh = rd.ldh;
    int *si = fit_scal_i(f);
    double *sd = fit_scal_d(f);
    if (si[SI_DONE] || !si[SI_SOLVE] || !rd.live) return;
    const bool retry = si[SI_RETRY] != 0;  // list rebuild only: H does not cover the strong set yet
    extern __shared__ double sm[];
    double *b = sm, *g = b + ni, *hxs = g + ni, *xv = hxs + ni, *vq = xv + ni, *red = vq + ni;  // red[32]
    int *ixx = reinterpret_cast<int *>(red + 32), *mm = ixx + ni, *ml = mm + ni, *hp = ml + ni,
        *sl = hp + ni, *spos = sl + f.ldh;
    // the standardised strong block, nS x nS, in shared memory when it fits (hs_cap): every coordinate update
    // reads one row of it, and from L2 that load was most of the update's latency
    double *Hs = reinterpret_cast<double *>((reinterpret_cast<uintptr_t>(spos + ni) + 15) & ~(uintptr_t)15);
    double *Gb = f.st_d, *Gbs = Gb + ni, *Gxv = Gbs + ni, *Gxm = Gxv + 2 * ni, *Gxs = Gxm + ni,
           *Gvq = Gxs + ni;
    int *Gixx = f.st_i, *Gmm = Gixx + ni, *Gml = Gmm + ni, *Gju = Gml + ni, *Gspos = Gju + ni,
        *Ggpos = Gspos + ni;
    const int *slist = fit_slist(f);
    int *glist = fit_glist(f);
    double *H = rd.H;
    const double *gr = rd.sums, *hx = rd.sums + ni;
    const double xmz = rd.sums[3 * ni + SU_XMZ], sr_in = rd.sums[3 * ni + SU_SR];
    const int nS = si[SI_NS], ilm = si[SI_ILM];
    const int *hmap = Ggpos;
    const bool in_smem = nS <= hs_cap;
    int nin = si[SI_NIN], nlp = si[SI_NLP];
    double b0 = sd[SD_B0];
    const double al1 = sd[SD_AL], shr = sd[SD_SHR];

    for (int j = tid; j < ni; j += T) {
        const double m = Gxm[j], s = Gxs[j];
        const bool ok = Gju[j] != 0;
        b[j] = Gb[j];
        vq[j] = Gvq[j];
        ixx[j] = Gixx[j];
        mm[j] = Gmm[j];
        ml[j] = Gml[j];
        hp[j] = hmap[j];
        spos[j] = Gspos[j];
        g[j] = ok ? (gr[j] - m * sr_in) / s : 0.0;
        hxs[j] = ok ? (hx[j] - m * xmz) / s : 0.0;
        xv[j] = 0.0;
        if (!retry) Gbs[j] = b[j];
    }
    for (int c = tid; c < nS; c += T) sl[c] = slist[c];
    __syncthreads();
    bool fail = false;
    if (!retry) {
    // standardise the strong block of H in place: lstandard1's centring and scaling are affine in it
    for (int e = tid; e < nS * nS; e += T) {
        const int a = e / nS, c = e - a * nS;
        const int ja = sl[a], jc = sl[c];
        const double ma = Gxm[ja], mc = Gxm[jc];
        double *h = H + (size_t)hp[ja] * ldh + hp[jc];
        const double t = (*h - ma * hx[jc] - mc * hx[ja] + ma * mc * xmz) / (Gxs[ja] * Gxs[jc]);
        *h = t;
        if (in_smem) Hs[(size_t)a * nS + c] = t;
        if (a == c) {  // the curvature xv_j = sum v x_j^2 of the standardised column
            xv[ja] = t;
            Gxv[ja] = t;
        }
    }
    if (tid == 0) {
        sd[SD_SUMS] += (double)nS;
        sd[SD_SUMS2] += (double)nS * nS;
        sd[SD_SUMA] += (double)nin;
        sd[SD_BS0] = b0;
    }
    __syncthreads();
    double sr = sr_in;
    for (;;) {
        ++nlp;
        double dlx = 0.0;
        for (int k = 0; k < ni; ++k) {  // the strong set, in ascending column order as glmnet does
            if (!ixx[k]) continue;
            const double bk = b[k], gk = g[k], xk = xv[k];
            const double u = gk + xk * bk, au = fabs(u) - vq[k] * al1;
            const double bn = au <= 0.0 ? 0.0 : copysign(au, u) / xk;
            const double d = bn - bk;
            if (fabs(d) <= 0.0) continue;
            dlx = fmax(dlx, xk * d * d);
            const bool fresh = (mm[k] == 0);
            if (fresh) ++nin;
            sr -= d * hxs[k];
            __syncthreads();
            if (tid == 0) {
                b[k] = bn;
                if (fresh) {
                    mm[k] = nin;
                    ml[nin - 1] = k;
                }
            }
            if (in_smem) {
                const double *hr = Hs + (size_t)spos[k] * nS;
                for (int c = tid; c < nS; c += T) g[sl[c]] -= d * hr[c];
            } else {
                const double *hk = H + (size_t)hp[k] * ldh;
                for (int c = tid; c < nS; c += T) {
                    const int jj = sl[c];
                    g[jj] -= d * hk[hp[jj]];
                }
            }
            __syncthreads();
        }
        {
            const double d = sr / xmz;
            if (d != 0.0) {
                b0 += d;
                dlx = fmax(dlx, xmz * d * d);
                __syncthreads();
                for (int c = tid; c < nS; c += T) g[sl[c]] -= d * hxs[sl[c]];
                sr -= d * xmz;
                __syncthreads();
            }
        }
        if (dlx < shr) break;
        if (nlp > f.maxit) {
            fail = true;
            break;
        }
        for (;;) {
            ++nlp;
            dlx = 0.0;
            for (int l = 0; l < nin; ++l) {
                const int k = ml[l];
                const double bk = b[k], gk = g[k], xk = xv[k];
                const double u = gk + xk * bk, au = fabs(u) - vq[k] * al1;
                const double bn = au <= 0.0 ? 0.0 : copysign(au, u) / xk;
                const double d = bn - bk;
                if (fabs(d) <= 0.0) continue;
                dlx = fmax(dlx, xk * d * d);
                sr -= d * hxs[k];
                __syncthreads();
                if (tid == 0) b[k] = bn;
                if (in_smem) {
                    const double *hr = Hs + (size_t)spos[k] * nS;
                    for (int c = tid; c < nS; c += T) g[sl[c]] -= d * hr[c];
                } else {
                    const double *hk = H + (size_t)hp[k] * ldh;
                    for (int c = tid; c < nS; c += T) {
                        const int jj = sl[c];
                        g[jj] -= d * hk[hp[jj]];
                    }
                }
                __syncthreads();
            }
            const double d = sr / xmz;
            if (d != 0.0) {
                b0 += d;
                dlx = fmax(dlx, xmz * d * d);
                __syncthreads();
                for (int c = tid; c < nS; c += T) g[sl[c]] -= d * hxs[sl[c]];
                sr -= d * xmz;
                __syncthreads();
            }
            if (dlx < shr) break;
            if (nlp > f.maxit) {
                fail = true;
                break;
            }
        }
        if (fail) break;
    }
    }  // !retry
    // ---- coefficients for the next evaluation pass, on the raw x scale ----
    double part = 0.0;
    if (!retry)
        for (int j = tid; j < ni; j += T) {
            const double bj = (b[j] != 0.0) ? b[j] / Gxs[j] : 0.0;
            f.coef[j] = bj;
            part += bj * Gxm[j];
            Gb[j] = b[j];
            Gmm[j] = mm[j];
            Gml[j] = ml[j];
        }
    const double dotbm = bsum_all(part, red);
    // ---- this round's H is done with: the list lognet_pre_kernel built for the next step becomes current ----
    __syncthreads();
    if (si[SI_NEWLIST]) {
        const int *gn = fit_glist_next(f), *pn = fit_gpos_next(f);
        for (int c = tid; c < f.ldh; c += T) glist[c] = gn[c];
        for (int j = tid; j < ni; j += T) Ggpos[j] = pn[j];
    }
    __syncthreads();
    if (tid == 0) {
        si[SI_NEWLIST] = 0;
        if (!retry) {
            f.coef[ni] = b0 - dotbm;
            sd[SD_B0] = b0;
            si[SI_NIN] = nin;
            si[SI_NLP] = nlp;
            si[SI_ROUNDS] += 1;
            if (fail) {
                si[SI_JERR] = -(ilm + 1);
                si[SI_DONE] = 1;  // (the host reads it from lognet_pre_kernel's flags of the next round)
            }
        }
    }
}

// ---- held-out binomial deviance of a CV fit at every lambda ---------------------------------
// rows = the fit's hole (its held-out inner fold).  grid.y = chunks of LC lambdas (LC a multiple of 8).
// eta of eight rows at eight lambdas is one 8 x 8 DMMA accumulation over the columns: A = the rows (read
// straight from the design, one 32-byte sector per row and step), B = the coefficients of the chunk in shared
// memory (row stride nip = 4 mod 16 doubles).  A warp takes eight rows at a time; lane (g, k4) then holds
// eta of row g at lambdas 8 n + 2 k4 + {0, 1} and evaluates their deviance terms -- every lane busy, where
// the one-row-per-warp form spent a warp reduction per (row, lambda).
template <int NT>  // lambda tiles per chunk: LC = 8 NT (compile-time: a guarded mma.sync pays a re-convergence each)
__global__ void __launch_bounds__(256)
lognet_cvdev_kernel(const LognetFit *__restrict__ fits, int fid, int nip,
                    double *__restrict__ partials /*[gridDim.x][nlam]*/) {
    constexpr int LC = 8 * NT;
    const LognetFit f = fits[fid];
    const double *X = f.X;
    const long long ld = f.ld;
    extern __shared__ double sm[];  // beta chunk [LC][nip] + a0[LC] + red [warps][LC]
    const int ni = f.ni;
    const int lmu = fit_scal_i(f)[SI_LMU];
    const int m0 = blockIdx.y * LC;
    const int mc = max(0, min(LC, lmu - m0));
    double *sb = sm, *sa0 = sb + (size_t)LC * nip, *sred = sa0 + LC;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int g = lane >> 2, k4 = lane & 3;
    for (int e = threadIdx.x; e < LC * nip; e += blockDim.x) {
        const int m = e / nip, c = e - m * nip;
        sb[e] = (m < mc && c < ni) ? f.beta[(size_t)(m0 + m) * ni + c] : 0.0;  // zero beyond ni: the columns there (w, y, 1) drop out
    }
    for (int e = threadIdx.x; e < LC; e += blockDim.x) sa0[e] = e < mc ? f.a0[m0 + e] : 0.0;
    __syncthreads();
    const int nks = (ni + 3) >> 2;  // 4-column steps
    double dev[NT][2];
#pragma unroll
    for (int n = 0; n < NT; ++n) dev[n][0] = dev[n][1] = 0.0;
    const long long wstride = (long long)gridDim.x * nw * 8;
    for (long long t8 = ((long long)blockIdx.x * nw + wid) * 8; t8 < f.hole_len; t8 += wstride) {
        const long long t = t8 + g;
        const bool valid = t < f.hole_len;
        const long long pp = f.hole_begin + (valid ? t : t8);  // (a lane without a row re-reads the first one)
        const long long row = f.list ? (long long)__ldg(f.list + pp) : pp;
        const double *xr = X + (size_t)row * (size_t)ld;
        double c[NT][2];
#pragma unroll
        for (int n = 0; n < NT; ++n) c[n][0] = c[n][1] = 0.0;
        const double *bp = sb + (size_t)g * nip + k4;
#pragma unroll 2
        for (int kk = 0; kk < nks; ++kk) {
            const double a = __ldg(xr + 4 * kk + k4);
#pragma unroll
            for (int n = 0; n < NT; ++n) atk::dmma884(c[n][0], c[n][1], a, bp[(size_t)n * 8 * nip + 4 * kk]);
        }
        const double yi = __ldg(xr + f.iy);
#pragma unroll
        for (int n = 0; n < NT; ++n) {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int m = 8 * n + 2 * k4 + j;
                const double eta = sa0[m] + c[n][j];
                double pr = 1.0 / (1.0 + exp(-eta));
                pr = fmin(fmax(pr, 1.0e-5), 1.0 - 1.0e-5);
                const double dv = -2.0 * (yi * log(pr) + (1.0 - yi) * log(1.0 - pr));
                if (valid && m < mc) dev[n][j] += dv;
            }
        }
    }
    // lanes with the same k4 hold the same lambdas for different rows
#pragma unroll
    for (int n = 0; n < NT; ++n)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            double v = dev[n][j];
            v += __shfl_xor_sync(0xffffffffu, v, 4);
            v += __shfl_xor_sync(0xffffffffu, v, 8);
            v += __shfl_xor_sync(0xffffffffu, v, 16);
            const int m = 8 * n + 2 * k4 + j;
            if (g == 0) sred[(size_t)wid * LC + m] = v;
        }
    __syncthreads();
    for (int m = threadIdx.x; m < mc; m += blockDim.x) {
        double s = 0.0;
        for (int q = 0; q < nw; ++q) s += sred[(size_t)q * LC + m];
        partials[(size_t)blockIdx.x * f.nlam + m0 + m] = s;
    }
}

// ---- prediction on a row list ----------------------------------------------------------------
template <int ITER>
__global__ void __launch_bounds__(256)
predict_rows_kernel(const double *__restrict__ X, long long ld, int ni, const int32_t *__restrict__ rows,
                    long long row_begin, long long nrows, const double *__restrict__ beta, double a0,
                    double extra, int mode, int scatter, double *__restrict__ out0,
                    double *__restrict__ out1) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    double b[2 * ITER];
#pragma unroll
    for (int i = 0; i < ITER; ++i) {
        const int c = 2 * lane + 64 * i;
        b[2 * i] = c < ni ? beta[c] : 0.0;
        b[2 * i + 1] = c + 1 < ni ? beta[c + 1] : 0.0;
    }
    const long long wstride = (long long)gridDim.x * nw;
    for (long long t = (long long)blockIdx.x * nw + wid; t < nrows; t += wstride) {
        const long long row = rows ? (long long)__ldg(rows + row_begin + t) : row_begin + t;
        const double *xr = X + (size_t)row * (size_t)ld;
        double dot = 0.0;
#pragma unroll
        for (int i = 0; i < ITER; ++i) {
            const int c = 2 * lane + 64 * i;
            if (c + 1 < ld) {
                const double2 t2 = __ldg(reinterpret_cast<const double2 *>(xr + c));
                dot += t2.x * b[2 * i] + t2.y * b[2 * i + 1];
            }
        }
        const double eta = a0 + warp_sum(dot);
        if (lane == 0) {
            const long long o = scatter ? row : t;
            if (mode == 0)
                out0[o] = eta;
            else if (mode == 1)
                out0[o] = 1.0 / (1.0 + exp(-eta));
            else {
                out0[o] = eta;
                out1[o] = eta + extra;
            }
        }
    }
}

template <typename F>
int dispatch_iter(int ni, F &&fn) {
    const int it = (ni + 63) / 64;
    if (it <= 1) return fn(std::integral_constant<int, 1>());
    if (it <= 2) return fn(std::integral_constant<int, 2>());
    if (it <= 4) return fn(std::integral_constant<int, 4>());
    if (it <= 8) return fn(std::integral_constant<int, 8>());
    if (it <= 16) return fn(std::integral_constant<int, 16>());
    return -1;
}

}  // namespace

int ate_i_predict_rows(ate_ctx *ctx, const double *X, long long ld, int ni, const int32_t *rows,
                       long long row_begin, long long nrows, const double *d_beta, double a0, double extra,
                       int mode, int scatter, double *out0, double *out1) {
    if (nrows <= 0) return 0;
    int blocks = ctx->num_sms * 4;
    long long need = (nrows + 7) / 8;
    if (need < blocks) blocks = (int)need;
    int rc = dispatch_iter(ni, [&](auto it) {
        constexpr int IT = decltype(it)::value;
        predict_rows_kernel<IT><<<blocks, 256, 0, ctx->stream>>>(X, ld, ni, rows, row_begin, nrows, d_beta, a0,
                                                                 extra, mode, scatter, out0, out1);
        return 0;
    });
    if (rc < 0) return ate_fail(ctx, -1, "predict: ni = %d too large", ni);
    ATE_LAUNCHED(ctx);
    return 0;
}

// bump allocator over one cached device block: the per-call state of the driver costs no cudaMalloc
struct Carver {
    char *base = nullptr;
    size_t off = 0;
    template <typename T>
    T *take(size_t count) {
        off = (off + 255) / 256 * 256;
        T *p = base ? reinterpret_cast<T *>(base + off) : nullptr;
        off += sizeof(T) * (count ? count : 1);
        return p;
    }
};

int ate_i_cv_binomial(ate_ctx *ctx, const double *X, long long ld, int ni, int iy,
                      const double *vp_host, const PathOpts &po, const std::vector<CvProblem> &probs,
                      std::vector<CvOut> &outs, IrlsStats *stats) {
    const int P = (int)probs.size();
    if (P == 0) return 0;
    if (ni > 1024) return ate_fail(ctx, -1, "cv_binomial: ni = %d too large", ni);
    const int nlam = po.flmin >= 1.0 ? po.nulam : po.nlam;
    if (nlam < 1) return ate_fail(ctx, -1, "cv_binomial: need at least one lambda");
    const int F = probs[0].nseg;
    for (auto &pr : probs)
        if (pr.nseg != F) return ate_fail(ctx, -1, "cv_binomial: problems must share nfolds");
    const bool sharded = po.shard_count > 1;
    const int nlead = P, nfol = P * F, nfits = nlead + nfol;
    const int nsegp = F > 0 ? F : 1;
    const int plen = 3 * ni + SU_N;
    const int ldh = (ni + 31) / 32 * 32;
    // the step kernel stages [0, ncs) of every row: the features and the response column
    const int ncs = ((ni > iy ? ni : iy) + 1 + 7) / 8 * 8;
    if (ldh > ld || ncs > ld) return ate_fail(ctx, -1, "cv_binomial: design ld (%lld) too small for %d features", ld, ni);
    const size_t sd_len = 7 * (size_t)ni + SD_N, si_len = 7 * (size_t)ni + 3 * (size_t)ldh + SI_N;
    const size_t si_scal = 6 * (size_t)ni + 2 * (size_t)ldh;
    const size_t out_len = 3 * (size_t)nlam + (size_t)ni * nlam;
    const int ev_blocks = ctx->num_sms * 2;

    // ---- device state: one cached block, carved ----
    Carver cv;
    double *dcoef, *dstd, *dout, *dvp, *dulam, *dysum, *dpart, *dcv, *dseq;
    int *dsti, *dflags, *dnl;
    LognetFit *dfits;
    RoundDev *drdev;
    SegDesc *dsegs;
    for (int pass = 0; pass < 2; ++pass) {
        cv.off = 0;
        dcoef = cv.take<double>((size_t)(ni + 1) * nfits);
        dstd = cv.take<double>(sd_len * nfits);
        dout = cv.take<double>(out_len * nfits);
        dvp = cv.take<double>(ni);
        dulam = cv.take<double>(nlam);
        dseq = cv.take<double>((size_t)nlam * nlead);
        dysum = cv.take<double>((size_t)P * nsegp);
        dpart = cv.take<double>((size_t)ev_blocks * nlam);
        dcv = cv.take<double>((size_t)(nfol > 0 ? nfol : 1) * nlam);
        dsti = cv.take<int>(si_len * nfits);
        dflags = cv.take<int>(4 * (size_t)nfits);
        dnl = cv.take<int>(1);
        dfits = cv.take<LognetFit>(nfits);
        drdev = cv.take<RoundDev>(nfits);
        dsegs = cv.take<SegDesc>((size_t)P * nsegp);
        if (pass == 0) {
            void *blk = nullptr;
            ATE_TRY(ate_cached(ctx, "binom.state", cv.off + 256, &blk));
            cv.base = (char *)blk;
        }
    }
    ATE_CUDA(ctx, cudaMemsetAsync(dout, 0, sizeof(double) * out_len * nfits, ctx->stream));
    ATE_CUDA(ctx, cudaMemsetAsync(dflags, 0, sizeof(int) * 4 * nfits, ctx->stream));
    if (vp_host) ATE_CUDA(ctx, cudaMemcpyAsync(dvp, vp_host, sizeof(double) * ni, cudaMemcpyHostToDevice, ctx->stream));
    if (po.flmin >= 1.0) {
        ATE_CUDA(ctx, cudaMemcpyAsync(dulam, po.ulam, sizeof(double) * nlam, cudaMemcpyHostToDevice, ctx->stream));
        ATE_CUDA(ctx, cudaMemcpyAsync(dnl, &nlam, sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    }
    // Private contiguous copy of every problem's rows (in list order, pitch = staged width + 4): the
    // step kernel then stages a whole slot with one bulk copy instead of one per row, which is what
    // lets it reach HBM speed; ~100 passes per fit amortise the one gather.  Falls back to the shared
    // design + row lists when the copies would not fit comfortably.
    const long long ldp = ncs + 4;
    std::vector<double *> Xp(P, nullptr);
    {
        size_t need = 0, free_b = 0, total_b = 0;
        for (int p = 0; p < P; ++p) need += (size_t)probs[p].nt * (size_t)ldp;
        ATE_CUDA(ctx, cudaMemGetInfo(&free_b, &total_b));
        const char *off = getenv("ATE_NO_PRIVATE_COPY");
        const auto itc = ctx->cache.find("binom.rows");
        const size_t have = itc == ctx->cache.end() ? 0 : itc->second.second;  // already ours
        if (!(off && off[0] == '1') && need * sizeof(double) < (free_b + have) / 2) {
            void *pool = nullptr;
            ATE_TRY(ate_cached(ctx, "binom.rows", need * sizeof(double), &pool));
            size_t o = 0;
            for (int p = 0; p < P; ++p) {
                Xp[p] = (double *)pool + o;
                ATE_TRY(ate_i_gather_rows(ctx, X, ld, probs[p].d_list, probs[p].nt, ncs, Xp[p], ldp));
                o += (size_t)probs[p].nt * (size_t)ldp;
            }
        }
    }
    std::vector<RoundProblem> rprobs(P);
    for (int p = 0; p < P; ++p) {
        rprobs[p].X = Xp[p] ? Xp[p] : X;
        rprobs[p].ld = Xp[p] ? ldp : ld;
        rprobs[p].rows = Xp[p] ? nullptr : probs[p].d_list;
        rprobs[p].nrows = probs[p].nt;
        rprobs[p].padded = Xp[p] ? 1 : 0;
    }
    // fit ids: [0, nlead) the full-data fits, then problem-major the cv-fold fits
    std::vector<LognetFit> hf(nfits);
    std::vector<double> fit_nt(nfits), fit_hl_glob(nfits, 0.0);
    for (int id = 0; id < nfits; ++id) {
        const int p = id < nlead ? id : (id - nlead) / F;
        const int sg = id < nlead ? -1 : (id - nlead) % F;
        const CvProblem &pr = probs[p];
        const double ntg = (double)(pr.nt_global > 0 ? pr.nt_global : pr.nt);
        LognetFit &g = hf[id];
        g.X = rprobs[p].X;
        g.ld = rprobs[p].ld;
        g.list = rprobs[p].rows;
        g.hole_begin = sg < 0 ? 0 : pr.seg[sg];
        g.hole_len = sg < 0 ? 0 : pr.seg[sg + 1] - pr.seg[sg];
        if (sg >= 0) fit_hl_glob[id] = (double)(pr.seg_global.empty() ? g.hole_len : pr.seg_global[sg]);
        g.nt = ntg - fit_hl_glob[id];
        fit_nt[id] = g.nt;
        g.prob = p; g.seg = sg;
        g.ni = ni; g.iy = iy; g.ldh = ldh; g.ncs = ncs;
        g.vp = vp_host ? dvp : nullptr;
        g.nlam = nlam; g.thr = po.thr; g.maxit = po.maxit;
        g.coef = dcoef + (size_t)(ni + 1) * id;
        g.st_d = dstd + sd_len * id;
        g.st_i = dsti + si_len * id;
        double *o = dout + out_len * id;
        g.alm = o; g.a0 = o + nlam; g.dev = o + 2 * nlam; g.beta = o + 3 * nlam;
        if (id < nlead) {
            g.flmin = po.flmin;
            g.ulam = po.flmin >= 1.0 ? dulam : nullptr;
            g.nlam_ptr = po.flmin >= 1.0 ? dnl : nullptr;
            g.lead_done = nullptr;
            g.seq = dseq + (size_t)nlam * id;
        } else {
            g.flmin = 1.0;
            g.ulam = po.flmin >= 1.0 ? dulam : hf[p].seq;  // the caller's sequence, or the one the full-data fit publishes
            g.nlam_ptr = hf[p].st_i + si_scal + SI_LMU;
            g.lead_done = hf[p].st_i + si_scal + SI_DONE;
            g.seq = nullptr;
        }
        if (g.nt < 2.0) return ate_fail(ctx, -1, "cv_binomial: a fit with %g observations", g.nt);
    }
    ATE_CUDA(ctx, cudaMemcpyAsync(dfits, hf.data(), sizeof(LognetFit) * nfits, cudaMemcpyHostToDevice, ctx->stream));
    {
        std::vector<SegDesc> hs((size_t)P * nsegp);
        for (int p = 0; p < P; ++p)
            for (int sg = 0; sg < nsegp; ++sg) {
                SegDesc &d = hs[(size_t)p * nsegp + sg];
                d.X = rprobs[p].X; d.ld = rprobs[p].ld; d.list = rprobs[p].rows; d.iy = iy; d.pad = 0;
                d.begin = F > 0 ? probs[p].seg[sg] : 0;
                d.end = F > 0 ? probs[p].seg[sg + 1] : probs[p].nt;
            }
        ATE_CUDA(ctx, cudaMemcpyAsync(dsegs, hs.data(), sizeof(SegDesc) * hs.size(), cudaMemcpyHostToDevice, ctx->stream));
        lognet_ysum_kernel<<<P * nsegp, 256, 0, ctx->stream>>>(dsegs, dysum);
        ATE_LAUNCHED(ctx);
        if (sharded) ATE_TRY(ate_i_allreduce(ctx, dysum, (long long)P * nsegp));
        ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // hs is a host temporary
    }
    lognet_init_kernel<<<nfits, 256, 0, ctx->stream>>>(dfits, dysum, nsegp);
    ATE_LAUNCHED(ctx);

    const size_t pre_smem = sizeof(double) * ((size_t)ni + 32) + sizeof(int) * 2 * (size_t)ldh;
    const size_t cd_base = sizeof(double) * (5 * (size_t)ni + 32) + sizeof(int) * (5 * (size_t)ni + (size_t)ldh) + 32;
    if (cd_base > 200 * 1024) return ate_fail(ctx, -1, "cv_binomial: ni too large");
    // the rest of the SM's shared memory holds the standardised strong block (hs_cap x hs_cap doubles)
    int hs_cap = (int)sqrt((double)(224 * 1024 - cd_base) / sizeof(double));
    if (hs_cap > ldh) hs_cap = ldh;
    const size_t cd_smem = cd_base + sizeof(double) * (size_t)hs_cap * hs_cap;
    if (cd_smem > 48 * 1024)
        ATE_CUDA(ctx, cudaFuncSetAttribute(lognet_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cd_smem));
    int rc = 0;

    const bool timing = getenv("ATE_TIMING") != nullptr;
    auto tnow = []() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    double t_host = 0.0, t_wait = 0.0, t_mark = tnow();

    std::vector<int> hdone(nfits, 0), hng(nfits, 0), hlmu(nfits, 0), hones(nfits, -1), age(nfits, -1);
    void *hflags_v = nullptr, *hrdev_v = nullptr;
    ATE_TRY(ate_cached_pinned(ctx, "binom.hflags", sizeof(int) * 4 * (size_t)nfits, &hflags_v));
    ATE_TRY(ate_cached_pinned(ctx, "binom.hrdev", sizeof(RoundDev) * (size_t)nfits, &hrdev_v));
    int *hflags = (int *)hflags_v;
    RoundDev *hrdev = (RoundDev *)hrdev_v;
    std::vector<RoundFit> rfits(nfits);
    std::vector<RoundSlot> slots;
    struct EventGuard {
        cudaEvent_t e = nullptr;
        ~EventGuard() { if (e) cudaEventDestroy(e); }
    } evg;
    ATE_CUDA(ctx, cudaEventCreateWithFlags(&evg.e, cudaEventDisableTiming));
    cudaEvent_t ev_flags = evg.e;
    int rounds = 0;
    const int max_rounds = nlam * 60 + 100;
    RoundStats rst;
    // Lock-step rounds over ALL fits.  One round = step kernel for every live fit (evaluation + Gram over
    // the predicted lists) -> (ranks sharing the fit: all-reduce of the packed results) -> bookkeeping ->
    // coordinate-descent solve.  The full-data fit of a problem fixes the lambda sequence of its cv-fold
    // fits; the sequence is determined once three lambdas are recorded (lognet_pre_kernel publishes it
    // then), so the fold fits join a few rounds behind and share the CTAs of their full-data fit instead
    // of waiting for it to finish.  ATE_TWO_PHASE=1: the fold fits start when the full-data fit is done.
    static const bool two_phase = getenv("ATE_TWO_PHASE") && getenv("ATE_TWO_PHASE")[0] == '1';
    for (;;) {
        int live = 0, starting = 0;
        for (int id = 0; id < nfits; ++id) {
            RoundFit &r = rfits[id];
            r.live = 0;
            if (hdone[id]) continue;
            if (age[id] < 0) {  // not started yet
                bool go = id < nlead;
                if (!go) {
                    const int p = hf[id].prob;
                    go = hdone[p] || (!two_phase && (po.flmin >= 1.0 || hlmu[p] >= 3));
                }
                if (!go) continue;
                age[id] = 0;
            }
            r.live = 1;
            ++live;
            starting += (age[id] == 0);
            r.prob = hf[id].prob;
            r.hole_begin = hf[id].hole_begin;
            r.hole_len = hf[id].hole_len;
            r.w = 1.0 / fit_nt[id];
            r.coef = hf[id].coef;
            r.colidx = hf[id].st_i + 6 * (size_t)ni + ldh;  // the list the last solve left
            r.nG = hng[id];
            r.ones = hones[id];
        }
        if (!live) break;
        // the first step of a fit also needs sum v x^2 of every column: a launch of its own (other template
        // instantiation, other result bank) beside the launch of the fits that are under way
        for (int bank = 0; bank < 2; ++bank) {
            const int want_init = bank;  // bank 0: fits under way, bank 1: fits on their first step
            if ((want_init ? starting : live - starting) == 0) continue;
            std::vector<RoundFit> sel = rfits;
            for (int id = 0; id < nfits; ++id)
                if (sel[id].live && (age[id] == 0) != (want_init != 0)) sel[id].live = 0;
            double *packed = nullptr;
            long long count = 0;
            ATE_TRY(ate_i_irls_round(ctx, ncs, ni, iy, want_init, bank, rprobs, sel, slots, &packed, &count, &rst));
            for (int id = 0; id < nfits; ++id) {
                if (!sel[id].live) continue;
                hrdev[id].sums = packed + slots[id].sums_off;
                hrdev[id].H = packed + slots[id].H_off;
                hrdev[id].ldh = slots[id].ldh;
            }
            if (sharded && count > 0) ATE_TRY(ate_i_allreduce(ctx, packed, count));
            if (stats) {
                stats->step_launches += 1;
                stats->row_fits += (double)rst.row_fits;
                stats->rows_staged += (double)rst.rows_staged;
            }
        }
        for (int id = 0; id < nfits; ++id) hrdev[id].live = rfits[id].live;
        ATE_CUDA(ctx, cudaMemcpyAsync(drdev, hrdev, sizeof(RoundDev) * nfits, cudaMemcpyHostToDevice, ctx->stream));
        // the full-data fits first: what they record this round (lmu, done, the sequence) is what their
        // fold fits read in the second launch
        lognet_pre_kernel<<<nlead, CD_THREADS, pre_smem, ctx->stream>>>(dfits, drdev, dflags);
        ATE_LAUNCHED(ctx);
        if (nfol > 0) {
            lognet_pre_kernel<<<nfol, CD_THREADS, pre_smem, ctx->stream>>>(dfits + nlead, drdev + nlead, dflags + 4 * (size_t)nlead);
            ATE_LAUNCHED(ctx);
        }
        // what the host needs for the next launch (who is done, how long the next lists are) is known now: it
        // is read back while the solve runs, and the next round is planned and enqueued behind the solve
        ATE_CUDA(ctx, cudaMemcpyAsync(hflags, dflags, sizeof(int) * 4 * nfits, cudaMemcpyDeviceToHost, ctx->stream));
        ATE_CUDA(ctx, cudaEventRecord(ev_flags, ctx->stream));
        lognet_solve_kernel<<<nfits, CD_THREADS, cd_smem, ctx->stream>>>(dfits, drdev, hs_cap);
        ATE_LAUNCHED(ctx);
        if (timing) { const double t = tnow(); t_host += t - t_mark; t_mark = t; }
        ATE_CUDA(ctx, cudaEventSynchronize(ev_flags));
        if (timing) { const double t = tnow(); t_wait += t - t_mark; t_mark = t; }
        for (int id = 0; id < nfits; ++id) {
            if (!rfits[id].live) continue;
            hdone[id] = hflags[4 * (size_t)id];
            hng[id] = hflags[4 * (size_t)id + 1];
            hlmu[id] = hflags[4 * (size_t)id + 2];
            hones[id] = hflags[4 * (size_t)id + 3];
            ++age[id];
        }
        if (++rounds > max_rounds) return ate_fail(ctx, -2, "cv_binomial: IRLS round cap hit");
    }
    if (timing)
        fprintf(stderr, "[ate timing] cv_binomial: %d rounds, host %.1f ms (planning + launches), waiting on the GPU %.1f ms\n",
                rounds, t_host, t_wait);

    // ---- held-out deviance of every CV fold at every lambda (this rank's rows of the fold) ----
    std::vector<double> hcv;
    if (F > 0) {
        const int nip = (ni + 15) / 16 * 16 + 4;  // row stride of the coefficient chunk: 4 mod 16 doubles
        int LC = (int)((128 * 1024) / (sizeof(double) * (size_t)nip)) / 8 * 8;
        if (LC > 64) LC = 64;
        if (LC < 8) LC = 8;
        const int nchunk = (nlam + LC - 1) / LC;
        const size_t cv_smem = sizeof(double) * ((size_t)LC * nip + LC + 8 * (size_t)LC);
        if (cv_smem > 220 * 1024) return ate_fail(ctx, -1, "cv_binomial: ni = %d too large for the cv kernel", ni);
        auto cv_launch = [&](auto nt_tag, int id, bool size_only) -> int {
            constexpr int NTt = decltype(nt_tag)::value;
            if (size_only)
                return cudaFuncSetAttribute(lognet_cvdev_kernel<NTt>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cv_smem) == cudaSuccess ? 0 : -1;
            lognet_cvdev_kernel<NTt><<<dim3(ev_blocks, nchunk), 256, cv_smem, ctx->stream>>>(dfits, id, nip, dpart);
            return 0;
        };
        auto cv_dispatch = [&](int id, bool size_only) -> int {
            switch (LC / 8) {
                case 1: return cv_launch(std::integral_constant<int, 1>{}, id, size_only);
                case 2: return cv_launch(std::integral_constant<int, 2>{}, id, size_only);
                case 3: return cv_launch(std::integral_constant<int, 3>{}, id, size_only);
                case 4: return cv_launch(std::integral_constant<int, 4>{}, id, size_only);
                case 5: return cv_launch(std::integral_constant<int, 5>{}, id, size_only);
                case 6: return cv_launch(std::integral_constant<int, 6>{}, id, size_only);
                case 7: return cv_launch(std::integral_constant<int, 7>{}, id, size_only);
                default: return cv_launch(std::integral_constant<int, 8>{}, id, size_only);
            }
        };
        if (cv_dispatch(0, true) != 0) return ate_fail(ctx, -1, "cv_binomial: cannot size cv kernel");
        ATE_CUDA(ctx, cudaMemsetAsync(dcv, 0, sizeof(double) * (size_t)nfol * nlam, ctx->stream));
        for (int q = 0; q < nfol; ++q) {
            const int id = nlead + q;
            ATE_CUDA(ctx, cudaMemsetAsync(dpart, 0, sizeof(double) * (size_t)ev_blocks * nlam, ctx->stream));
            cv_dispatch(id, false);
            ATE_LAUNCHED(ctx);
            fold_partials_kernel<<<(nlam + 7) / 8, 256, 0, ctx->stream>>>(dpart, ev_blocks, nlam, dcv + (size_t)q * nlam);
            ATE_LAUNCHED(ctx);
        }
        if (sharded) ATE_TRY(ate_i_allreduce(ctx, dcv, (long long)nfol * nlam));
        hcv.resize((size_t)nfol * nlam);
        ATE_CUDA(ctx, cudaMemcpyAsync(hcv.data(), dcv, sizeof(double) * hcv.size(), cudaMemcpyDeviceToHost, ctx->stream));
    }
    std::vector<double> hout(out_len * (size_t)nlead), hsd(sd_len * (size_t)nfits);
    std::vector<int> hsi(si_len * (size_t)nfits);
    ATE_CUDA(ctx, cudaMemcpyAsync(hout.data(), dout, sizeof(double) * hout.size(), cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaMemcpyAsync(hsd.data(), dstd, sizeof(double) * hsd.size(), cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaMemcpyAsync(hsi.data(), dsti, sizeof(int) * hsi.size(), cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));

    if (stats) {
        stats->rounds = rounds;
        for (int id = 0; id < nfits; ++id) {
            const double *dv = hsd.data() + sd_len * (size_t)id + 7 * (size_t)ni;
            const int *iv = hsi.data() + si_len * (size_t)id + si_scal;
            stats->solves += iv[SI_ROUNDS];
            stats->sum_strong += dv[SD_SUMS];
            stats->sum_strong2 += dv[SD_SUMS2];
            stats->list_misses += iv[SI_MISS];
        }
    }

    outs.assign(P, CvOut());
    for (int p = 0; p < P; ++p) {
        CvOut &o = outs[p];
        const int *iv = hsi.data() + si_len * (size_t)p + si_scal;
        const double *dv = hsd.data() + sd_len * (size_t)p + 7 * (size_t)ni;
        o.lmu = iv[SI_LMU];
        o.nlp = iv[SI_NLP];
        o.jerr = iv[SI_JERR];
        const double *ob = hout.data() + out_len * (size_t)p;
        o.lambda.assign(ob, ob + o.lmu);
        o.a0.assign(ob + nlam, ob + nlam + o.lmu);
        o.dev.assign(ob + 2 * nlam, ob + 2 * nlam + o.lmu);
        o.beta.assign(ob + 3 * nlam, ob + 3 * nlam + (size_t)ni * o.lmu);
        o.nulldev = 2.0 * fit_nt[p] * dv[SD_DEV0];
        if (F > 0 && o.lmu > 0) {
            std::vector<double> cvk((size_t)F * o.lmu), wk(F);
            std::vector<int> nlk(F);
            for (int f = 0; f < F; ++f) {
                const int id = nlead + p * F + f;
                const int *fv = hsi.data() + si_len * (size_t)id + si_scal;
                nlk[f] = fv[SI_LMU] < o.lmu ? fv[SI_LMU] : o.lmu;
                if (fv[SI_JERR] != 0 && o.jerr == 0) o.jerr = fv[SI_JERR];
                o.nlp += fv[SI_NLP];
                wk[f] = fit_hl_glob[id];
                for (int l = 0; l < o.lmu; ++l)
                    cvk[(size_t)f * o.lmu + l] = hcv[((size_t)p * F + f) * nlam + l] / wk[f];
            }
            ate_host_cv_summary(cvk, wk, nlk, F, o.lmu, o.lambda.data(), o.cvm, o.cvsd, &o.idx_min, &o.idx_1se);
        }
    }
    return 0;
}
