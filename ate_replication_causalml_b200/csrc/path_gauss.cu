// path_gauss.cu -- a3: gaussian lasso path + cv.glmnet (ate_functions.R:101-103,123-125,304-305)
// solved entirely in Gram space.
//
// glmnet's covariance-updating coordinate descent (type.gaussian = "covariance") touches the
// data only through <x_j, x_k> and <x_j, y> on standardised columns.  Those are affine
// functions of the raw moment matrix S = [X y 1]^T [X y 1] of the fit's rows, and the moment
// matrix of "all training rows but inner fold f" is S_total - S_f.  So one DMMA Gram pass per
// inner fold (gram.cu) feeds the full-data fit, all nfolds CV fits AND the held-out squared
// errors (a quadratic form in S_f) -- no further pass over the n rows is needed.
//
//   gauss_prep_kernel : S -> means, 1/N sds, standardised C (ni x ni) and g = <x_j, y>
//   gauss_path_kernel : one CTA per fit walks the lambda path: cyclic CD over the strong set,
//                       active-set inner loop, KKT re-check, early stopping -- the same loop
//                       order as glmnet's elnet1, so the iterates agree to rounding.
//                       Scalars are computed redundantly by every thread (uniform control
//                       flow); the length-ni gradient update is spread over the CTA.
//   gauss_cv_kernel   : held-out MSE of every (fold, lambda) from S_f.
#include <math.h>

#include "path.cuh"

namespace {

struct GaussFit {
    const double *S_total, *S_sub;  // mS x mS raw moments; S = S_total - S_sub (S_sub may be NULL)
    int mS, ni, iy, i1;
    const double *vp;  // ni raw penalty factors
    double flmin;      // < 1: automatic sequence
    const double *ulam;
    const int *nlam_ptr;  // device count when ulam is another fit's output
    int nlam;
    double thr;
    int maxit;
    double *C, *g, *xm, *xs, *scal;  // workspace: scal = {ym, ys, nt}
    int *ju;
    int *lmu, *nlp, *jerr;  // outputs
    double *alm, *rsq, *a0, *beta;
};

constexpr int PATH_THREADS = 256;
constexpr int PREP_SPLIT = 8;

__device__ __forceinline__ double S_at(const GaussFit &f, int r, int c) {
    const size_t o = (size_t)r * f.mS + c;
    return f.S_sub ? f.S_total[o] - f.S_sub[o] : f.S_total[o];
}

__global__ void __launch_bounds__(PATH_THREADS)
gauss_prep_kernel(const GaussFit *__restrict__ fits) {
    const GaussFit f = fits[blockIdx.x];
    extern __shared__ double sm[];  // xm[ni], xs[ni]
    double *xm = sm, *xs = sm + f.ni;
    const double nt = S_at(f, f.i1, f.i1);
    const double ym = S_at(f, f.iy, f.i1) / nt;
    const double ys = sqrt(S_at(f, f.iy, f.iy) / nt - ym * ym);
    for (int j = threadIdx.x; j < f.ni; j += blockDim.x) {
        const double m = S_at(f, j, f.i1) / nt;
        const double sq = S_at(f, j, j) / nt;
        const double var = sq - m * m;
        const bool ok = var > 1.0e-14 * (sq > 1.0e-300 ? sq : 1.0e-300);
        xm[j] = m;
        xs[j] = ok ? sqrt(var) : 1.0;
        if (blockIdx.y == 0) {
            f.xm[j] = m;
            f.xs[j] = xs[j];
            f.ju[j] = ok ? 1 : 0;
            f.g[j] = ok ? (S_at(f, j, f.iy) / nt - m * ym) / (xs[j] * ys) : 0.0;
        }
    }
    if (blockIdx.y == 0 && threadIdx.x == 0) {
        f.scal[0] = ym;
        f.scal[1] = ys;
        f.scal[2] = nt;
    }
    __syncthreads();
    const int rows_per = (f.ni + gridDim.y - 1) / gridDim.y;
    const int r0 = blockIdx.y * rows_per, r1 = min(f.ni, r0 + rows_per);
    for (int r = r0; r < r1; ++r)
        for (int c = threadIdx.x; c < f.ni; c += blockDim.x)
            f.C[(size_t)r * f.ni + c] = (S_at(f, r, c) / nt - xm[r] * xm[c]) / (xs[r] * xs[c]);
}

// block-wide max, result broadcast to every thread
__device__ __forceinline__ double block_max_all(double v, double *red) {
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
__device__ __forceinline__ double block_sum_all(double v, double *red) {
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

__global__ void __launch_bounds__(PATH_THREADS)
gauss_path_kernel(const GaussFit *__restrict__ fits) {
    const GaussFit f = fits[blockIdx.x];
    const int ni = f.ni, tid = threadIdx.x, T = blockDim.x;
    extern __shared__ double sm[];
    double *a = sm, *g = a + ni, *da = g + ni, *vq = da + ni, *red = vq + ni;  // red[32]
    int *ix = reinterpret_cast<int *>(red + 32), *mm = ix + ni, *ia = mm + ni, *ju = ia + ni;
    const double *C = f.C;

    // ---- set-up: penalty factors rescaled to sum ni, state cleared ----
    double vsum = 0.0;
    for (int j = tid; j < ni; j += T) {
        const double t = f.vp ? fmax(f.vp[j], 0.0) : 1.0;
        vq[j] = t;
        vsum += t;
        a[j] = 0.0;
        g[j] = f.g[j];
        da[j] = 0.0;
        ix[j] = 0;
        mm[j] = 0;
        ia[j] = 0;
        ju[j] = f.ju[j];
    }
    vsum = block_sum_all(vsum, red);
    for (int j = tid; j < ni; j += T) vq[j] = vq[j] * (double)ni / vsum;
    __syncthreads();

    const double ym = f.scal[0], ys = f.scal[1];
    const int nlam = (f.flmin < 1.0) ? f.nlam : min(f.nlam, *f.nlam_ptr);
    const int mnl = min(GLM_MNLAM, nlam);
    double alf = 1.0;
    if (f.flmin < 1.0) alf = pow(fmax(GLM_EPS, f.flmin), 1.0 / (double)(nlam - 1));
    double alm = 0.0, rsq = 0.0;
    int nin = 0, iz = 0, nlp = 0, lmu = 0, jerr = 0;
    const double thr = f.thr;

    for (int m = 0; m < nlam; ++m) {
        const double alm0_in = alm;
        double alm0 = alm0_in;
        if (f.flmin >= 1.0)
            alm = f.ulam[m] / ys;
        else if (m > 1)
            alm *= alf;
        else if (m == 0)
            alm = GLM_BIG;
        else {
            double t = 0.0;
            for (int j = tid; j < ni; j += T)
                if (ju[j] && vq[j] > 0.0) t = fmax(t, fabs(g[j]) / vq[j]);
            alm0 = block_max_all(t, red);  // alpha = 1: no division
            alm = alf * alm0;
        }
        const double ab = alm;  // alm * alpha
        const double rsq0 = rsq;
        int jz = 1;
        const double tlam = 2.0 * alm - alm0;
        for (int j = tid; j < ni; j += T)
            if (!ix[j] && ju[j] && fabs(g[j]) > tlam * vq[j]) ix[j] = 1;
        __syncthreads();
        bool fail = false;
        for (;;) {
            if (!(iz && jz)) {
                bool converged = false;
                for (;;) {  // full sweeps over the strong set
                    ++nlp;
                    double dlx = 0.0;
                    for (int k = 0; k < ni; ++k) {
                        if (!ix[k]) continue;
                        const double ak = a[k], gk = g[k];
                        const double u = gk + ak, v = fabs(u) - vq[k] * ab;
                        const double an = v > 0.0 ? copysign(v, u) : 0.0;
                        if (an == ak) continue;
                        const bool fresh = (mm[k] == 0);
                        if (fresh) ++nin;
                        __syncthreads();  // everyone has read a[k], g[k], mm[k]
                        if (tid == 0) {
                            a[k] = an;
                            if (fresh) {
                                mm[k] = nin;
                                ia[nin - 1] = k;
                            }
                        }
                        const double del = an - ak;
                        rsq += del * (2.0 * gk - del);
                        dlx = fmax(dlx, del * del);
                        const double *ck = C + (size_t)k * ni;
                        for (int j = tid; j < ni; j += T)
                            if (ju[j]) g[j] -= ck[j] * del;
                        __syncthreads();
                    }
                    __syncthreads();  // nobody is still scanning ix[] when the KKT pass edits it
                    if (dlx < thr) {
                        int add = 0;
                        for (int j = tid; j < ni; j += T)
                            if (!ix[j] && ju[j] && fabs(g[j]) > ab * vq[j]) {
                                ix[j] = 1;
                                add = 1;
                            }
                        if (__syncthreads_or(add)) continue;
                        converged = true;
                    }
                    break;
                }
                if (converged) break;
                if (nlp > f.maxit) {
                    fail = true;
                    break;
                }
            }
            iz = 1;
            for (int l = tid; l < nin; l += T) da[l] = a[ia[l]];
            __syncthreads();
            for (;;) {  // active-set sweeps
                ++nlp;
                double dlx = 0.0;
                for (int l = 0; l < nin; ++l) {
                    const int k = ia[l];
                    const double ak = a[k], gk = g[k];
                    const double u = gk + ak, v = fabs(u) - vq[k] * ab;
                    const double an = v > 0.0 ? copysign(v, u) : 0.0;
                    if (an == ak) continue;
                    __syncthreads();
                    if (tid == 0) a[k] = an;
                    const double del = an - ak;
                    rsq += del * (2.0 * gk - del);
                    dlx = fmax(dlx, del * del);
                    const double *ck = C + (size_t)k * ni;
                    for (int j = tid; j < nin; j += T) {
                        const int jj = ia[j];
                        g[jj] -= ck[jj] * del;
                    }
                    __syncthreads();
                }
                __syncthreads();
                if (dlx < thr) break;
                if (nlp > f.maxit) {
                    fail = true;
                    break;
                }
            }
            if (fail) break;
            for (int l = tid; l < nin; l += T) da[l] = a[ia[l]] - da[l];
            __syncthreads();
            for (int j = tid; j < ni; j += T) {
                if (mm[j] != 0 || !ju[j]) continue;
                double s = 0.0;
                for (int l = 0; l < nin; ++l) s += da[l] * C[(size_t)ia[l] * ni + j];
                g[j] -= s;
            }
            __syncthreads();
            jz = 0;
        }
        if (fail) {
            jerr = -(m + 1);
            break;
        }
        // ---- record lambda m on the caller's scale ----
        double part = 0.0;
        int nz = 0;
        for (int j = tid; j < ni; j += T) {
            const double b = (a[j] != 0.0) ? ys * a[j] / f.xs[j] : 0.0;
            f.beta[(size_t)m * ni + j] = b;
            part += b * f.xm[j];
            nz += (a[j] != 0.0);
        }
        const double dotbm = block_sum_all(part, red);
        const int me = (int)block_sum_all((double)nz, red);
        if (tid == 0) {
            f.a0[m] = ym - dotbm;
            f.alm[m] = alm * ys;
            f.rsq[m] = rsq;
        }
        lmu = m + 1;
        if (m + 1 < mnl) continue;
        if (f.flmin >= 1.0) continue;
        if (me > ni + 1) break;
        if (rsq - rsq0 < GLM_FDEV * rsq) break;
        if (rsq > GLM_DEVMAX) break;
    }
    __syncthreads();
    if (tid == 0) {
        *f.lmu = lmu;
        *f.nlp = nlp;
        *f.jerr = jerr;
        // fix.lam: the first lambda was `big`; R replaces it by log-linear extrapolation
        if (f.flmin < 1.0 && lmu > 2) f.alm[0] = exp(2.0 * log(f.alm[1]) - log(f.alm[2]));
    }
}

// held-out mean squared error of fit `fid` at lambda blockIdx.y, from the fold's own moments
__global__ void __launch_bounds__(128)
gauss_cv_kernel(const GaussFit *__restrict__ fits, int first_fit, int lmu_cap,
                double *__restrict__ cv_out /*[nfits][lmu_cap]*/) {
    const GaussFit f = fits[first_fit + blockIdx.x];
    const int m = blockIdx.y, ni = f.ni;
    __shared__ double red[32];
    if (m >= *f.lmu) {
        if (threadIdx.x == 0) cv_out[(size_t)blockIdx.x * lmu_cap + m] = NAN;
        return;
    }
    const double *S = f.S_sub;  // moments of the held-out rows
    const double *b = f.beta + (size_t)m * ni;
    const double a0 = f.a0[m];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    double acc = 0.0;
    for (int j = wid; j < ni; j += nw) {
        const double bj = b[j];
        if (bj == 0.0) continue;
        const double *Sj = S + (size_t)j * f.mS;
        double q = 0.0;
        for (int k = lane; k < ni; k += 32) q += b[k] * Sj[k];
        q = warp_sum(q);
        if (lane == 0) acc += bj * q + 2.0 * bj * (a0 * Sj[f.i1] - Sj[f.iy]);
    }
    const double tot = block_sum_all(acc, red);
    if (threadIdx.x == 0) {
        const double nf = S[(size_t)f.i1 * f.mS + f.i1];
        const double syy = S[(size_t)f.iy * f.mS + f.iy], sy1 = S[(size_t)f.iy * f.mS + f.i1];
        const double sse = syy - 2.0 * a0 * sy1 + a0 * a0 * nf + tot;
        cv_out[(size_t)blockIdx.x * lmu_cap + m] = sse / nf;
    }
}

__global__ void sum_mats_kernel(const double *__restrict__ parts, int nparts, size_t len,
                                double *__restrict__ out) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < len; e += stride) {
        double s = 0.0;
        for (int q = 0; q < nparts; ++q) s += parts[(size_t)q * len + e];
        out[e] = s;
    }
}

}  // namespace

void ate_host_cv_summary(const std::vector<double> &cvk, const std::vector<double> &wk,
                         const std::vector<int> &nlk, int nfolds, int lmu, const double *lambda,
                         std::vector<double> &cvm, std::vector<double> &cvsd, int *idx_min,
                         int *idx_1se) {
    cvm.assign(lmu, 0.0);
    cvsd.assign(lmu, 0.0);
    for (int l = 0; l < lmu; ++l) {
        double sw = 0.0, sm = 0.0;
        int nn = 0;
        for (int f = 0; f < nfolds; ++f)
            if (l < nlk[f]) {
                sw += wk[f];
                sm += wk[f] * cvk[(size_t)f * lmu + l];
                ++nn;
            }
        const double mean = sm / sw;
        double sv = 0.0;
        for (int f = 0; f < nfolds; ++f)
            if (l < nlk[f]) {
                const double d = cvk[(size_t)f * lmu + l] - mean;
                sv += wk[f] * d * d;
            }
        cvm[l] = mean;
        cvsd[l] = sqrt((sv / sw) / (double)(nn - 1));
    }
    // getmin(): largest lambda attaining the minimum; largest lambda within one se of it
    double best = cvm[0];
    for (int l = 1; l < lmu; ++l) best = cvm[l] < best ? cvm[l] : best;
    int imin = -1;
    for (int l = 0; l < lmu; ++l)
        if (cvm[l] <= best && (imin < 0 || lambda[l] > lambda[imin])) imin = l;
    const double bar = cvm[imin] + cvsd[imin];
    int i1 = -1;
    for (int l = 0; l < lmu; ++l)
        if (cvm[l] <= bar && (i1 < 0 || lambda[l] > lambda[i1])) i1 = l;
    *idx_min = imin;
    *idx_1se = i1;
}

extern "C" int ate_cv_select(int nfolds, int lmu, const double *cvraw, const double *fold_weights, const int *fold_nlam,
                             const double *lambda, double *cvm, double *cvsd, int *idx_min, int *idx_1se) {
    if (nfolds < 2 || lmu < 1 || !cvraw || !fold_weights || !lambda || !cvm || !cvsd || !idx_min || !idx_1se) return -1;
    std::vector<double> cvk(cvraw, cvraw + (size_t)nfolds * lmu), wk(fold_weights, fold_weights + nfolds), m, sd;
    std::vector<int> nlk(nfolds, lmu);
    if (fold_nlam)
        for (int f = 0; f < nfolds; ++f) nlk[f] = fold_nlam[f] < lmu ? fold_nlam[f] : lmu;
    ate_host_cv_summary(cvk, wk, nlk, nfolds, lmu, lambda, m, sd, idx_min, idx_1se);
    for (int l = 0; l < lmu; ++l) {
        cvm[l] = m[l];
        cvsd[l] = sd[l];
    }
    return 0;
}

int ate_i_cv_gaussian(ate_ctx *ctx, const double *X, long long ld, int ni, int iy, int i1,
                      const double *vp_host, const PathOpts &po, const std::vector<CvProblem> &probs,
                      std::vector<CvOut> &outs) {
    const int P = (int)probs.size();
    if (P == 0) return 0;
    const int mS = i1 + 1;
    if (iy >= mS || ni > iy) return ate_fail(ctx, -1, "cv_gaussian: bad column layout");
    const int nlam = po.flmin >= 1.0 ? po.nulam : po.nlam;
    if (nlam < 1) return ate_fail(ctx, -1, "cv_gaussian: need at least one lambda");
    int F = probs[0].nseg;
    for (auto &pr : probs)
        if (pr.nseg != F) return ate_fail(ctx, -1, "cv_gaussian: problems must share nfolds");
    const int nlead = P, nfol = P * F, nfits = nlead + nfol;
    const size_t m2 = (size_t)mS * mS;

    // ---- device state: one block cached in the ctx, carved (a second call of the same shape allocates
    //      nothing: with one process per GPU, cudaMalloc / cudaFree of eight processes serialise in the
    //      driver and showed up as tens of ms of the outcome stage at 8 GPUs) ----
    const size_t n2 = (size_t)ni * ni;
    struct Ptr { double *p; };
    struct IPtr { int *p; };
    Ptr dSf{nullptr}, dStot{nullptr}, dC{nullptr}, dvec{nullptr}, dout{nullptr}, dvp{nullptr}, dulam{nullptr}, dcv{nullptr};
    IPtr dint{nullptr}, dnl{nullptr};
    struct { GaussFit *p; } dfits{nullptr};
    {
        char *base = nullptr;
        for (int pass = 0; pass < 2; ++pass) {
            size_t off = 0;
            auto take = [&](size_t bytes) {
                off = (off + 255) / 256 * 256;
                char *q = base ? base + off : nullptr;
                off += bytes ? bytes : 8;
                return q;
            };
            dStot.p = (double *)take(sizeof(double) * m2 * P);
            dSf.p = (double *)take(sizeof(double) * m2 * (size_t)(F > 0 ? nfol : 0));
            dC.p = (double *)take(sizeof(double) * n2 * nfits);
            dvec.p = (double *)take(sizeof(double) * (size_t)nfits * (3 * (size_t)ni + 4));                       // g, xm, xs, scal
            dout.p = (double *)take(sizeof(double) * (size_t)nfits * (3 * (size_t)nlam + (size_t)ni * nlam));     // alm, rsq, a0, beta
            dvp.p = (double *)take(sizeof(double) * ni);
            dulam.p = (double *)take(sizeof(double) * nlam);
            dcv.p = (double *)take(sizeof(double) * (size_t)(F > 0 ? nfol : 0) * nlam);
            dint.p = (int *)take(sizeof(int) * (size_t)nfits * ((size_t)ni + 4));                                 // ju, lmu, nlp, jerr
            dnl.p = (int *)take(sizeof(int));
            dfits.p = (GaussFit *)take(sizeof(GaussFit) * nfits);
            if (pass == 0) {
                void *blk = nullptr;
                ATE_TRY(ate_cached(ctx, "gauss.state", off + 256, &blk));
                base = (char *)blk;
            }
        }
    }

    // ---- moment matrices: one Gram per (problem, inner fold), then totals ----
    if (F > 0) {
        for (int p = 0; p < P; ++p)
            for (int f = 0; f < F; ++f)
                ATE_TRY(ate_i_gram(ctx, X, ld, mS, probs[p].d_list, probs[p].seg[f],
                                   probs[p].seg[f + 1] - probs[p].seg[f], nullptr,
                                   dSf.p + m2 * ((size_t)p * F + f)));
        // ranks sharing the fit hold row slices: the moment matrices are additive over rows
        if (po.shard_count > 1) ATE_TRY(ate_i_allreduce(ctx, dSf.p, (long long)(m2 * (size_t)nfol)));
        for (int p = 0; p < P; ++p) {
            sum_mats_kernel<<<ctx->num_sms, 256, 0, ctx->stream>>>(dSf.p + m2 * (size_t)p * F, F, m2, dStot.p + m2 * p);
            ATE_LAUNCHED(ctx);
        }
    } else {
        for (int p = 0; p < P; ++p)
            ATE_TRY(ate_i_gram(ctx, X, ld, mS, probs[p].d_list, 0, probs[p].nt, nullptr, dStot.p + m2 * p));
        if (po.shard_count > 1) ATE_TRY(ate_i_allreduce(ctx, dStot.p, (long long)(m2 * (size_t)P)));
    }

    // ---- per-fit workspace and outputs ----
    if (vp_host) ATE_CUDA(ctx, cudaMemcpyAsync(dvp.p, vp_host, sizeof(double) * ni, cudaMemcpyHostToDevice, ctx->stream));
    if (po.flmin >= 1.0) ATE_CUDA(ctx, cudaMemcpyAsync(dulam.p, po.ulam, sizeof(double) * nlam, cudaMemcpyHostToDevice, ctx->stream));
    std::vector<GaussFit> hf(nfits);
    auto fill = [&](int id, const double *St, const double *Ss) {
        GaussFit &g = hf[id];
        g.S_total = St;
        g.S_sub = Ss;
        g.mS = mS; g.ni = ni; g.iy = iy; g.i1 = i1;
        g.vp = vp_host ? dvp.p : nullptr;
        g.nlam = nlam; g.thr = po.thr; g.maxit = po.maxit;
        g.C = dC.p + n2 * id;
        double *v = dvec.p + (size_t)id * (3 * (size_t)ni + 4);
        g.g = v; g.xm = v + ni; g.xs = v + 2 * ni; g.scal = v + 3 * ni;
        int *iv = dint.p + (size_t)id * ((size_t)ni + 4);
        g.ju = iv; g.lmu = iv + ni; g.nlp = iv + ni + 1; g.jerr = iv + ni + 2;
        double *o = dout.p + (size_t)id * (3 * (size_t)nlam + (size_t)ni * nlam);
        g.alm = o; g.rsq = o + nlam; g.a0 = o + 2 * nlam; g.beta = o + 3 * nlam;
    };
    for (int p = 0; p < P; ++p) {
        fill(p, dStot.p + m2 * p, nullptr);
        hf[p].flmin = po.flmin;
        hf[p].ulam = po.flmin >= 1.0 ? dulam.p : nullptr;
        hf[p].nlam_ptr = nullptr;
        if (po.flmin >= 1.0) {  // user lambdas: nlam fixed
            hf[p].nlam_ptr = nullptr;
        }
    }
    for (int p = 0; p < P; ++p)
        for (int f = 0; f < F; ++f) {
            const int id = nlead + p * F + f;
            fill(id, dStot.p + m2 * p, dSf.p + m2 * ((size_t)p * F + f));
            hf[id].flmin = 1.0;  // cv folds run on the full fit's lambda sequence
            hf[id].ulam = hf[p].alm;
            hf[id].nlam_ptr = hf[p].lmu;
        }
    // leaders with user lambdas need a device count too
    if (po.flmin >= 1.0) {
        ATE_CUDA(ctx, cudaMemcpyAsync(dnl.p, &nlam, sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
        for (int p = 0; p < P; ++p) hf[p].nlam_ptr = dnl.p;
    }
    ATE_CUDA(ctx, cudaMemcpyAsync(dfits.p, hf.data(), sizeof(GaussFit) * nfits, cudaMemcpyHostToDevice, ctx->stream));

    const size_t prep_smem = sizeof(double) * 2 * (size_t)ni;
    const size_t path_smem = sizeof(double) * (4 * (size_t)ni + 32) + sizeof(int) * 4 * (size_t)ni;
    if (path_smem > 200 * 1024) return ate_fail(ctx, -1, "cv_gaussian: ni = %d too large", ni);
    if (path_smem > 48 * 1024)
        ATE_CUDA(ctx, cudaFuncSetAttribute(gauss_path_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)path_smem));
    gauss_prep_kernel<<<dim3(nfits, PREP_SPLIT), PATH_THREADS, prep_smem, ctx->stream>>>(dfits.p);
    ATE_LAUNCHED(ctx);
    gauss_path_kernel<<<nlead, PATH_THREADS, path_smem, ctx->stream>>>(dfits.p);
    ATE_LAUNCHED(ctx);
    if (F > 0) {
        gauss_path_kernel<<<nfol, PATH_THREADS, path_smem, ctx->stream>>>(dfits.p + nlead);
        ATE_LAUNCHED(ctx);
        gauss_cv_kernel<<<dim3(nfol, nlam), 128, 0, ctx->stream>>>(dfits.p, nlead, nlam, dcv.p);
        ATE_LAUNCHED(ctx);
    }

    // ---- results back ----
    std::vector<double> hout((size_t)nlead * (3 * (size_t)nlam + (size_t)ni * nlam));
    std::vector<int> hint((size_t)nfits * ((size_t)ni + 4));
    std::vector<double> hcv((size_t)nfol * nlam), hscal((size_t)nfits * (3 * (size_t)ni + 4));
    ATE_CUDA(ctx, cudaMemcpyAsync(hout.data(), dout.p, sizeof(double) * hout.size(), cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaMemcpyAsync(hint.data(), dint.p, sizeof(int) * hint.size(), cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaMemcpyAsync(hscal.data(), dvec.p, sizeof(double) * hscal.size(), cudaMemcpyDeviceToHost, ctx->stream));
    if (F > 0) ATE_CUDA(ctx, cudaMemcpyAsync(hcv.data(), dcv.p, sizeof(double) * hcv.size(), cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));

    outs.assign(P, CvOut());
    for (int p = 0; p < P; ++p) {
        CvOut &o = outs[p];
        const int *iv = hint.data() + (size_t)p * ((size_t)ni + 4);
        o.lmu = iv[ni];
        o.nlp = iv[ni + 1];
        o.jerr = iv[ni + 2];
        const double *ob = hout.data() + (size_t)p * (3 * (size_t)nlam + (size_t)ni * nlam);
        o.lambda.assign(ob, ob + o.lmu);
        o.dev.assign(ob + nlam, ob + nlam + o.lmu);
        o.a0.assign(ob + 2 * nlam, ob + 2 * nlam + o.lmu);
        o.beta.assign(ob + 3 * nlam, ob + 3 * nlam + (size_t)ni * o.lmu);
        const double *sc = hscal.data() + (size_t)p * (3 * (size_t)ni + 4) + 3 * ni;
        o.nulldev = sc[1] * sc[1] * sc[2];  // sum (y - ybar)^2 = ys^2 * nt
        if (F > 0 && o.lmu > 0) {
            std::vector<double> cvk((size_t)F * o.lmu), wk(F);
            std::vector<int> nlk(F);
            for (int f = 0; f < F; ++f) {
                const int id = nlead + p * F + f;
                const int *fv = hint.data() + (size_t)id * ((size_t)ni + 4);
                nlk[f] = fv[ni] < o.lmu ? fv[ni] : o.lmu;
                if (fv[ni + 2] != 0 && o.jerr == 0) o.jerr = fv[ni + 2];
                o.nlp += fv[ni + 1];
                wk[f] = probs[p].seg_global.empty() ? (double)(probs[p].seg[f + 1] - probs[p].seg[f])
                                                     : (double)probs[p].seg_global[f];
                for (int l = 0; l < o.lmu; ++l) cvk[(size_t)f * o.lmu + l] = hcv[((size_t)p * F + f) * nlam + l];
            }
            ate_host_cv_summary(cvk, wk, nlk, F, o.lmu, o.lambda.data(), o.cvm, o.cvsd, &o.idx_min, &o.idx_1se);
        }
    }
    return 0;
}
