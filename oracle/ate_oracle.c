/*
 * ate_oracle.c -- CPU restatement of the reference's hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this file.  The product (libate_b200.so) never links, loads or calls it.
 *
 * PARITY UNPINNED.  The reference (/root/reference, pure R) keeps all of this arithmetic
 * in un-vendored, un-pinned CRAN packages (glmnet ~2.0-13..2.0-16, base R 3.4/3.5; see
 * ate_replication.Rmd:20-24 and :4 for the date) and holds no tests, fixtures or golden
 * vectors.  R is not installed here, so the restatement below is written from the
 * published algorithms of those packages and pinned only by (i) closed forms,
 * (ii) scikit-learn's independent coordinate descent, (iii) remembered known answers
 * of R's Mersenne-Twister (set.seed(1); runif(3), sample(10)).  See DESIGN.md.
 *
 * What follows which reference line:
 *   orc_scale            ate_replication.Rmd:72-74   scale(): centre, divide by sd (n-1)
 *   orc_expand_deg2      ate_functions.R:288-296     ordered pairwise products, c1-major
 *   orc_glmnet_gaussian  ate_functions.R:101-103,123-125,304-305   glmnet(family="gaussian")
 *   orc_glmnet_binomial  ate_functions.R:139-141                   glmnet(family="binomial")
 *   orc_cv_glmnet        same call sites (cv.glmnet wrapper: folds, cvm, cvsd, getmin)
 *   orc_predict          ate_functions.R:144 (response), :106,:128,:308-309 (coef at s)
 *   orc_aipw             ate_functions.R:241-243 (tau), :255-256 (sandwich se)
 *   orc_bootstrap_idx    ate_functions.R:245-252, :267-283 (nuisances held fixed)
 *   orc_r_*              base R set.seed / unif_rand / sample()  (ate_functions.R:269,
 *                        cv.glmnet's foldid = sample(rep(seq(nfolds), length = N)))
 *   orc_crossfit_aipw    the K-fold composite of SURVEY.md section 8 (a DECISION of this
 *                        build; the reference's only precedent is ate_functions.R:372-389)
 *
 * All matrices are column-major doubles (R's layout); indices crossing the API are
 * 1-based 32-bit ints as in R.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))

/* Thread control for the timed CPU-baseline legs of bench.py.  The composite has 2K independent
 * cv.glmnet jobs (outer level) of nfolds independent path fits each (inner level); R itself runs
 * all of them on one thread.  orc_set_threads(t) deals t threads over the two levels; t = 1 is the
 * faithful single-threaded timing.  Returns the (outer * inner) thread count in effect. */
static int g_outer_threads = 0, g_inner_threads = 1; /* 0: OpenMP's default at the outer level */
ORC_API int orc_set_threads(int t) {
#ifdef _OPENMP
    if (t < 1) t = 1;
    g_outer_threads = t < 10 ? t : 10;
    g_inner_threads = t / g_outer_threads;
    if (g_inner_threads < 1) g_inner_threads = 1;
    if (g_inner_threads > 10) g_inner_threads = 10;
    omp_set_max_active_levels(g_inner_threads > 1 ? 2 : 1);
    omp_set_dynamic(0);
    omp_set_num_threads(g_outer_threads);
    return g_outer_threads * g_inner_threads;
#else
    (void)t;
    return 1;
#endif
}
ORC_API int orc_get_threads(void) {
#ifdef _OPENMP
    return (g_outer_threads > 0 ? g_outer_threads : omp_get_max_threads()) * g_inner_threads;
#else
    return 1;
#endif
}

/* glmnet.control() internal constants */
#define G_FDEV 1.0e-5
#define G_DEVMAX 0.999
#define G_EPS 1.0e-6
#define G_BIG 9.9e35
#define G_MNLAM 5
#define G_PMIN 1.0e-9

/* ------------------------------------------------------------------------------------ */
/* base R numerics                                                                       */
/* ------------------------------------------------------------------------------------ */

/* R's mean(): long double accumulate, then one refinement pass (summary.c). */
ORC_API double orc_r_mean(const double *x, long n) {
    if (n <= 0) return NAN;
    long double s = 0.0L;
    for (long i = 0; i < n; ++i) s += x[i];
    s /= (long double)n;
    long double t = 0.0L;
    for (long i = 0; i < n; ++i) t += (x[i] - s);
    s += t / (long double)n;
    return (double)s;
}

/* mean(x, na.rm = TRUE): drops NA and NaN (not +-Inf) and shrinks the denominator. */
ORC_API double orc_r_mean_narm(const double *x, long n, long *n_kept) {
    double *tmp = (double *)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
    long m = 0;
    for (long i = 0; i < n; ++i)
        if (!isnan(x[i])) tmp[m++] = x[i];
    double r = orc_r_mean(tmp, m);
    free(tmp);
    if (n_kept) *n_kept = m;
    return r;
}

/* R's sd(): two-pass, n-1 denominator (cov.c). */
ORC_API double orc_r_sd(const double *x, long n) {
    if (n < 2) return NAN;
    double mu = orc_r_mean(x, n);
    long double ss = 0.0L;
    for (long i = 0; i < n; ++i) {
        long double d = (long double)x[i] - mu;
        ss += d * d;
    }
    return sqrt((double)(ss / (long double)(n - 1)));
}

/* scale(x): per column centre by mean and divide by sd (n-1).  In place.  `cols` selects
 * columns (0-based) or NULL for all.  center/scale out-vectors are optional. */
ORC_API void orc_scale(double *x, long n, int p, const int *cols, int ncols, double *center,
                       double *scl) {
    int m = cols ? ncols : p;
    for (int c = 0; c < m; ++c) {
        int j = cols ? cols[c] : c;
        double *xj = x + (size_t)j * n;
        double mu = orc_r_mean(xj, n);
        for (long i = 0; i < n; ++i) xj[i] -= mu;
        /* scale() uses sqrt(sum(x^2)/(n-1)) on the centred column */
        long double ss = 0.0L;
        for (long i = 0; i < n; ++i) ss += (long double)xj[i] * xj[i];
        double sd = sqrt((double)(ss / (long double)(n - 1)));
        for (long i = 0; i < n; ++i) xj[i] /= sd;
        if (center) center[c] = mu;
        if (scl) scl[c] = sd;
    }
}

/* belloni's expansion: for c1 in covs, for c2 in covs: col = c1*c2, appended after the
 * mains (include_main != 0) -> p0 + p0^2 columns, else just the p0^2 products. */
ORC_API void orc_expand_deg2(const double *x, long n, int p0, int include_main, double *out) {
    size_t o = 0;
    if (include_main) {
        memcpy(out, x, sizeof(double) * (size_t)n * p0);
        o = (size_t)n * p0;
    }
    for (int c1 = 0; c1 < p0; ++c1)
        for (int c2 = 0; c2 < p0; ++c2) {
            const double *a = x + (size_t)c1 * n, *b = x + (size_t)c2 * n;
            double *d = out + o;
            for (long i = 0; i < n; ++i) d[i] = a[i] * b[i];
            o += (size_t)n;
        }
}

/* ---- Mersenne-Twister exactly as R seeds and draws it ---- */
typedef struct {
    uint32_t mt[624];
    int mti;
} orc_rng;

static orc_rng g_rng;

ORC_API void orc_r_set_seed(uint32_t seed) {
    for (int j = 0; j < 50; ++j) seed = 69069u * seed + 1u;
    /* i_seed[0] is mti (overwritten with 624 by FixupSeeds), i_seed[1..624] the state */
    seed = 69069u * seed + 1u;
    for (int j = 0; j < 624; ++j) {
        seed = 69069u * seed + 1u;
        g_rng.mt[j] = seed;
    }
    g_rng.mti = 624;
}

static uint32_t mt_next(orc_rng *r) {
    static const uint32_t mag01[2] = {0x0u, 0x9908b0dfu};
    uint32_t y;
    if (r->mti >= 624) {
        int kk;
        for (kk = 0; kk < 624 - 397; kk++) {
            y = (r->mt[kk] & 0x80000000u) | (r->mt[kk + 1] & 0x7fffffffu);
            r->mt[kk] = r->mt[kk + 397] ^ (y >> 1) ^ mag01[y & 1u];
        }
        for (; kk < 623; kk++) {
            y = (r->mt[kk] & 0x80000000u) | (r->mt[kk + 1] & 0x7fffffffu);
            r->mt[kk] = r->mt[kk + (397 - 624)] ^ (y >> 1) ^ mag01[y & 1u];
        }
        y = (r->mt[623] & 0x80000000u) | (r->mt[0] & 0x7fffffffu);
        r->mt[623] = r->mt[396] ^ (y >> 1) ^ mag01[y & 1u];
        r->mti = 0;
    }
    y = r->mt[r->mti++];
    y ^= (y >> 11);
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= (y >> 18);
    return y;
}

ORC_API double orc_r_unif_rand(void) {
    const double i2_32m1 = 2.328306437080797e-10;
    double v = (double)mt_next(&g_rng) * 2.3283064365386963e-10;
    if (v <= 0.0) return 0.5 * i2_32m1;
    if ((1.0 - v) <= 0.0) return 1.0 - 0.5 * i2_32m1;
    return v;
}

/* R_unif_index: kind 0 = "Rounding" (R <= 3.5), kind 1 = "Rejection" (R >= 3.6). */
static double r_unif_index(double dn, int kind) {
    if (kind == 0) return floor(dn * orc_r_unif_rand());
    if (dn <= 0) return 0.0;
    int bits = (int)ceil(log2(dn));
    double dv;
    do {
        int64_t v = 0;
        for (int nb = 0; nb <= bits; nb += 16) {
            int v1 = (int)floor(orc_r_unif_rand() * 65536);
            v = 65536 * v + v1;
        }
        if (bits < 64) v &= (((int64_t)1 << bits) - 1);
        dv = (double)v;
    } while (dn <= dv);
    return dv;
}

/* sample(n, k, replace = TRUE): 1-based indices */
ORC_API void orc_r_sample_replace(int n, long k, int kind, int32_t *out) {
    for (long i = 0; i < k; ++i) out[i] = (int32_t)(r_unif_index((double)n, kind) + 1.0);
}

/* sample.int(n): a permutation of 1..n (do_sample without replacement, k = n) */
ORC_API void orc_r_sample_perm(int n, int kind, int32_t *out) {
    int32_t *x = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < n; ++i) x[i] = i;
    int m = n;
    for (int i = 0; i < n; ++i) {
        int j = (int)r_unif_index((double)m, kind);
        out[i] = x[j] + 1;
        x[j] = x[--m];
    }
    free(x);
}

/* cv.glmnet's default: foldid = sample(rep(seq(nfolds), length = N)) */
ORC_API void orc_r_foldid(int n, int nfolds, int kind, int32_t *foldid) {
    int32_t *perm = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    orc_r_sample_perm(n, kind, perm);
    for (int i = 0; i < n; ++i) foldid[i] = ((perm[i] - 1) % nfolds) + 1;
    free(perm);
}

/* ------------------------------------------------------------------------------------ */
/* glmnet: gaussian, covariance updating                                                 */
/* ------------------------------------------------------------------------------------ */

static double dotp(const double *a, const double *b, long n) {
    double s = 0.0;
    for (long i = 0; i < n; ++i) s += a[i] * b[i];
    return s;
}

static void chkvars(long no, int ni, const double *x, int *ju) {
    for (int j = 0; j < ni; ++j) {
        const double *xj = x + (size_t)j * no;
        ju[j] = 0;
        for (long i = 1; i < no; ++i)
            if (xj[i] != xj[0]) {
                ju[j] = 1;
                break;
            }
    }
}

static void norm_vp(int ni, const double *vp_in, double *vq) {
    double s = 0.0;
    for (int j = 0; j < ni; ++j) {
        vq[j] = vp_in ? (vp_in[j] > 0 ? vp_in[j] : 0.0) : 1.0;
        s += vq[j];
    }
    for (int j = 0; j < ni; ++j) vq[j] = vq[j] * ni / s;
}

/*
 * The Fortran entry `elnet` with ka = 1 (covariance updates), as glmnet() drives it.
 *   x  no*ni col-major (not modified), y no, w no or NULL (ones), vp ni or NULL (ones).
 *   flmin < 1: generate nlam lambdas down to flmin*lambda_max;  flmin >= 1: use ulam[].
 * Outputs: lmu, a0[nlam], beta[ni*nlam] (dense, original x scale), rsq[nlam], alm[nlam],
 * nlp (total passes).  Return value is glmnet's jerr (0 ok, <0 non-fatal, >0 fatal).
 */
ORC_API int orc_elnet(long no, int ni, const double *x_in, const double *y_in, const double *w_in,
                      const double *vp_in, double alpha, int ne, int nx, int nlam, double flmin,
                      const double *ulam, double thr, int isd, int intr, int maxit, int *lmu_out,
                      double *a0, double *beta, double *rsqo, double *almo, int *nlp_out) {
    int jerr = 0;
    size_t szx = (size_t)no * ni;
    double *x = (double *)malloc(sizeof(double) * szx);
    double *y = (double *)malloc(sizeof(double) * no);
    double *w = (double *)malloc(sizeof(double) * no);
    double *v = (double *)malloc(sizeof(double) * no);
    double *vq = (double *)malloc(sizeof(double) * ni);
    double *g = (double *)calloc(ni, sizeof(double));
    double *xm = (double *)calloc(ni, sizeof(double));
    double *xs = (double *)malloc(sizeof(double) * ni);
    double *xv = (double *)malloc(sizeof(double) * ni);
    double *a = (double *)calloc(ni, sizeof(double));
    double *da = (double *)calloc(ni, sizeof(double));
    double *c = (double *)malloc(sizeof(double) * (size_t)ni * (nx > 0 ? nx : 1));
    double *ao = (double *)calloc((size_t)(nx > 0 ? nx : 1) * nlam, sizeof(double));
    double *vlam = (double *)malloc(sizeof(double) * nlam);
    int *ju = (int *)malloc(sizeof(int) * ni);
    int *mm = (int *)calloc(ni, sizeof(int));
    int *ia = (int *)calloc(nx > 0 ? nx : 1, sizeof(int));
    int *ix = (int *)calloc(ni, sizeof(int));
    int *kin = (int *)calloc(nlam, sizeof(int));
    int lmu = 0, nlp = 0;
    double ym = 0.0, ys = 1.0;

    memcpy(x, x_in, sizeof(double) * szx);
    memcpy(y, y_in, sizeof(double) * no);
    for (long i = 0; i < no; ++i) w[i] = w_in ? w_in[i] : 1.0;
    norm_vp(ni, vp_in, vq);
    for (int j = 0; j < ni; ++j) xs[j] = 1.0;

    chkvars(no, ni, x, ju);
    {
        int any = 0;
        for (int j = 0; j < ni; ++j) any |= ju[j];
        if (!any) {
            jerr = 7777;
            goto done;
        }
    }
    /* ---- standard() ---- */
    {
        double sw = 0.0;
        for (long i = 0; i < no; ++i) sw += w[i];
        for (long i = 0; i < no; ++i) {
            w[i] /= sw;
            v[i] = sqrt(w[i]);
        }
        if (intr == 0) {
            ym = 0.0;
            for (long i = 0; i < no; ++i) y[i] = v[i] * y[i];
            double vy = dotp(v, y, no);
            ys = sqrt(dotp(y, y, no) - vy * vy);
            for (long i = 0; i < no; ++i) y[i] /= ys;
            for (int j = 0; j < ni; ++j) {
                if (!ju[j]) continue;
                double *xj = x + (size_t)j * no;
                xm[j] = 0.0;
                for (long i = 0; i < no; ++i) xj[i] = v[i] * xj[i];
                xv[j] = dotp(xj, xj, no);
                if (isd) {
                    double xb = dotp(v, xj, no);
                    double xbq = xb * xb, vc = xv[j] - xbq;
                    xs[j] = sqrt(vc);
                    for (long i = 0; i < no; ++i) xj[i] /= xs[j];
                    xv[j] = 1.0 + xbq / vc;
                }
            }
        } else {
            for (int j = 0; j < ni; ++j) {
                if (!ju[j]) continue;
                double *xj = x + (size_t)j * no;
                xm[j] = dotp(w, xj, no);
                for (long i = 0; i < no; ++i) xj[i] = v[i] * (xj[i] - xm[j]);
                xv[j] = dotp(xj, xj, no);
                if (isd) xs[j] = sqrt(xv[j]);
            }
            if (isd) {
                for (int j = 0; j < ni; ++j) {
                    if (!ju[j]) continue;
                    double *xj = x + (size_t)j * no;
                    for (long i = 0; i < no; ++i) xj[i] /= xs[j];
                    xv[j] = 1.0;
                }
            }
            ym = dotp(w, y, no);
            for (long i = 0; i < no; ++i) y[i] = v[i] * (y[i] - ym);
            ys = sqrt(dotp(y, y, no));
            for (long i = 0; i < no; ++i) y[i] /= ys;
        }
        for (int j = 0; j < ni; ++j)
            if (ju[j]) g[j] = dotp(y, x + (size_t)j * no, no);
    }
    if (flmin >= 1.0)
        for (int m = 0; m < nlam; ++m) vlam[m] = ulam[m] / ys;

    /* ---- elnet1 ---- */
    {
        const double bta = alpha, omb = 1.0 - alpha;
        double alf = 1.0, alm = 0.0, alm0, rsq = 0.0, rsq0;
        int nin = 0, iz = 0, jz;
        int mnl = G_MNLAM < nlam ? G_MNLAM : nlam;
        if (flmin < 1.0) {
            double eqs = G_EPS > flmin ? G_EPS : flmin;
            alf = pow(eqs, 1.0 / (nlam - 1));
        }
        for (int m = 0; m < nlam; ++m) {
            alm0 = alm;
            if (flmin >= 1.0)
                alm = vlam[m];
            else if (m > 1)
                alm *= alf;
            else if (m == 0)
                alm = G_BIG;
            else {
                alm0 = 0.0;
                for (int j = 0; j < ni; ++j)
                    if (ju[j] && vq[j] > 0.0) {
                        double t = fabs(g[j]) / vq[j];
                        if (t > alm0) alm0 = t;
                    }
                alm0 /= (bta > 1.0e-3 ? bta : 1.0e-3);
                alm = alf * alm0;
            }
            const double dem = alm * omb, ab = alm * bta;
            rsq0 = rsq;
            jz = 1;
            const double tlam = bta * (2.0 * alm - alm0);
            for (int k = 0; k < ni; ++k)
                if (!ix[k] && ju[k] && fabs(g[k]) > tlam * vq[k]) ix[k] = 1;
            int overflow = 0;
            for (;;) {
                if (!(iz && jz)) {
                    int converged_full = 0;
                    for (;;) { /* :again: */
                        ++nlp;
                        double dlx = 0.0;
                        for (int k = 0; k < ni; ++k) {
                            if (!ix[k]) continue;
                            double ak = a[k], u = g[k] + ak * xv[k], vv = fabs(u) - vq[k] * ab;
                            a[k] = 0.0;
                            if (vv > 0.0) a[k] = copysign(vv, u) / (xv[k] + vq[k] * dem);
                            if (a[k] == ak) continue;
                            if (mm[k] == 0) {
                                ++nin;
                                if (nin > nx) break;
                                double *cn = c + (size_t)(nin - 1) * ni;
                                for (int j = 0; j < ni; ++j) {
                                    if (!ju[j]) continue;
                                    if (mm[j] != 0)
                                        cn[j] = c[(size_t)(mm[j] - 1) * ni + k];
                                    else if (j == k)
                                        cn[j] = xv[j];
                                    else
                                        cn[j] = dotp(x + (size_t)j * no, x + (size_t)k * no, no);
                                }
                                mm[k] = nin;
                                ia[nin - 1] = k;
                            }
                            double del = a[k] - ak;
                            rsq += del * (2.0 * g[k] - del * xv[k]);
                            double t = xv[k] * del * del;
                            if (t > dlx) dlx = t;
                            const double *ck = c + (size_t)(mm[k] - 1) * ni;
                            for (int j = 0; j < ni; ++j)
                                if (ju[j]) g[j] -= ck[j] * del;
                        }
                        if (nin > nx) {
                            overflow = 1;
                            break;
                        }
                        if (dlx < thr) {
                            int ixx = 0;
                            for (int k = 0; k < ni; ++k)
                                if (!ix[k] && ju[k] && fabs(g[k]) > ab * vq[k]) {
                                    ix[k] = 1;
                                    ixx = 1;
                                }
                            if (ixx) continue;
                            converged_full = 1;
                        }
                        break;
                    }
                    if (overflow || converged_full) break;
                    if (nlp > maxit) {
                        jerr = -(m + 1);
                        goto unscale;
                    }
                }
                iz = 1;
                for (int l = 0; l < nin; ++l) da[l] = a[ia[l]];
                for (;;) {
                    ++nlp;
                    double dlx = 0.0;
                    for (int l = 0; l < nin; ++l) {
                        int k = ia[l];
                        double ak = a[k], u = g[k] + ak * xv[k], vv = fabs(u) - vq[k] * ab;
                        a[k] = 0.0;
                        if (vv > 0.0) a[k] = copysign(vv, u) / (xv[k] + vq[k] * dem);
                        if (a[k] == ak) continue;
                        double del = a[k] - ak;
                        rsq += del * (2.0 * g[k] - del * xv[k]);
                        double t = xv[k] * del * del;
                        if (t > dlx) dlx = t;
                        const double *ck = c + (size_t)(mm[k] - 1) * ni;
                        for (int j = 0; j < nin; ++j) g[ia[j]] -= ck[ia[j]] * del;
                    }
                    if (dlx < thr) break;
                    if (nlp > maxit) {
                        jerr = -(m + 1);
                        goto unscale;
                    }
                }
                for (int l = 0; l < nin; ++l) da[l] = a[ia[l]] - da[l];
                for (int j = 0; j < ni; ++j) {
                    if (mm[j] != 0 || !ju[j]) continue;
                    double s = 0.0;
                    for (int l = 0; l < nin; ++l) s += da[l] * c[(size_t)l * ni + j];
                    g[j] -= s;
                }
                jz = 0;
            }
            if (overflow) {
                jerr = -10000 - (m + 1);
                break;
            }
            for (int l = 0; l < nin; ++l) ao[(size_t)m * nx + l] = a[ia[l]];
            kin[m] = nin;
            rsqo[m] = rsq;
            almo[m] = alm;
            lmu = m + 1;
            if (m + 1 < mnl) continue;
            if (flmin >= 1.0) continue;
            int me = 0;
            for (int l = 0; l < nin; ++l)
                if (ao[(size_t)m * nx + l] != 0.0) ++me;
            if (me > ne) break;
            if (rsq - rsq0 < G_FDEV * rsq) break;
            if (rsq > G_DEVMAX) break;
        }
    }
unscale:
    /* elnetu tail: back to the caller's scale */
    for (int k = 0; k < lmu; ++k) {
        almo[k] *= ys;
        double *bk = beta + (size_t)k * ni;
        for (int j = 0; j < ni; ++j) bk[j] = 0.0;
        double s = 0.0;
        for (int l = 0; l < kin[k]; ++l) {
            int j = ia[l];
            double cv = ys * ao[(size_t)k * nx + l] / xs[j];
            bk[j] = cv;
            s += cv * xm[j];
        }
        a0[k] = intr ? ym - s : 0.0;
    }
done:
    *lmu_out = lmu;
    *nlp_out = nlp;
    free(x); free(y); free(w); free(v); free(vq); free(g); free(xm); free(xs); free(xv);
    free(a); free(da); free(c); free(ao); free(vlam); free(ju); free(mm); free(ia); free(ix);
    free(kin);
    return jerr;
}

/* ------------------------------------------------------------------------------------ */
/* glmnet: two-class binomial, Newton (lognet -> lstandard1 -> lognet2n)                 */
/* ------------------------------------------------------------------------------------ */

static double dev2(long n, const double *w, const double *y, const double *p, double pmin) {
    double pmax = 1.0 - pmin, s = 0.0;
    for (long i = 0; i < n; ++i) {
        double pi = p[i] < pmin ? pmin : (p[i] > pmax ? pmax : p[i]);
        s -= w[i] * (y[i] * log(pi) + (1.0 - y[i]) * log(1.0 - pi));
    }
    return s;
}

/*
 * y is 0/1 (the modelled class is 1), w_in observation weights or NULL.
 * Outputs as orc_elnet, with dev[] = fraction of null deviance explained and *dev0_out the
 * null deviance on R's scale (2 * sum(w) * dev0).
 */
ORC_API int orc_lognet(long no, int ni, const double *x_in, const double *y_in,
                       const double *w_in, const double *vp_in, double alpha, int ne, int nx,
                       int nlam, double flmin, const double *ulam, double shri, int isd, int intr,
                       int maxit, int *lmu_out, double *a0, double *beta, double *devo,
                       double *almo, int *nlp_out, double *dev0_out) {
    int jerr = 0;
    size_t szx = (size_t)no * ni;
    double *x = (double *)malloc(sizeof(double) * szx);
    double *w = (double *)malloc(sizeof(double) * no);
    double *q = (double *)malloc(sizeof(double) * no);
    double *r = (double *)malloc(sizeof(double) * no);
    double *v = (double *)malloc(sizeof(double) * no);
    double *vq = (double *)malloc(sizeof(double) * ni);
    double *xm = (double *)calloc(ni, sizeof(double));
    double *xs = (double *)malloc(sizeof(double) * ni);
    double *xv = (double *)calloc(ni, sizeof(double));
    double *ga = (double *)calloc(ni, sizeof(double));
    double *b = (double *)calloc(ni + 1, sizeof(double));  /* b[0] intercept, b[1+j] */
    double *bs = (double *)calloc(ni + 1, sizeof(double));
    double *ao = (double *)calloc((size_t)(nx > 0 ? nx : 1) * nlam, sizeof(double));
    int *ju = (int *)malloc(sizeof(int) * ni);
    int *mm = (int *)calloc(ni, sizeof(int));
    int *mlist = (int *)calloc(nx > 0 ? nx : 1, sizeof(int));
    int *ixx = (int *)calloc(ni, sizeof(int));
    int *kin = (int *)calloc(nlam, sizeof(int));
    const double *y = y_in;
    int lmu = 0, nlp = 0;
    double sw = 0.0, dev0 = 0.0;

    memcpy(x, x_in, sizeof(double) * szx);
    for (long i = 0; i < no; ++i) {
        w[i] = w_in ? w_in[i] : 1.0;
        sw += w[i];
    }
    for (long i = 0; i < no; ++i) w[i] /= sw;
    norm_vp(ni, vp_in, vq);
    for (int j = 0; j < ni; ++j) xs[j] = 1.0;
    chkvars(no, ni, x, ju);
    {
        int any = 0;
        for (int j = 0; j < ni; ++j) any |= ju[j];
        if (!any) {
            jerr = 7777;
            goto done;
        }
    }
    /* ---- lstandard1 ---- */
    for (int j = 0; j < ni; ++j) {
        if (!ju[j]) continue;
        double *xj = x + (size_t)j * no;
        if (intr == 0) {
            xm[j] = 0.0;
            if (isd) {
                double s2 = 0.0, s1 = 0.0;
                for (long i = 0; i < no; ++i) {
                    s2 += w[i] * xj[i] * xj[i];
                    s1 += w[i] * xj[i];
                }
                xs[j] = sqrt(s2 - s1 * s1);
                for (long i = 0; i < no; ++i) xj[i] /= xs[j];
            }
        } else {
            xm[j] = dotp(w, xj, no);
            for (long i = 0; i < no; ++i) xj[i] -= xm[j];
            if (isd) {
                double s2 = 0.0;
                for (long i = 0; i < no; ++i) s2 += w[i] * xj[i] * xj[i];
                xs[j] = sqrt(s2);
                for (long i = 0; i < no; ++i) xj[i] /= xs[j];
            }
        }
    }
    /* ---- lognet2n ---- */
    {
        const double pmin = G_PMIN;
        const double fmax = log(1.0 / pmin - 1.0), fmin = -fmax;
        const double vmin = (1.0 + pmin) * pmin * (1.0 - pmin);
        const double bta = alpha, omb = 1.0 - alpha;
        double q0 = dotp(w, y, no);
        if (q0 <= pmin) {
            jerr = 8001;
            goto done;
        }
        if (q0 >= 1.0 - pmin) {
            jerr = 9001;
            goto done;
        }
        if (intr == 0) q0 = 0.5;
        double bz = intr ? log(q0 / (1.0 - q0)) : 0.0;
        double vi = q0 * (1.0 - q0);
        b[0] = bz;
        for (long i = 0; i < no; ++i) {
            v[i] = vi * w[i];
            r[i] = w[i] * (y[i] - q0);
            q[i] = q0;
        }
        double xmz = vi;
        double dev1 = -(bz * q0 + log(1.0 - q0));
        dev0 = dev1;
        for (long i = 0; i < no; ++i) {
            if (y[i] > 0.0) dev0 += w[i] * y[i] * log(y[i]);
            if (y[i] < 1.0) dev0 += w[i] * (1.0 - y[i]) * log(1.0 - y[i]);
        }
        double alf = 1.0;
        if (flmin < 1.0) {
            double eqs = G_EPS > flmin ? G_EPS : flmin;
            alf = pow(eqs, 1.0 / (nlam - 1));
        }
        int nin = 0;
        int mnl = G_MNLAM < nlam ? G_MNLAM : nlam;
        const double shr = shri * dev0;
        for (int j = 0; j < ni; ++j)
            if (ju[j]) ga[j] = fabs(dotp(r, x + (size_t)j * no, no));
        double al = 0.0, al0;
        for (int ilm = 0; ilm < nlam; ++ilm) {
            al0 = al;
            if (flmin >= 1.0)
                al = ulam[ilm];
            else if (ilm > 1)
                al *= alf;
            else if (ilm == 0)
                al = G_BIG;
            else {
                al0 = 0.0;
                for (int j = 0; j < ni; ++j)
                    if (ju[j] && vq[j] > 0.0) {
                        double t = ga[j] / vq[j];
                        if (t > al0) al0 = t;
                    }
                al0 /= (bta > 1.0e-3 ? bta : 1.0e-3);
                al = alf * al0;
            }
            const double al2 = al * omb, al1 = al * bta, tlam = bta * (2.0 * al - al0);
            for (int k = 0; k < ni; ++k)
                if (!ixx[k] && ju[k] && ga[k] > tlam * vq[k]) ixx[k] = 1;
            int overflow = 0, stop_path = 0;
            for (;;) { /* IRLS outer loop */
                bs[0] = b[0];
                for (int l = 0; l < nin; ++l) bs[1 + mlist[l]] = b[1 + mlist[l]];
                for (int j = 0; j < ni; ++j) {
                    if (!ixx[j]) continue;
                    const double *xj = x + (size_t)j * no;
                    double s = 0.0;
                    for (long i = 0; i < no; ++i) s += v[i] * xj[i] * xj[i];
                    xv[j] = s;
                }
                for (;;) { /* middle loop: full sweep over the strong set */
                    ++nlp;
                    double dlx = 0.0;
                    for (int k = 0; k < ni; ++k) {
                        if (!ixx[k]) continue;
                        const double *xk = x + (size_t)k * no;
                        double bk = b[1 + k], gk = dotp(r, xk, no);
                        double u = gk + xv[k] * bk, au = fabs(u) - vq[k] * al1;
                        if (au <= 0.0)
                            b[1 + k] = 0.0;
                        else
                            b[1 + k] = copysign(au, u) / (xv[k] + vq[k] * al2);
                        double d = b[1 + k] - bk;
                        if (fabs(d) <= 0.0) continue;
                        double t = xv[k] * d * d;
                        if (t > dlx) dlx = t;
                        for (long i = 0; i < no; ++i) r[i] -= d * v[i] * xk[i];
                        if (mm[k] == 0) {
                            ++nin;
                            if (nin > nx) break;
                            mm[k] = nin;
                            mlist[nin - 1] = k;
                        }
                    }
                    if (nin > nx) {
                        overflow = 1;
                        break;
                    }
                    double d = 0.0;
                    if (intr) {
                        double sr = 0.0;
                        for (long i = 0; i < no; ++i) sr += r[i];
                        d = sr / xmz;
                    }
                    if (d != 0.0) {
                        b[0] += d;
                        double t = xmz * d * d;
                        if (t > dlx) dlx = t;
                        for (long i = 0; i < no; ++i) r[i] -= d * v[i];
                    }
                    if (dlx < shr) break;
                    if (nlp > maxit) {
                        jerr = -(ilm + 1);
                        goto unscale;
                    }
                    for (;;) { /* active-set loop */
                        ++nlp;
                        dlx = 0.0;
                        for (int l = 0; l < nin; ++l) {
                            int k = mlist[l];
                            const double *xk = x + (size_t)k * no;
                            double bk = b[1 + k], gk = dotp(r, xk, no);
                            double u = gk + xv[k] * bk, au = fabs(u) - vq[k] * al1;
                            if (au <= 0.0)
                                b[1 + k] = 0.0;
                            else
                                b[1 + k] = copysign(au, u) / (xv[k] + vq[k] * al2);
                            double dd = b[1 + k] - bk;
                            if (fabs(dd) <= 0.0) continue;
                            double t = xv[k] * dd * dd;
                            if (t > dlx) dlx = t;
                            for (long i = 0; i < no; ++i) r[i] -= dd * v[i] * xk[i];
                        }
                        d = 0.0;
                        if (intr) {
                            double sr = 0.0;
                            for (long i = 0; i < no; ++i) sr += r[i];
                            d = sr / xmz;
                        }
                        if (d != 0.0) {
                            b[0] += d;
                            double t = xmz * d * d;
                            if (t > dlx) dlx = t;
                            for (long i = 0; i < no; ++i) r[i] -= d * v[i];
                        }
                        if (dlx < shr) break;
                        if (nlp > maxit) {
                            jerr = -(ilm + 1);
                            goto unscale;
                        }
                    }
                }
                if (overflow) break;
                /* refresh the quadratic approximation */
                for (long i = 0; i < no; ++i) {
                    double fi = b[0];
                    for (int l = 0; l < nin; ++l) {
                        int k = mlist[l];
                        fi += b[1 + k] * x[(size_t)k * no + i];
                    }
                    if (fi < fmin)
                        q[i] = 0.0;
                    else if (fi > fmax)
                        q[i] = 1.0;
                    else
                        q[i] = 1.0 / (1.0 + exp(-fi));
                }
                xmz = 0.0;
                for (long i = 0; i < no; ++i) {
                    v[i] = w[i] * q[i] * (1.0 - q[i]);
                    xmz += v[i];
                }
                if (xmz <= vmin) {
                    stop_path = 1;
                    break;
                }
                for (long i = 0; i < no; ++i) r[i] = w[i] * (y[i] - q[i]);
                if (xmz * (b[0] - bs[0]) * (b[0] - bs[0]) < shr) {
                    int ix = 0;
                    for (int l = 0; l < nin; ++l) {
                        int k = mlist[l];
                        double dd = b[1 + k] - bs[1 + k];
                        if (xv[k] * dd * dd < shr) continue;
                        ix = 1;
                        break;
                    }
                    if (ix == 0) {
                        for (int k = 0; k < ni; ++k) {
                            if (ixx[k] || !ju[k]) continue;
                            ga[k] = fabs(dotp(r, x + (size_t)k * no, no));
                            if (ga[k] > al1 * vq[k]) {
                                ixx[k] = 1;
                                ix = 1;
                            }
                        }
                        if (ix == 1) continue;
                        break;
                    }
                }
            }
            if (overflow) {
                jerr = -10000 - (ilm + 1);
                break;
            }
            for (int l = 0; l < nin; ++l) ao[(size_t)ilm * nx + l] = b[1 + mlist[l]];
            kin[ilm] = nin;
            a0[ilm] = b[0];
            almo[ilm] = al;
            lmu = ilm + 1;
            devo[ilm] = (dev1 - dev2(no, w, y, q, pmin)) / dev0;
            if (stop_path || xmz <= vmin) break;
            if (ilm + 1 < mnl) continue;
            if (flmin >= 1.0) continue;
            int me = 0;
            for (int l = 0; l < nin; ++l)
                if (ao[(size_t)ilm * nx + l] != 0.0) ++me;
            if (me > ne) break;
            if (devo[ilm] > G_DEVMAX) break;
            if (devo[ilm] - devo[ilm - 1] < G_FDEV) break; /* [EXT] absolute, as in 2.0-x */
        }
    }
unscale:
    for (int k = 0; k < lmu; ++k) {
        double *bk = beta + (size_t)k * ni;
        for (int j = 0; j < ni; ++j) bk[j] = 0.0;
        double s = 0.0;
        for (int l = 0; l < kin[k]; ++l) {
            int j = mlist[l];
            double cv = ao[(size_t)k * nx + l] / xs[j];
            bk[j] = cv;
            s += cv * xm[j];
        }
        if (intr)
            a0[k] -= s;
        else
            a0[k] = 0.0;
    }
done:
    *lmu_out = lmu;
    *nlp_out = nlp;
    if (dev0_out) *dev0_out = 2.0 * sw * dev0;
    free(x); free(w); free(q); free(r); free(v); free(vq); free(xm); free(xs); free(xv);
    free(ga); free(b); free(bs); free(ao); free(ju); free(mm); free(mlist); free(ixx); free(kin);
    return jerr;
}

/* ------------------------------------------------------------------------------------ */
/* glmnet() R-level wrapper: defaults, lambda handling, fix.lam                          */
/* ------------------------------------------------------------------------------------ */

/* family: 0 gaussian, 1 binomial.  lambda_in == NULL -> automatic sequence. */
ORC_API int orc_glmnet(int family, long n, int p, const double *x, const double *y,
                       const double *w, const double *vp, double alpha, int nlambda,
                       double lambda_min_ratio, const double *lambda_in, int nlambda_in,
                       double thresh, int maxit, int standardize, int intercept, int *lmu,
                       double *lambda, double *a0, double *beta, double *dev_ratio, int *nlp,
                       double *nulldev) {
    int ne = p + 1;
    int nx = 2 * ne + 20 < p ? 2 * ne + 20 : p;
    double flmin;
    int nlam;
    const double *ulam;
    double dummy = 0.0;
    if (lambda_in) {
        flmin = 1.0;
        nlam = nlambda_in;
        ulam = lambda_in; /* caller passes a decreasing sequence, as rev(sort(lambda)) would */
    } else {
        flmin = lambda_min_ratio > 0 ? lambda_min_ratio : (n < p ? 0.01 : 1.0e-4);
        nlam = nlambda;
        ulam = &dummy;
    }
    int jerr;
    if (family == 0) {
        jerr = orc_elnet(n, p, x, y, w, vp, alpha, ne, nx, nlam, flmin, ulam, thresh, standardize,
                         intercept, maxit, lmu, a0, beta, dev_ratio, lambda, nlp);
        if (nulldev) {
            double sw = 0, sy = 0, ss = 0;
            for (long i = 0; i < n; ++i) {
                double wi = w ? w[i] : 1.0;
                sw += wi;
                sy += wi * y[i];
            }
            double yb = sy / sw;
            for (long i = 0; i < n; ++i) ss += (w ? w[i] : 1.0) * (y[i] - yb) * (y[i] - yb);
            *nulldev = ss;
        }
    } else {
        jerr = orc_lognet(n, p, x, y, w, vp, alpha, ne, nx, nlam, flmin, ulam, thresh,
                          standardize, intercept, maxit, lmu, a0, beta, dev_ratio, lambda, nlp,
                          nulldev);
    }
    /* fix.lam: lambda[1] was `big`; R rewrites it on the log-linear extrapolation */
    if (!lambda_in && *lmu > 2) lambda[0] = exp(2.0 * log(lambda[1]) - log(lambda[2]));
    return jerr;
}

/* ------------------------------------------------------------------------------------ */
/* cv.glmnet                                                                             */
/* ------------------------------------------------------------------------------------ */

static void take_rows(const double *x, long n, int p, const int32_t *foldid, int fold, int keep_eq,
                      double *out, long m) {
    for (int j = 0; j < p; ++j) {
        const double *xj = x + (size_t)j * n;
        double *oj = out + (size_t)j * m;
        long t = 0;
        for (long i = 0; i < n; ++i)
            if ((foldid[i] == fold) == (keep_eq != 0)) oj[t++] = xj[i];
    }
}

/*
 * Outputs (length *lmu each unless noted): lambda, a0, beta[p * lmu] of the full-data fit,
 * cvm, cvsd; idx_min / idx_1se are 0-based positions of lambda.min / lambda.1se.
 * type.measure = "deviance" (the default of both families).  weights = 1, grouped = TRUE.
 */
ORC_API int orc_cv_glmnet(int family, long n, int p, const double *x, const double *y,
                          const double *vp, double alpha, int nfolds, const int32_t *foldid,
                          int nlambda, double lambda_min_ratio, double thresh, int maxit,
                          int standardize, int intercept, int *lmu_out, double *lambda, double *a0,
                          double *beta, double *cvm, double *cvsd, int *idx_min, int *idx_1se,
                          double *dev_ratio) {
    int lmu = 0, nlp = 0;
    double nulldev;
    int jerr = orc_glmnet(family, n, p, x, y, NULL, vp, alpha, nlambda, lambda_min_ratio, NULL, 0,
                          thresh, maxit, standardize, intercept, &lmu, lambda, a0, beta, dev_ratio,
                          &nlp, &nulldev);
    *lmu_out = lmu;
    if (jerr > 0 || lmu == 0) return jerr ? jerr : 1;
    double *cvk = (double *)calloc((size_t)nfolds * lmu, sizeof(double)); /* per-fold mean */
    double *wk = (double *)calloc(nfolds, sizeof(double));
    int *nlk = (int *)calloc(nfolds, sizeof(int));
    int worst = 0;
#ifdef _OPENMP
    const int cv_threads = omp_in_parallel() ? g_inner_threads : (g_outer_threads > 0 ? g_outer_threads * g_inner_threads : omp_get_max_threads());
#else
    const int cv_threads = 1;
#endif
#pragma omp parallel for schedule(dynamic, 1) num_threads(cv_threads)
    for (int f = 1; f <= nfolds; ++f) {
        long nin_ = 0;
        for (long i = 0; i < n; ++i) nin_ += (foldid[i] == f);
        long ntr = n - nin_;
        double *xt = (double *)malloc(sizeof(double) * (size_t)ntr * p);
        double *yt = (double *)malloc(sizeof(double) * ntr);
        double *xh = (double *)malloc(sizeof(double) * (size_t)(nin_ > 0 ? nin_ : 1) * p);
        double *yh = (double *)malloc(sizeof(double) * (nin_ > 0 ? nin_ : 1));
        take_rows(x, n, p, foldid, f, 0, xt, ntr);
        take_rows(y, n, 1, foldid, f, 0, yt, ntr);
        take_rows(x, n, p, foldid, f, 1, xh, nin_);
        take_rows(y, n, 1, foldid, f, 1, yh, nin_);
        double *fa0 = (double *)calloc(lmu, sizeof(double));
        double *fb = (double *)calloc((size_t)p * lmu, sizeof(double));
        double *fl = (double *)calloc(lmu, sizeof(double));
        double *fd = (double *)calloc(lmu, sizeof(double));
        int flmu = 0, fnlp = 0;
        double fnd;
        int je = orc_glmnet(family, ntr, p, xt, yt, NULL, vp, alpha, lmu, 0.0, lambda, lmu, thresh,
                            maxit, standardize, intercept, &flmu, fl, fa0, fb, fd, &fnlp, &fnd);
        if (je > 0) {
#pragma omp critical
            worst = je;
        }
        nlk[f - 1] = flmu;
        wk[f - 1] = (double)nin_;
        for (int l = 0; l < flmu; ++l) {
            const double *bl = fb + (size_t)l * p;
            double s = 0.0;
            for (long i = 0; i < nin_; ++i) {
                double eta = fa0[l];
                for (int j = 0; j < p; ++j)
                    if (bl[j] != 0.0) eta += bl[j] * xh[(size_t)j * nin_ + i];
                if (family == 0) {
                    double e = yh[i] - eta;
                    s += e * e;
                } else {
                    double pr = 1.0 / (1.0 + exp(-eta));
                    pr = pr < 1.0e-5 ? 1.0e-5 : (pr > 1.0 - 1.0e-5 ? 1.0 - 1.0e-5 : pr);
                    s += -2.0 * (yh[i] * log(pr) + (1.0 - yh[i]) * log(1.0 - pr));
                }
            }
            cvk[(size_t)(f - 1) * lmu + l] = s / (double)nin_;
        }
        free(xt); free(yt); free(xh); free(yh); free(fa0); free(fb); free(fl); free(fd);
    }
    if (worst > 0) {
        free(cvk); free(wk); free(nlk);
        return worst;
    }
    for (int l = 0; l < lmu; ++l) {
        double sw = 0.0, sm = 0.0;
        int nn = 0;
        for (int f = 0; f < nfolds; ++f)
            if (l < nlk[f]) {
                sw += wk[f];
                sm += wk[f] * cvk[(size_t)f * lmu + l];
                ++nn;
            }
        double m = sm / sw, sv = 0.0;
        for (int f = 0; f < nfolds; ++f)
            if (l < nlk[f]) {
                double d = cvk[(size_t)f * lmu + l] - m;
                sv += wk[f] * d * d;
            }
        cvm[l] = m;
        cvsd[l] = sqrt((sv / sw) / (double)(nn - 1));
    }
    /* getmin */
    {
        double cvmin = cvm[0];
        for (int l = 1; l < lmu; ++l)
            if (cvm[l] < cvmin) cvmin = cvm[l];
        int imin = -1;
        for (int l = 0; l < lmu; ++l)
            if (cvm[l] <= cvmin && (imin < 0 || lambda[l] > lambda[imin])) imin = l;
        double semin = cvm[imin] + cvsd[imin];
        int i1 = -1;
        for (int l = 0; l < lmu; ++l)
            if (cvm[l] <= semin && (i1 < 0 || lambda[l] > lambda[i1])) i1 = l;
        *idx_min = imin;
        *idx_1se = i1;
    }
    free(cvk); free(wk); free(nlk);
    return jerr;
}

/* predict(): link (type 0) or response (type 1; plogis for binomial) at one coefficient set */
ORC_API void orc_predict(long n, int p, const double *x, double a0, const double *beta,
                         int response_binomial, double *out) {
    for (long i = 0; i < n; ++i) out[i] = a0;
    for (int j = 0; j < p; ++j) {
        if (beta[j] == 0.0) continue;
        const double *xj = x + (size_t)j * n;
        for (long i = 0; i < n; ++i) out[i] += beta[j] * xj[i];
    }
    if (response_binomial)
        for (long i = 0; i < n; ++i) out[i] = 1.0 / (1.0 + exp(-out[i]));
}

/* ------------------------------------------------------------------------------------ */
/* AIPW score, sandwich SE, bootstrap                                                    */
/* ------------------------------------------------------------------------------------ */

/* ate_functions.R:241-243 and :255-256.  est1/est2 are optional outputs (length n). */
ORC_API void orc_aipw(long n, const double *w, const double *y, const double *p,
                      const double *m0, const double *m1, double *tau_out, double *se_out,
                      double *est1_out, double *est2_out) {
    double *e1 = (double *)malloc(sizeof(double) * n);
    double *e2 = (double *)malloc(sizeof(double) * n);
    for (long i = 0; i < n; ++i) {
        e1[i] = w[i] * (y[i] - m1[i]) / p[i] + (1.0 - w[i]) * (y[i] - m0[i]) / (1.0 - p[i]);
        e2[i] = m1[i] - m0[i];
    }
    double tau = orc_r_mean_narm(e1, n, NULL) + orc_r_mean(e2, n);
    long double ss = 0.0L;
    for (long i = 0; i < n; ++i) {
        double Ii = (w[i] * y[i]) / p[i] - m1[i] * (w[i] - p[i]) / p[i] -
                    (((1.0 - w[i]) * y[i] / (1.0 - p[i])) + (m0[i] * (w[i] - p[i]) / (1.0 - p[i]))) -
                    tau;
        ss += (long double)Ii * Ii;
    }
    *tau_out = tau;
    *se_out = sqrt((double)ss * pow((double)n, -2.0));
    if (est1_out) memcpy(est1_out, e1, sizeof(double) * n);
    if (est2_out) memcpy(est2_out, e2, sizeof(double) * n);
    free(e1);
    free(e2);
}

/* One replicate of tau_hat_dr_est (ate_functions.R:267-283) for a given 1-based index
 * vector; the five vectors are gathered and the score recomputed exactly as written. */
ORC_API double orc_tau_hat_dr_est(long n, const double *w, const double *y, const double *p,
                                  const double *m0, const double *m1, const int32_t *idx) {
    double *e1 = (double *)malloc(sizeof(double) * n);
    double *e2 = (double *)malloc(sizeof(double) * n);
    for (long t = 0; t < n; ++t) {
        long i = idx[t] - 1;
        e1[t] = w[i] * (y[i] - m1[i]) / p[i] + (1.0 - w[i]) * (y[i] - m0[i]) / (1.0 - p[i]);
        e2[t] = m1[i] - m0[i];
    }
    double r = orc_r_mean_narm(e1, n, NULL) + orc_r_mean(e2, n);
    free(e1);
    free(e2);
    return r;
}

/* The bootstrap driver (:245-252) with a supplied B x n index matrix (replicate-major:
 * idx[b*n + t]); boot[B] receives the replicate estimates, return value is sd(boot). */
ORC_API double orc_bootstrap_idx(long n, const double *w, const double *y, const double *p,
                                 const double *m0, const double *m1, const int32_t *idx, int B,
                                 double *boot) {
#ifdef _OPENMP
    const int bt = g_outer_threads > 0 ? g_outer_threads * g_inner_threads : omp_get_max_threads();
#else
    const int bt = 1;
#endif
#pragma omp parallel for schedule(static) num_threads(bt)
    for (int b = 0; b < B; ++b)
        boot[b] = orc_tau_hat_dr_est(n, w, y, p, m0, m1, idx + (size_t)b * n);
    return orc_r_sd(boot, B);
}

/* Same, drawing the indices with R's sample() from the current RNG state. */
ORC_API double orc_bootstrap_r(long n, const double *w, const double *y, const double *p,
                               const double *m0, const double *m1, int B, int kind, double *boot) {
    int32_t *idx = (int32_t *)malloc(sizeof(int32_t) * n);
    for (int b = 0; b < B; ++b) {
        orc_r_sample_replace((int)n, n, kind, idx);
        boot[b] = orc_tau_hat_dr_est(n, w, y, p, m0, m1, idx);
    }
    free(idx);
    return orc_r_sd(boot, B);
}

/* ------------------------------------------------------------------------------------ */
/* The cross-fitted composite                           */
/* ------------------------------------------------------------------------------------ */

/*
 * x: n*p design (col-major), w, y: n.  foldid: n outer folds (1..K).
 * inner_ps / inner_out: K*n inner fold ids (row k holds the ids cv.glmnet would draw for the
 * training rows of outer fold k, in training-row order, the tail unused).
 * use_1se: 1 -> lambda.1se (the reference's default s), 0 -> lambda.min.
 * Outputs: phat, m0, m1 (n, stitched out-of-fold), tau, se (sandwich).
 */
ORC_API int orc_crossfit_aipw(long n, int p, const double *x, const double *w, const double *y,
                              int K, const int32_t *foldid, int nfolds_inner,
                              const int32_t *inner_ps, const int32_t *inner_out, int nlambda,
                              double thresh, int maxit, int use_1se, double *phat, double *m0,
                              double *m1, double *tau, double *se, double *lam_ps,
                              double *lam_out) {
    int rc = 0;
#pragma omp parallel for schedule(dynamic, 1)
    for (int job = 0; job < 2 * K; ++job) {
        int k = job / 2 + 1, which = job % 2; /* 0 propensity, 1 outcome */
        long nh = 0;
        for (long i = 0; i < n; ++i) nh += (foldid[i] == k);
        long nt = n - nh;
        int pp = which ? p + 1 : p;
        double *xt = (double *)malloc(sizeof(double) * (size_t)nt * pp);
        double *xh = (double *)malloc(sizeof(double) * (size_t)(nh > 0 ? nh : 1) * pp);
        double *yt = (double *)malloc(sizeof(double) * nt);
        take_rows(x, n, p, foldid, k, 0, xt, nt);
        take_rows(x, n, p, foldid, k, 1, xh, nh);
        double *vp = NULL;
        if (which) { /* [covariates, W], penalty.factor = c(1..1, 0): ate_functions.R:91-103 */
            take_rows(w, n, 1, foldid, k, 0, xt + (size_t)p * nt, nt);
            take_rows(y, n, 1, foldid, k, 0, yt, nt);
            vp = (double *)malloc(sizeof(double) * pp);
            for (int j = 0; j < p; ++j) vp[j] = 1.0;
            vp[p] = 0.0;
        } else {
            take_rows(w, n, 1, foldid, k, 0, yt, nt);
        }
        double *lambda = (double *)calloc(nlambda, sizeof(double));
        double *a0 = (double *)calloc(nlambda, sizeof(double));
        double *beta = (double *)calloc((size_t)pp * nlambda, sizeof(double));
        double *cvm = (double *)calloc(nlambda, sizeof(double));
        double *cvsd = (double *)calloc(nlambda, sizeof(double));
        double *dr = (double *)calloc(nlambda, sizeof(double));
        int lmu, imin, i1se;
        const int32_t *inner = (which ? inner_out : inner_ps) + (size_t)(k - 1) * n;
        int je = orc_cv_glmnet(which ? 0 : 1, nt, pp, xt, yt, vp, 1.0, nfolds_inner, inner, nlambda,
                               0.0, thresh, maxit, 1, 1, &lmu, lambda, a0, beta, cvm, cvsd, &imin,
                               &i1se, dr);
        if (je > 0) {
#pragma omp critical
            rc = je;
        } else {
            int s = use_1se ? i1se : imin;
            const double *bs = beta + (size_t)s * pp;
            double *tmp = (double *)malloc(sizeof(double) * (nh > 0 ? nh : 1));
            if (!which) {
                orc_predict(nh, p, xh, a0[s], bs, 1, tmp);
                long t = 0;
                for (long i = 0; i < n; ++i)
                    if (foldid[i] == k) phat[i] = tmp[t++];
                if (lam_ps) lam_ps[k - 1] = lambda[s];
            } else {
                /* predict with W := 1 and W := 0 (ate_functions.R:221-228) */
                orc_predict(nh, p, xh, a0[s], bs, 0, tmp);
                long t = 0;
                for (long i = 0; i < n; ++i)
                    if (foldid[i] == k) {
                        m0[i] = tmp[t];
                        m1[i] = tmp[t] + bs[p];
                        ++t;
                    }
                if (lam_out) lam_out[k - 1] = lambda[s];
            }
            free(tmp);
        }
        free(xt); free(xh); free(yt); free(vp); free(lambda); free(a0); free(beta); free(cvm);
        free(cvsd); free(dr);
    }
    if (rc > 0) return rc;
    orc_aipw(n, w, y, phat, m0, m1, tau, se, NULL, NULL);
    return 0;
}

/* ------------------------------------------------------------------------------------ */
/* stats::glm(family = binomial("logit")) -- the nuisance fits of doubly_robust_glm      */
/* (ate_functions.R:218-220, :231-234; also :156-158 and ate_replication.Rmd:165-168)     */
/* ------------------------------------------------------------------------------------ */

/* binomial()$linkinv as R's C code computes it (family.c: logit_linkinv) */
static double r_logit_linkinv(double eta) {
    const double thresh = 30.0, mthresh = -30.0, inveps = 1.0 / 2.220446049250313e-16;
    double tmp = (eta < mthresh) ? 2.220446049250313e-16 : ((eta > thresh) ? inveps : exp(eta));
    return tmp / (1.0 + tmp);
}

static double binom_dev(long n, const double *y, const double *mu) {
    /* sum(binomial()$dev.resids(y, mu, 1)) with y in {0,1}: y log(y/mu) terms vanish at y = 0 */
    double s = 0.0;
    for (long i = 0; i < n; ++i) {
        double a = y[i] > 0.0 ? y[i] * log(y[i] / mu[i]) : 0.0;
        double b = y[i] < 1.0 ? (1.0 - y[i]) * log((1.0 - y[i]) / (1.0 - mu[i])) : 0.0;
        s += 2.0 * (a + b);
    }
    return s;
}

/* ------------------------------------------------------------------------------------------
 * Householder QR least squares as R's lm.fit / lm.wfit / glm.fit do it: Cdqrls -> LINPACK dqrdc2
 * (Householder with LIMITED column pivoting: a column whose remaining norm has fallen below
 * tol * its original norm is moved to the end and the rank reduced; tol = 1e-7) followed by dqrsl
 * (Q'y, back-substitution).  [EXT] restated from the published LINPACK/R algorithm; lm call sites:
 * ate_functions.R:28 (ate_condmean_ols), :53-54 (prop_score_weight), :74-75 (prop_score_ols),
 * :320 (belloni); glm call sites :218-220, :231-234.
 * X: n x q col-major, overwritten; y: n, overwritten with Q'y.  pivot[q]: original column of each
 * position.  coef[q]: by ORIGINAL column, NaN for the aliased ones.  rinv_diag[q] (optional):
 * diag((X'X)^-1) by original column over the estimable ones (summary.lm's chol2inv(R)).
 * Returns the rank.
 */
static int qr_ls(long n, int q, double *X, double *y, double tol, double *coef, double *xtx_inv_diag) {
    int *pivot = (int *)malloc(sizeof(int) * q);
    double *qraux = (double *)malloc(sizeof(double) * q), *w0 = (double *)malloc(sizeof(double) * q),
           *w1 = (double *)malloc(sizeof(double) * q), *tmpc = (double *)malloc(sizeof(double) * n);
    for (int j = 0; j < q; ++j) {
        pivot[j] = j;
        double s = 0.0;
        for (long i = 0; i < n; ++i) s += X[(size_t)j * n + i] * X[(size_t)j * n + i];
        qraux[j] = sqrt(s);
        w0[j] = qraux[j];
        w1[j] = qraux[j];
        if (w1[j] == 0.0) w1[j] = 1.0;
    }
    const int lup = q < n ? q : (int)n;
    int k = q + 1; /* 1-based: rank + 1 */
    for (int l = 1; l <= lup; ++l) {
        /* cycle negligible columns to the end */
        while (l < k && qraux[l - 1] < w1[l - 1] * tol) {
            const int lp1 = l + 1;
            memcpy(tmpc, X + (size_t)(l - 1) * n, sizeof(double) * n);
            const int pi = pivot[l - 1];
            const double qa = qraux[l - 1], a0 = w0[l - 1], a1 = w1[l - 1];
            for (int j = lp1; j <= q; ++j) {
                memcpy(X + (size_t)(j - 2) * n, X + (size_t)(j - 1) * n, sizeof(double) * n);
                pivot[j - 2] = pivot[j - 1];
                qraux[j - 2] = qraux[j - 1];
                w0[j - 2] = w0[j - 1];
                w1[j - 2] = w1[j - 1];
            }
            memcpy(X + (size_t)(q - 1) * n, tmpc, sizeof(double) * n);
            pivot[q - 1] = pi;
            qraux[q - 1] = qa;
            w0[q - 1] = a0;
            w1[q - 1] = a1;
            --k;
        }
        if (l == n) break;
        double *xl = X + (size_t)(l - 1) * n;
        double nrm = 0.0;
        for (long i = l - 1; i < n; ++i) nrm += xl[i] * xl[i];
        nrm = sqrt(nrm);
        if (nrm != 0.0) {
            if (xl[l - 1] != 0.0) nrm = xl[l - 1] < 0.0 ? -fabs(nrm) : fabs(nrm);
            for (long i = l - 1; i < n; ++i) xl[i] /= nrm;
            xl[l - 1] += 1.0;
            for (int j = l + 1; j <= q; ++j) {
                double *xj = X + (size_t)(j - 1) * n;
                double t = 0.0;
                for (long i = l - 1; i < n; ++i) t -= xl[i] * xj[i];
                t /= xl[l - 1];
                for (long i = l - 1; i < n; ++i) xj[i] += t * xl[i];
                if (qraux[j - 1] != 0.0) {
                    double tt = 1.0 - (fabs(xj[l - 1]) / qraux[j - 1]) * (fabs(xj[l - 1]) / qraux[j - 1]);
                    if (tt < 0.0) tt = 0.0;
                    if (fabs(tt) < 1e-6) {
                        double s = 0.0;
                        for (long i = l; i < n; ++i) s += xj[i] * xj[i];
                        qraux[j - 1] = sqrt(s);
                        w0[j - 1] = qraux[j - 1];
                    } else
                        qraux[j - 1] *= sqrt(tt);
                }
            }
            qraux[l - 1] = xl[l - 1];
            xl[l - 1] = -nrm;
        }
    }
    int rank = k - 1;
    if (rank > n) rank = (int)n;
    /* dqrsl: y <- Q'y over the first `rank` reflectors */
    for (int l = 1; l <= rank && l < n + 0; ++l) {
        if (qraux[l - 1] == 0.0) continue;
        double *xl = X + (size_t)(l - 1) * n;
        const double save = xl[l - 1];
        xl[l - 1] = qraux[l - 1];
        double t = 0.0;
        for (long i = l - 1; i < n; ++i) t -= xl[i] * y[i];
        t /= xl[l - 1];
        for (long i = l - 1; i < n; ++i) y[i] += t * xl[i];
        xl[l - 1] = save;
    }
    for (int j = 0; j < q; ++j) coef[j] = NAN;
    double *b = (double *)malloc(sizeof(double) * (rank > 0 ? rank : 1));
    for (int a = rank - 1; a >= 0; --a) {
        double s = y[a];
        for (int c = a + 1; c < rank; ++c) s -= X[(size_t)c * n + a] * b[c];
        b[a] = s / X[(size_t)a * n + a];
    }
    for (int a = 0; a < rank; ++a) coef[pivot[a]] = b[a];
    if (xtx_inv_diag) {
        /* (R'R)^-1 = R^-1 R^-T; diagonal entry a = sum_c (R^-1)[a,c]^2 over c >= a */
        double *Ri = (double *)calloc((size_t)(rank > 0 ? rank : 1) * (rank > 0 ? rank : 1), sizeof(double));
        for (int c = 0; c < rank; ++c) {
            Ri[(size_t)c * rank + c] = 1.0 / X[(size_t)c * n + c];
            for (int a = c - 1; a >= 0; --a) {
                double s = 0.0;
                for (int m = a + 1; m <= c; ++m) s -= X[(size_t)m * n + a] * Ri[(size_t)c * rank + m];
                Ri[(size_t)c * rank + a] = s / X[(size_t)a * n + a];
            }
        }
        for (int j = 0; j < q; ++j) xtx_inv_diag[j] = NAN;
        for (int a = 0; a < rank; ++a) {
            double s = 0.0;
            for (int c = a; c < rank; ++c) s += Ri[(size_t)c * rank + a] * Ri[(size_t)c * rank + a];
            xtx_inv_diag[pivot[a]] = s;
        }
        free(Ri);
    }
    free(b); free(pivot); free(qraux); free(w0); free(w1); free(tmpc);
    return rank;
}

/*
 * lm(y ~ x [, weights = w]) followed by summary(): lm.fit / lm.wfit (rows scaled by sqrt(w), zero
 * weights dropped from the degrees of freedom), coefficient table `Estimate`, `Std. Error` =
 * sqrt(resvar * diag((X'X)^-1)), resvar = rss / rdf, rss = sum w e^2 (summary.lm).
 * x: n*p col-major WITHOUT intercept; coef/se [p+1], (Intercept) first, NaN = NA (aliased).
 * resid (optional): y - fitted.  Returns the rank.
 */
ORC_API int orc_lm(long n, int p, const double *x, const double *y, const double *w, double *coef,
                   double *se, double *sigma, double *rss_out, long *rdf_out, double *resid) {
    const int q = p + 1;
    double *X = (double *)malloc(sizeof(double) * (size_t)n * q), *yy = (double *)malloc(sizeof(double) * n);
    double *dinv = (double *)malloc(sizeof(double) * q);
    long nok = 0;
    for (long i = 0; i < n; ++i) {
        const double sw = w ? sqrt(w[i]) : 1.0;
        if (!w || w[i] != 0.0) ++nok;
        X[i] = sw;
        for (int j = 0; j < p; ++j) X[(size_t)(j + 1) * n + i] = sw * x[(size_t)j * n + i];
        yy[i] = sw * y[i];
    }
    const int rank = qr_ls(n, q, X, yy, 1e-7, coef, dinv);
    double rss = 0.0;
    for (long i = 0; i < n; ++i) {
        double f = 0.0;
        for (int j = 0; j < q; ++j)
            if (coef[j] == coef[j]) f += coef[j] * (j == 0 ? 1.0 : x[(size_t)(j - 1) * n + i]);
        const double e = y[i] - f;
        if (resid) resid[i] = e;
        rss += (w ? w[i] : 1.0) * e * e;
    }
    const long rdf = nok - rank;
    const double resvar = rdf > 0 ? rss / (double)rdf : NAN;
    for (int j = 0; j < q; ++j) se[j] = sqrt(resvar * dinv[j]);
    if (sigma) *sigma = sqrt(resvar);
    if (rss_out) *rss_out = rss;
    if (rdf_out) *rdf_out = rdf;
    free(X); free(yy); free(dinv);
    return rank;
}

/*
 * glm.fit's IRLS: mustart = (y + 0.5)/2, weighted least squares each step, stop when
 * |dev - devold| / (|dev| + 0.1) < epsilon (1e-8), at most maxit (25) steps.
 * Each step is solved as R does (Cdqrls: rows scaled by sqrt(w), Householder QR, qr_ls above).
 * x: n*p col-major WITHOUT intercept; coef[p+1]: (Intercept) first, then the columns of x.
 * Returns 0 if converged, 1 otherwise.
 */
ORC_API int orc_glm_binomial(long n, int p, const double *x, const double *y, int maxit,
                             double epsilon, double *coef, double *deviance, int *iters) {
    const int q = p + 1;
    double *eta = (double *)malloc(sizeof(double) * n), *mu = (double *)malloc(sizeof(double) * n);
    double *Xw = (double *)malloc(sizeof(double) * (size_t)n * q), *zw = (double *)malloc(sizeof(double) * n);
    for (long i = 0; i < n; ++i) {
        mu[i] = (y[i] + 0.5) / 2.0;
        eta[i] = log(mu[i] / (1.0 - mu[i]));
    }
    double devold = binom_dev(n, y, mu), dev = devold;
    int conv = 0, it = 0;
    for (it = 1; it <= maxit; ++it) {
        for (long i = 0; i < n; ++i) {
            const double w = mu[i] * (1.0 - mu[i]);
            const double z = eta[i] + (y[i] - mu[i]) / w;
            const double sw = sqrt(w);
            Xw[i] = sw;
            for (int a = 1; a < q; ++a) Xw[(size_t)a * n + i] = sw * x[(size_t)(a - 1) * n + i];
            zw[i] = sw * z;
        }
        const int rank = qr_ls(n, q, Xw, zw, fmin(1e-7, epsilon / 1000.0) /* glm.fit: min(1e-07, control$epsilon/1000) */, coef, NULL);
        if (rank < q) { /* glm would carry NA coefficients on; this path treats it as a failure */
            free(eta); free(mu); free(Xw); free(zw);
            return 2;
        }
        for (long i = 0; i < n; ++i) {
            double e = coef[0];
            for (int a = 1; a < q; ++a) e += coef[a] * x[(size_t)(a - 1) * n + i];
            eta[i] = e;
            mu[i] = r_logit_linkinv(e);
        }
        dev = binom_dev(n, y, mu);
        if (fabs(dev - devold) / (fabs(dev) + 0.1) < epsilon) {
            conv = 1;
            break;
        }
        devold = dev;
    }
    *deviance = dev;
    *iters = it > maxit ? maxit : it;
    free(eta); free(mu); free(Xw); free(zw);
    return conv ? 0 : 1;
}
