// path.cuh -- internal interfaces of the lambda-path solvers (gaussian + binomial) and of the
// cv.glmnet orchestration built on them.
#pragma once
#include <vector>

#include "common.cuh"

// glmnet.control() constants (glmnet 2.0-x)
#define GLM_FDEV 1.0e-5
#define GLM_DEVMAX 0.999
#define GLM_EPS 1.0e-6
#define GLM_BIG 9.9e35
#define GLM_MNLAM 5
#define GLM_PMIN 1.0e-9

// One cv.glmnet call on a subset of the rows of the row-major design.
//   d_list : device row ids of the subset, grouped by inner fold (segment f = rows whose
//            cv.glmnet foldid is f+1).  NULL = rows 0..nt-1 in order (only with nseg <= 1).
//   seg    : nseg+1 offsets into d_list.  nseg == 0: plain glmnet() on the subset, no CV.
// When several ranks share the fit (PathOpts::shard_count > 1) every rank holds a slice of the rows:
// d_list / nt / seg describe THIS rank's rows; nt_global / seg_global (fold sizes) the whole problem.
struct CvProblem {
    const int32_t *d_list = nullptr;
    long long nt = 0;
    int nseg = 0;
    std::vector<long long> seg;
    long long nt_global = 0;              // 0: same as nt
    std::vector<long long> seg_global;    // nseg fold sizes over all ranks; empty: same as seg
};

struct CvOut {
    int lmu = 0, idx_min = 0, idx_1se = 0, jerr = 0, nlp = 0;
    double nulldev = 0.0;
    std::vector<double> lambda, a0, beta /* ni x lmu, column per lambda */, cvm, cvsd, dev;
};

struct PathOpts {
    int nlam = 100;
    double flmin = 1.0e-4;         // lambda.min.ratio; >= 1 means "use ulam"
    const double *ulam = nullptr;  // host, decreasing
    int nulam = 0;
    double thr = 1.0e-7;
    int maxit = 100000;
    int shard_rank = 0, shard_count = 1;  // ranks sharing the fit, each holding a slice of the rows
};

// Sum a device buffer of doubles over the ranks sharing the fit, in place, on the ctx stream: NCCL when
// the ctx has a communicator (ate_ctx_nccl_init), else through the host hook (ate_ctx_set_allreduce).
int ate_i_allreduce(ate_ctx *ctx, double *dbuf, long long count);
// every rank contributes `chunk` doubles at dsend; drecv receives nranks * chunk doubles, rank-major
int ate_i_allgather(ate_ctx *ctx, const double *dsend, double *drecv, long long chunk, int rank, int nranks);

// X: row-major design, features are columns [0, ni); iy / i1: response and ones columns.
int ate_i_cv_gaussian(ate_ctx *ctx, const double *X, long long ld, int ni, int iy, int i1,
                      const double *vp_host, const PathOpts &po, const std::vector<CvProblem> &probs,
                      std::vector<CvOut> &outs);
struct IrlsStats {
    int rounds = 0;
    double solves = 0.0, sum_strong = 0.0, sum_strong2 = 0.0;
    double step_launches = 0.0, row_fits = 0.0, rows_staged = 0.0, list_misses = 0.0;
};
int ate_i_cv_binomial(ate_ctx *ctx, const double *X, long long ld, int ni, int iy,
                      const double *vp_host, const PathOpts &po, const std::vector<CvProblem> &probs,
                      std::vector<CvOut> &outs, IrlsStats *stats);

// cvm / cvsd / getmin from per-fold means (host).  cvk[f*lmu + l], wk[f] fold sizes, nlk[f] number
// of lambdas fold f returned.
void ate_host_cv_summary(const std::vector<double> &cvk, const std::vector<double> &wk,
                         const std::vector<int> &nlk, int nfolds, int lmu, const double *lambda,
                         std::vector<double> &cvm, std::vector<double> &cvsd, int *idx_min,
                         int *idx_1se);

// eta = a0 + x_row . beta for the rows of a list; link / response / both-arm outcome prediction.
//   mode 0: out0[row] = eta                      (link)
//   mode 1: out0[row] = plogis(eta)              (binomial response)
//   mode 2: out0[row] = eta, out1[row] = eta + extra   (m0 / m1 with W := 0 / 1; extra = beta_W)
// rows NULL = 0..nrows-1.  `scatter`: write at out[row id] (stitched) instead of out[t].
int ate_i_predict_rows(ate_ctx *ctx, const double *X, long long ld, int ni, const int32_t *rows,
                       long long row_begin, long long nrows, const double *d_beta, double a0, double extra,
                       int mode, int scatter, double *out0, double *out1);
