// irls_round.cuh -- interface of the multi-fit IRLS round (irls_round.cu): one launch advances EVERY
// live L1-logistic fit of a cv.glmnet batch by one IRLS step.  Fits of the same cv.glmnet problem
// (the full-data fit and its nfolds cv-fold fits, ate_functions.R:139-141) read the same rows; a CTA
// stages a slot of rows once and serves up to two fits from it, and CTAs that serve other fits of
// the same problem stream the same rows at the same time (L2 serves the repeats).
#pragma once
#include <vector>

#include "common.cuh"

// Lists of at least this many columns end with the staged rows' ONE column (RoundFit::ones): such a list
// is served by a CTA of its own, whose helpers then compute X^T r only.
constexpr int R_ONES_MIN = 65;

// one cv.glmnet problem = one set of rows, as this rank holds it
struct RoundProblem {
    const double *X;      // private contiguous copy of the rows (rows == NULL, ld == ncs + 4: a slot
    long long ld;         //   is ONE bulk copy) or the shared design read through a row list
    const int32_t *rows;  // device row ids or NULL
    long long nrows;      // rows held by this rank
    int padded;           // 1: a private copy whose rows end with the pad [0, 1, 0, 0] (zero and ONE column)
};

// one path fit, this round
struct RoundFit {
    int prob;                          // index into the problems
    long long hole_begin, hole_len;    // positions [hole_begin, hole_begin + hole_len) of the problem's
                                       // rows carry weight 0 in this fit (its held-out inner fold)
    double w;                          // observation weight 1 / nt (nt over ALL ranks)
    const double *coef;                // device: ni raw-scale coefficients, then the intercept
    const int *colidx;                 // device: the Gram's column list, padded to a multiple of 32 with ncs
    int nG;                            // columns in the list (0: evaluation only)
    int ones;                          // >= 0: list position ones holds the staged rows' ONE column (index
                                       // ncs + 1); X^T v of the list columns is then row `ones` of the Gram and
                                       // the helpers skip it.  -1: the helpers compute X^T v
    int live;                          // 0: skip this fit
};

// where one fit's results of the round sit inside the packed buffer (doubles from its start)
struct RoundSlot {
    long long sums_off;  // X^T r [ni], X^T v [ni], sum v x^2 [ni] (init rounds), sum v, sum r, deviance
    long long H_off;     // ldh x ldh Gram over the list, both triangles
    int ldh;             // round_up(nG, 32)
};

struct RoundStats {
    int ctas = 0;
    long long rows_staged = 0;      // sum over CTAs of the rows they staged
    long long row_fits = 0;         // sum over live fits of the rows they saw
    double flops = 0.0, alg_bytes = 0.0;
};

// Lays the round out (one CTA descriptor per SM, MMA warps dealt over the fits' Gram tiles), launches
// the step kernel and the fold kernel.  `packed` (device, grow-only, owned by the ctx cache) receives,
// for every live fit, its sums and its Gram at slots[fit]; *packed_count = doubles in use (what a
// multi-GPU driver all-reduces).  ncs: columns staged per row (multiple of 8 covering [0, ni) and iy).
// bank (0 / 1): which set of cached buffers the call uses -- two calls whose results are alive at the
// same time (the two launches of one round) take different banks.
int ate_i_irls_round(ate_ctx *ctx, int ncs, int ni, int iy, int init, int bank, const std::vector<RoundProblem> &probs,
                     const std::vector<RoundFit> &fits, std::vector<RoundSlot> &slots, double **packed,
                     long long *packed_count, RoundStats *stats);
