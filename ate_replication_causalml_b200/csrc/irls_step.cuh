// irls_step.cuh -- interface of the fused IRLS-step kernel (irls_step.cu).
#pragma once
#include "common.cuh"

struct IrlsStep {
    const double *X;       // row-major design
    long long ld;
    int ncols;             // columns staged per row: a multiple of 8 covering [0, ni) and iy; also the list's zero slot
    const int32_t *rows;   // row list (NULL = identity) and its logical window
    long long nrows, hole_begin, hole_len;
    int ni, iy;            // features are columns [0, ni); response column iy
    const double *coef;    // device: ni raw-scale coefficients, then the intercept
    double w;              // observation weight 1/nt
    const int *colidx;     // device column list of the Gram, padded to a multiple of 32 with `ncols`
    int nG;                // columns in the list (0: evaluation only)
    double *H;             // device out: nG x nG Gram, compact, leading dimension ldh
    int ldh;
    double *sums;          // device out: X^T r [ni], X^T v [ni], sum v x^2 [ni] (init only), sum v, sum r, deviance
    double *v_out;         // device out: IRLS weights by logical row position
    int init;              // also accumulate sum v x^2 (first step of a fit)
};

int ate_i_irls_step(ate_ctx *ctx, const IrlsStep &in);
// fold the row-chunk partials of a block-lower-triangular Gram into G (both triangles)
int ate_i_gram_fold(ate_ctx *ctx, const double *partials, int nchunks, int nblk, int p, int ldg, double *G);
