// design.cu -- a1 standardisation (ate_replication.Rmd:72-74) and a2 degree-2 expansion
// (ate_functions.R:288-296), plus the layout change into the engine's row-major design.
//
// HBM-bound passes.  Standardise: one moments pass (shifted sums, so a single read gives both
// mean and centred sum of squares) + one read/write pass = 3*n*p0*8 B.  Expand: the base
// covariates are read once into shared memory per row tile and every product column is
// written once = n*(p0 + p)*8 B.
//
// Device layouts
//   col-major  x[j*n + i]      -- what R hands over and gets back
//   row-major  X[i*ld + c]     -- the engine's design: every row is one contiguous ld*8-byte
//                                 record, which is what the Gram kernel's bulk row copies and
//                                 the IRLS row dot products want.
#include <math.h>

#include "common.cuh"

namespace {

constexpr int MOM_THREADS = 256;

// partial shifted moments of column cols[blockIdx.y]; shift = first element of the column
__global__ void __launch_bounds__(MOM_THREADS)
col_moments_kernel(const double *__restrict__ x, long long n, const int *__restrict__ cols,
                   double *__restrict__ partials /*[ncols][gridDim.x][2]*/) {
    __shared__ double red[2 * 32];
    const int j = cols ? cols[blockIdx.y] : blockIdx.y;
    const double *xj = x + (size_t)j * (size_t)n;
    const double c = __ldg(xj);
    double acc[2] = {0.0, 0.0};
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const double d = __ldg(xj + i) - c;
        acc[0] += d;
        acc[1] += d * d;
    }
    block_sum<2>(acc, red);
    if (threadIdx.x == 0) {
        double *o = partials + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 2;
        o[0] = acc[0];
        o[1] = acc[1];
    }
}

// one warp per column folds the partials in block order -> mean, sd (n-1)
__global__ void col_moments_finish_kernel(const double *__restrict__ x, long long n,
                                          const int *__restrict__ cols,
                                          const double *__restrict__ partials, int nparts,
                                          double *__restrict__ center, double *__restrict__ scale) {
    const int cidx = blockIdx.x;
    const int j = cols ? cols[cidx] : cidx;
    if (threadIdx.x != 0) return;
    double s1 = 0.0, s2 = 0.0;
    for (int b = 0; b < nparts; ++b) {
        s1 += partials[((size_t)cidx * nparts + b) * 2];
        s2 += partials[((size_t)cidx * nparts + b) * 2 + 1];
    }
    const double c = x[(size_t)j * (size_t)n];
    const double dn = (double)n;
    const double mean = c + s1 / dn;
    double ssc = s2 - s1 * s1 / dn;
    if (ssc < 0.0) ssc = 0.0;
    center[cidx] = mean;
    scale[cidx] = sqrt(ssc / (dn - 1.0));
}

__global__ void col_apply_kernel(double *__restrict__ x, long long n, const int *__restrict__ cols,
                                 const double *__restrict__ center,
                                 const double *__restrict__ scale) {
    const int cidx = blockIdx.y;
    const int j = cols ? cols[cidx] : cidx;
    double *xj = x + (size_t)j * (size_t)n;
    const double mu = center[cidx], sd = scale[cidx];
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        xj[i] = (xj[i] - mu) / sd;
}

// ---- expansion, col-major in -> col-major out (public ate_expand_deg2) ----------------------
constexpr int EXP_ROWS = 128;

__global__ void __launch_bounds__(256)
expand_colmajor_kernel(const double *__restrict__ x, long long n, int p0, int include_main,
                       double *__restrict__ out) {
    extern __shared__ double tile[];  // [p0][EXP_ROWS]
    const long long r0 = (long long)blockIdx.x * EXP_ROWS;
    const int rows = (int)((n - r0) < EXP_ROWS ? (n - r0) : EXP_ROWS);
    for (int e = threadIdx.x; e < p0 * EXP_ROWS; e += blockDim.x) {
        const int j = e / EXP_ROWS, r = e % EXP_ROWS;
        tile[e] = r < rows ? __ldg(x + (size_t)j * n + r0 + r) : 0.0;
    }
    __syncthreads();
    const int nmain = include_main ? p0 : 0;
    const int pout = nmain + p0 * p0;
    // consecutive threads -> consecutive rows of one output column (coalesced 8 B stores)
    for (int e = threadIdx.x; e < pout * EXP_ROWS; e += blockDim.x) {
        const int c = e / EXP_ROWS, r = e % EXP_ROWS;
        if (r >= rows) continue;
        double v;
        if (c < nmain)
            v = tile[c * EXP_ROWS + r];
        else {
            const int q = c - nmain, c1 = q / p0, c2 = q % p0;
            v = tile[c1 * EXP_ROWS + r] * tile[c2 * EXP_ROWS + r];
        }
        out[(size_t)c * n + r0 + r] = v;
    }
}

// ---- design builder: col-major base (+ w, y) -> row-major design with appended columns ------
// columns: [mains (nmain)] [products (nprod)] [w] [y] [1] [0 ... to ld]
constexpr int DES_ROWS = 32;

__global__ void __launch_bounds__(256)
build_design_kernel(const double *__restrict__ x, long long n, int p0, int nmain, int nprod,
                    const double *__restrict__ w, const double *__restrict__ y,
                    double *__restrict__ X, long long ld) {
    extern __shared__ double tile[];  // [DES_ROWS][p0 + 2]
    const int tw = p0 + 2;
    const long long r0 = (long long)blockIdx.x * DES_ROWS;
    const int rows = (int)((n - r0) < DES_ROWS ? (n - r0) : DES_ROWS);
    for (int e = threadIdx.x; e < tw * DES_ROWS; e += blockDim.x) {
        const int j = e / DES_ROWS, r = e % DES_ROWS;  // consecutive threads -> consecutive rows
        double v = 0.0;
        if (r < rows) {
            if (j < p0)
                v = __ldg(x + (size_t)j * n + r0 + r);
            else if (j == p0)
                v = w ? __ldg(w + r0 + r) : 0.0;
            else
                v = y ? __ldg(y + r0 + r) : 0.0;
        }
        tile[r * tw + j] = v;
    }
    __syncthreads();
    const int p = nmain + nprod;
    const int ldi = (int)ld;
    for (int e = threadIdx.x; e < rows * ldi; e += blockDim.x) {
        const int r = e / ldi, c = e % ldi;  // consecutive threads -> consecutive columns
        const double *t = tile + r * tw;
        double v;
        if (c < nmain)
            v = t[c];
        else if (c < p) {
            const int q = c - nmain;
            v = t[q / p0] * t[q % p0];
        } else if (c == p)
            v = t[p0];
        else if (c == p + 1)
            v = t[p0 + 1];
        else if (c == p + 2)
            v = 1.0;
        else
            v = 0.0;
        X[(size_t)(r0 + r) * ld + c] = v;
    }
}

// ---- plain col-major -> row-major transpose into columns [col_offset, col_offset + p) ------
__global__ void transpose_kernel(const double *__restrict__ x, long long n, int p,
                                 double *__restrict__ X, long long ld, int col_offset) {
    __shared__ double t[32][33];
    const long long r0 = (long long)blockIdx.x * 32;
    const int c0 = blockIdx.y * 32;
    for (int k = threadIdx.y; k < 32; k += blockDim.y) {
        const long long r = r0 + threadIdx.x;
        const int c = c0 + k;
        t[k][threadIdx.x] = (r < n && c < p) ? __ldg(x + (size_t)c * n + r) : 0.0;
    }
    __syncthreads();
    for (int k = threadIdx.y; k < 32; k += blockDim.y) {
        const long long r = r0 + k;
        const int c = c0 + threadIdx.x;
        if (r < n && c < p) X[(size_t)r * ld + col_offset + c] = t[threadIdx.x][k];
    }
}

// private, contiguous copy of a row subset: Xp[t][0..ncs) = X[list[t]][0..ncs), pad columns zero.
// One warp per row, 16-byte accesses.
__global__ void __launch_bounds__(256)
gather_rows_kernel(const double *__restrict__ X, long long ld, const int32_t *__restrict__ list,
                   long long nt, int ncs, double *__restrict__ Xp, long long ldp) {
    const int lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    const long long wstride = (long long)gridDim.x * nw;
    for (long long t = (long long)blockIdx.x * nw + (threadIdx.x >> 5); t < nt; t += wstride) {
        const long long r = list ? (long long)__ldg(list + t) : t;
        const double2 *src = reinterpret_cast<const double2 *>(X + (size_t)r * (size_t)ld);
        double2 *dst = reinterpret_cast<double2 *>(Xp + (size_t)t * (size_t)ldp);
        for (int c = lane; c < (int)(ldp / 2); c += 32) dst[c] = (2 * c < ncs) ? __ldg(src + c) : make_double2(0.0, 0.0);
    }
}

__global__ void fill_column_kernel(double *__restrict__ X, long long n, long long ld, int col,
                                   double value) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        X[(size_t)i * ld + col] = value;
}

}  // namespace

// ---- internal API ----------------------------------------------------------------------
int ate_i_standardize(ate_ctx *ctx, double *d_x, long long n, int p, const int *d_cols, int ncols,
                      double *d_center, double *d_scale) {
    if (ncols <= 0) return 0;
    int nb = ctx->num_sms * 4 / ncols;
    if (nb < 1) nb = 1;
    long long need = (n + MOM_THREADS - 1) / MOM_THREADS;
    if (need < nb) nb = (int)need;
    void *s = nullptr;
    ATE_TRY(ate_scratch(ctx, sizeof(double) * 2 * (size_t)nb * ncols, &s));
    double *partials = (double *)s;
    col_moments_kernel<<<dim3(nb, ncols), MOM_THREADS, 0, ctx->stream>>>(d_x, n, d_cols, partials);
    ATE_LAUNCHED(ctx);
    col_moments_finish_kernel<<<ncols, 32, 0, ctx->stream>>>(d_x, n, d_cols, partials, nb, d_center, d_scale);
    ATE_LAUNCHED(ctx);
    col_apply_kernel<<<dim3(nb, ncols), MOM_THREADS, 0, ctx->stream>>>(d_x, n, d_cols, d_center, d_scale);
    ATE_LAUNCHED(ctx);
    return 0;
}

int ate_i_build_design(ate_ctx *ctx, const double *d_xbase, long long n, int p0, int expand,
                       const double *d_w, const double *d_y, double *d_X, long long ld) {
    const int nmain = (expand == 1) ? 0 : p0;
    const int nprod = (expand == 0) ? 0 : p0 * p0;
    if (nmain + nprod + 3 > ld) return ate_fail(ctx, -1, "build_design: ld too small");
    const size_t smem = sizeof(double) * (size_t)(p0 + 2) * DES_ROWS;
    if (smem > 48 * 1024)
        ATE_CUDA(ctx, cudaFuncSetAttribute(build_design_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long blocks = (n + DES_ROWS - 1) / DES_ROWS;
    build_design_kernel<<<(unsigned)blocks, 256, smem, ctx->stream>>>(d_xbase, n, p0, nmain, nprod, d_w, d_y, d_X, ld);
    ATE_LAUNCHED(ctx);
    return 0;
}

int ate_i_transpose_to_rowmajor(ate_ctx *ctx, const double *x_colmajor, long long n, int p,
                                double *X, long long ld, int col_offset) {
    dim3 grid((unsigned)((n + 31) / 32), (unsigned)((p + 31) / 32));
    transpose_kernel<<<grid, dim3(32, 8), 0, ctx->stream>>>(x_colmajor, n, p, X, ld, col_offset);
    ATE_LAUNCHED(ctx);
    return 0;
}

int ate_i_gather_rows(ate_ctx *ctx, const double *X, long long ld, const int32_t *list, long long nt,
                      int ncs, double *Xp, long long ldp) {
    if (nt <= 0) return 0;
    if (ncs % 2 != 0 || ldp % 2 != 0 || ldp < ncs || ncs > ld) return ate_fail(ctx, -1, "gather_rows: bad widths");
    int blocks = ctx->num_sms * 8;
    if ((nt + 7) / 8 < blocks) blocks = (int)((nt + 7) / 8);
    gather_rows_kernel<<<blocks, 256, 0, ctx->stream>>>(X, ld, list, nt, ncs, Xp, ldp);
    ATE_LAUNCHED(ctx);
    return 0;
}

int ate_i_fill_column(ate_ctx *ctx, double *X, long long n, long long ld, int col, double value) {
    fill_column_kernel<<<ctx->num_sms * 2, 256, 0, ctx->stream>>>(X, n, ld, col, value);
    ATE_LAUNCHED(ctx);
    return 0;
}

// ---- public entry points ---------------------------------------------------------------
extern "C" int ate_standardize(ate_ctx *ctx, double *x, long long n, int p, const int *cols,
                               int ncols, double *center, double *scale) {
    if (!ctx) return -1;
    if (!x || n < 2 || p <= 0) return ate_fail(ctx, -1, "ate_standardize: bad arguments");
    ATE_CUDA(ctx, cudaSetDevice(ctx->device));
    const int m = cols ? ncols : p;
    if (m <= 0) return 0;
    for (int c = 0; cols && c < m; ++c)
        if (cols[c] < 0 || cols[c] >= p) return ate_fail(ctx, -1, "ate_standardize: column %d out of range", cols[c]);
    DevBuf<double> dx, dcs;
    DevBuf<int> dcols;
    const size_t N = (size_t)n * p;
    ATE_CUDA(ctx, dx.alloc(N));
    ATE_CUDA(ctx, dcs.alloc((size_t)2 * m));
    ATE_CUDA(ctx, cudaMemcpyAsync(dx.p, x, N * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    if (cols) {
        ATE_CUDA(ctx, dcols.alloc(m));
        ATE_CUDA(ctx, cudaMemcpyAsync(dcols.p, cols, sizeof(int) * m, cudaMemcpyHostToDevice, ctx->stream));
    }
    ATE_TRY(ate_i_standardize(ctx, dx.p, n, p, cols ? dcols.p : nullptr, m, dcs.p, dcs.p + m));
    ATE_CUDA(ctx, cudaMemcpyAsync(x, dx.p, N * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    if (center) ATE_CUDA(ctx, cudaMemcpyAsync(center, dcs.p, sizeof(double) * m, cudaMemcpyDeviceToHost, ctx->stream));
    if (scale) ATE_CUDA(ctx, cudaMemcpyAsync(scale, dcs.p + m, sizeof(double) * m, cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" int ate_expand_deg2(ate_ctx *ctx, const double *x, long long n, int p0, int include_main,
                               double *out) {
    if (!ctx) return -1;
    if (!x || !out || n <= 0 || p0 <= 0) return ate_fail(ctx, -1, "ate_expand_deg2: bad arguments");
    ATE_CUDA(ctx, cudaSetDevice(ctx->device));
    const int pout = p0 * p0 + (include_main ? p0 : 0);
    DevBuf<double> dx, dout;
    ATE_CUDA(ctx, dx.alloc((size_t)n * p0));
    ATE_CUDA(ctx, dout.alloc((size_t)n * pout));
    ATE_CUDA(ctx, cudaMemcpyAsync(dx.p, x, sizeof(double) * (size_t)n * p0, cudaMemcpyHostToDevice, ctx->stream));
    const size_t smem = sizeof(double) * (size_t)p0 * EXP_ROWS;
    if (smem > 200 * 1024) return ate_fail(ctx, -1, "ate_expand_deg2: p0 = %d too large", p0);
    if (smem > 48 * 1024)
        ATE_CUDA(ctx, cudaFuncSetAttribute(expand_colmajor_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long blocks = (n + EXP_ROWS - 1) / EXP_ROWS;
    expand_colmajor_kernel<<<(unsigned)blocks, 256, smem, ctx->stream>>>(dx.p, n, p0, include_main, dout.p);
    ATE_LAUNCHED(ctx);
    ATE_CUDA(ctx, cudaMemcpyAsync(out, dout.p, sizeof(double) * (size_t)n * pout, cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}
