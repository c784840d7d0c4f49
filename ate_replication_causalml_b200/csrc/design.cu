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

// partial shifted moments of column cols[blockIdx.y]; shift = first element of the column.
// The column is walked in 16-byte pairs from its first 16-byte-aligned element; the (at most one)
// element before it and the odd tail are handled by block 0 / the last pair's owner.
__global__ void __launch_bounds__(MOM_THREADS)
col_moments_kernel(const double *__restrict__ x, long long n, const int *__restrict__ cols,
                   double *__restrict__ partials /*[ncols][gridDim.x][2]*/) {
    __shared__ double red[2 * 32];
    const int j = cols ? cols[blockIdx.y] : blockIdx.y;
    const double *xj = x + (size_t)j * (size_t)n;
    const double c = __ldg(xj);
    double acc[2] = {0.0, 0.0};
    const long long head = (long long)(((uintptr_t)xj >> 3) & 1);  // elements before the aligned body
    const double2 *x2 = reinterpret_cast<const double2 *>(xj + head);
    const long long n2 = (n - head) >> 1;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
        const double2 v = __ldg(x2 + i);
        const double d0 = v.x - c, d1 = v.y - c;
        acc[0] += d0 + d1;
        acc[1] += d0 * d0 + d1 * d1;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (head) {
            const double d = __ldg(xj) - c;
            acc[0] += d;
            acc[1] += d * d;
        }
        if (head + 2 * n2 < n) {
            const double d = __ldg(xj + n - 1) - c;
            acc[0] += d;
            acc[1] += d * d;
        }
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

__global__ void __launch_bounds__(MOM_THREADS)
col_apply_kernel(double *__restrict__ x, long long n, const int *__restrict__ cols,
                 const double *__restrict__ center, const double *__restrict__ scale) {
    const int cidx = blockIdx.y;
    const int j = cols ? cols[cidx] : cidx;
    double *xj = x + (size_t)j * (size_t)n;
    const double mu = center[cidx], sd = scale[cidx];
    const long long head = (long long)(((uintptr_t)xj >> 3) & 1);
    double2 *x2 = reinterpret_cast<double2 *>(xj + head);
    const long long n2 = (n - head) >> 1;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
        double2 v = x2[i];
        v.x = (v.x - mu) / sd;
        v.y = (v.y - mu) / sd;
        x2[i] = v;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (head) xj[0] = (xj[0] - mu) / sd;
        if (head + 2 * n2 < n) xj[n - 1] = (xj[n - 1] - mu) / sd;
    }
}

// ---- expansion, col-major in -> col-major out (public ate_expand_deg2) ----------------------
constexpr int EXP_ROWS = 128;

__global__ void __launch_bounds__(256)
expand_colmajor_kernel(const double *__restrict__ x, long long n, int p0, int include_main,
                       double *__restrict__ out) {
    extern __shared__ double tile[];  // [p0][EXP_ROWS]
    const long long r0 = (long long)blockIdx.x * EXP_ROWS;
    const int rows = (int)((n - r0) < EXP_ROWS ? (n - r0) : EXP_ROWS);
    for (int j = threadIdx.x / EXP_ROWS; j < p0; j += 256 / EXP_ROWS) {
        const int r = threadIdx.x % EXP_ROWS;  // powers of two: shifts
        tile[j * EXP_ROWS + r] = r < rows ? __ldg(x + (size_t)j * n + r0 + r) : 0.0;
    }
    __syncthreads();
    const int nmain = include_main ? p0 : 0;
    const int pout = nmain + p0 * p0;
    // a thread owns a fixed pair of rows and walks the output columns; (c1, c2) advance by counters
    constexpr int PAIRS = EXP_ROWS / 2, CSTEP = 256 / PAIRS;
    const int rp = threadIdx.x % PAIRS, r = 2 * rp;
    int c = threadIdx.x / PAIRS;
    int q = c - nmain, c1 = 0, c2 = 0;
    if (q >= 0) {
        c1 = q / p0;  // once per thread
        c2 = q - c1 * p0;
    }
    const bool vec = ((n & 1) == 0);  // every column then starts 16-byte aligned (r0 and r are even)
    for (; c < pout; c += CSTEP) {
        double v0, v1;
        if (c < nmain) {
            v0 = tile[c * EXP_ROWS + r];
            v1 = tile[c * EXP_ROWS + r + 1];
            if (c + CSTEP >= nmain) {  // entering the products
                q = c + CSTEP - nmain;
                c1 = q / p0;
                c2 = q - c1 * p0;
            }
        } else {
            v0 = tile[c1 * EXP_ROWS + r] * tile[c2 * EXP_ROWS + r];
            v1 = tile[c1 * EXP_ROWS + r + 1] * tile[c2 * EXP_ROWS + r + 1];
            c2 += CSTEP;
            while (c2 >= p0) {
                c2 -= p0;
                ++c1;
            }
        }
        double *o = out + (size_t)c * n + r0 + r;
        if (vec && r + 1 < rows)
            *reinterpret_cast<double2 *>(o) = make_double2(v0, v1);
        else {
            if (r < rows) o[0] = v0;
            if (r + 1 < rows) o[1] = v1;
        }
    }
}

// ---- design builder: col-major base (+ w, y) -> row-major design with appended columns ------
// columns: [mains (nmain)] [products (nprod)] [w] [y] [1] [0 ... to ld]
// Every output column is the product of two entries of the staged row [base (p0) | w | y | 1 | 0]:
// a main is base * 1, the appended columns are w * 1, y * 1, 1 * 1, the padding 0 * 0.  The two
// factors of each column are looked up in a table built once per block (no division per element),
// and a thread writes two adjacent columns with one 16-byte store.
constexpr int DES_ROWS = 32;

__global__ void __launch_bounds__(256)
build_design_kernel(const double *__restrict__ x, long long ldx, long long n, int p0, int nmain, int nprod,
                    const double *__restrict__ w, const double *__restrict__ y,
                    double *__restrict__ X, long long ld) {
    extern __shared__ double tile[];  // [p0 + 4][DES_ROWS + 1], then the factor table
    const int tw = p0 + 4, tp = DES_ROWS + 1;
    short2 *lut = reinterpret_cast<short2 *>(tile + (size_t)tw * tp);
    const long long r0 = (long long)blockIdx.x * DES_ROWS;
    const int rows = (int)((n - r0) < DES_ROWS ? (n - r0) : DES_ROWS);
    for (int e = threadIdx.x; e < tw * DES_ROWS; e += blockDim.x) {
        const int j = e / DES_ROWS, r = e % DES_ROWS;  // consecutive threads -> consecutive rows
        double v = 0.0;
        if (j == p0 + 2)
            v = 1.0;
        else if (r < rows && j < p0 + 2) {
            if (j < p0)
                v = __ldg(x + (size_t)j * ldx + r0 + r);
            else if (j == p0)
                v = w ? __ldg(w + r0 + r) : 0.0;
            else
                v = y ? __ldg(y + r0 + r) : 0.0;
        }
        tile[j * tp + r] = v;
    }
    const int p = nmain + nprod, ldi = (int)ld, ONE = p0 + 2, ZERO = p0 + 3;
    for (int c = threadIdx.x; c < ldi; c += blockDim.x) {
        short2 f;
        if (c < nmain)
            f = make_short2((short)c, (short)ONE);
        else if (c < p) {
            const int q = c - nmain;
            f = make_short2((short)(q / p0), (short)(q % p0));
        } else if (c == p)
            f = make_short2((short)p0, (short)ONE);
        else if (c == p + 1)
            f = make_short2((short)(p0 + 1), (short)ONE);
        else if (c == p + 2)
            f = make_short2((short)ONE, (short)ONE);
        else
            f = make_short2((short)ZERO, (short)ZERO);
        lut[c] = f;
    }
    __syncthreads();
    const int half = ldi >> 1;  // ld is a multiple of 32
    for (int e = threadIdx.x; e < rows * half; e += blockDim.x) {
        const int r = e / half, c = 2 * (e - r * half);  // one division per 16-byte store
        const short2 f0 = lut[c], f1 = lut[c + 1];
        const double v0 = tile[f0.x * tp + r] * tile[f0.y * tp + r];
        const double v1 = tile[f1.x * tp + r] * tile[f1.y * tp + r];
        *reinterpret_cast<double2 *>(X + (size_t)(r0 + r) * ld + c) = make_double2(v0, v1);
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
        // beyond the staged columns: the lists' zero column (ncs) and ONE column (ncs + 1), then zeros
        for (int c = lane; c < (int)(ldp / 2); c += 32)
            dst[c] = (2 * c < ncs) ? __ldg(src + c) : make_double2(0.0, 2 * c == ncs ? 1.0 : 0.0);
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

int ate_i_build_design(ate_ctx *ctx, const double *d_xbase, long long ldx, long long n, int p0, int expand,
                       const double *d_w, const double *d_y, double *d_X, long long ld) {
    const int nmain = (expand == 1) ? 0 : p0;
    const int nprod = (expand == 0) ? 0 : p0 * p0;
    if (nmain + nprod + 3 > ld || (ld & 31) != 0) return ate_fail(ctx, -1, "build_design: bad ld");
    if (p0 + 4 > 32000) return ate_fail(ctx, -1, "build_design: p0 too large");
    const size_t smem = sizeof(double) * (size_t)(p0 + 4) * (DES_ROWS + 1) + sizeof(short2) * (size_t)ld;
    if (smem > 200 * 1024) return ate_fail(ctx, -1, "build_design: p0 = %d too large", p0);
    if (smem > 48 * 1024)
        ATE_CUDA(ctx, cudaFuncSetAttribute(build_design_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long blocks = (n + DES_ROWS - 1) / DES_ROWS;
    build_design_kernel<<<(unsigned)blocks, 256, smem, ctx->stream>>>(d_xbase, ldx, n, p0, nmain, nprod, d_w, d_y, d_X, ld);
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
