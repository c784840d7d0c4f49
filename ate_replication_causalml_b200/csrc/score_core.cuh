// score_core.cuh -- the device code shared by aipw_score_kernel (score.cu) and
// aipw_score_static_kernel (score_static.cu): a7/a8 of ate_functions.R:241-243, :255-256.
#pragma once
#include "common.cuh"

namespace score_core {

constexpr int NV = 6;  // sum est1 (non-NaN), count, sum est2, sum J, sum J^2, (spare)

// one observation: est1, est2 (ate_functions.R:241-242) and the summand of :255, term by term as
// written there; mean(est1, na.rm = TRUE) drops NaN but keeps +-Inf
__device__ __forceinline__ void one(double wi, double yi, double pi, double a0, double a1, double (&acc)[NV],
                                    double &e1, double &e2) {
    e1 = wi * (yi - a1) / pi + (1.0 - wi) * (yi - a0) / (1.0 - pi);
    e2 = a1 - a0;
    const double J = (wi * yi) / pi - a1 * (wi - pi) / pi -
                     (((1.0 - wi) * yi / (1.0 - pi)) + (a0 * (wi - pi) / (1.0 - pi)));
    if (!isnan(e1)) {
        acc[0] += e1;
        acc[1] += 1.0;
    }
    acc[2] += e2;
    acc[3] += J;
    acc[4] += J * J;
}

// grid-stride pass over the five vectors.  VEC: all pointers 16-byte aligned -> 16-byte loads and
// stores, two observations per step.  est1 / est2 / psi (= est1 + est2) are optional outputs.
template <bool VEC>
__device__ __forceinline__ void accumulate(long long n, const double *__restrict__ w, const double *__restrict__ y,
                                           const double *__restrict__ p, const double *__restrict__ m0,
                                           const double *__restrict__ m1, double *__restrict__ est1,
                                           double *__restrict__ est2, double *__restrict__ psi, double (&acc)[NV]) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (VEC) {
        const long long n2 = n >> 1;
        for (long long i = tid; i < n2; i += stride) {
            const double2 wi = __ldg(reinterpret_cast<const double2 *>(w) + i), yi = __ldg(reinterpret_cast<const double2 *>(y) + i);
            const double2 pi = __ldg(reinterpret_cast<const double2 *>(p) + i);
            const double2 a0 = __ldg(reinterpret_cast<const double2 *>(m0) + i), a1 = __ldg(reinterpret_cast<const double2 *>(m1) + i);
            double2 e1, e2;
            one(wi.x, yi.x, pi.x, a0.x, a1.x, acc, e1.x, e2.x);
            one(wi.y, yi.y, pi.y, a0.y, a1.y, acc, e1.y, e2.y);
            if (est1) reinterpret_cast<double2 *>(est1)[i] = e1;
            if (est2) reinterpret_cast<double2 *>(est2)[i] = e2;
            if (psi) reinterpret_cast<double2 *>(psi)[i] = make_double2(e1.x + e2.x, e1.y + e2.y);
        }
        if ((n & 1) && tid == 0) {
            const long long i = n - 1;
            double e1, e2;
            one(__ldg(w + i), __ldg(y + i), __ldg(p + i), __ldg(m0 + i), __ldg(m1 + i), acc, e1, e2);
            if (est1) est1[i] = e1;
            if (est2) est2[i] = e2;
            if (psi) psi[i] = e1 + e2;
        }
    } else {
        for (long long i = tid; i < n; i += stride) {
            double e1, e2;
            one(__ldg(w + i), __ldg(y + i), __ldg(p + i), __ldg(m0 + i), __ldg(m1 + i), acc, e1, e2);
            if (est1) est1[i] = e1;
            if (est2) est2[i] = e2;
            if (psi) psi[i] = e1 + e2;
        }
    }
}

// block partials -> scratch; the last block to finish folds them in block order (reproducible) and
// writes out[0..4] = tau, se, sum est1, count of non-NaN est1, sum est2.  red: NV * 32 doubles of
// shared memory; is_last: one shared flag.
__device__ __forceinline__ void finish(double (&acc)[NV], double *red, bool *is_last, long long n,
                                       double *__restrict__ partials, unsigned int *__restrict__ ticket,
                                       double *__restrict__ out) {
    block_sum<NV>(acc, red);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < NV; ++k) partials[(size_t)blockIdx.x * NV + k] = acc[k];
        __threadfence();
        *is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!*is_last) return;
    __threadfence();
    if (threadIdx.x < NV) {
        double s = 0.0;
        for (unsigned int b = 0; b < gridDim.x; ++b) s += ((volatile double *)partials)[(size_t)b * NV + threadIdx.x];
        red[threadIdx.x] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const double dn = (double)n;
        const double tau = red[0] / red[1] + red[2] / dn;
        // sum (J - tau)^2 = sum J^2 - 2 tau sum J + n tau^2
        double ss = red[4] - 2.0 * tau * red[3] + dn * tau * tau;
        if (ss < 0.0) ss = 0.0;
        out[0] = tau;
        out[1] = sqrt(ss) / dn;
        out[2] = red[0];
        out[3] = red[1];
        out[4] = red[2];
        *ticket = 0u;  // ready for the next launch
    }
}

__host__ inline bool aligned16(const void *a, const void *b, const void *c, const void *d, const void *e,
                               const void *f = nullptr, const void *g = nullptr, const void *h = nullptr) {
    const uintptr_t m = (uintptr_t)a | (uintptr_t)b | (uintptr_t)c | (uintptr_t)d | (uintptr_t)e | (uintptr_t)f |
                        (uintptr_t)g | (uintptr_t)h;
    return (m & 15u) == 0;
}

}  // namespace score_core
