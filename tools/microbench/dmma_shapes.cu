// dmma_shapes.cu -- throughput of the fp64 mma.sync shapes on sm_100a, per warp and per SM.
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o dmma_shapes dmma_shapes.cu
// Prints, for each shape and warps-per-SM, cycles per instruction (per warp) and FMA/clk/SM.
#include <cstdio>
#include <cuda_runtime.h>

template <int SHAPE>
__device__ __forceinline__ void mma(double (&c)[4], const double (&a)[8], const double (&b)[4]) {
    if constexpr (SHAPE == 0) {
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                     : "+d"(c[0]), "+d"(c[1]) : "d"(a[0]), "d"(b[0]));
    } else if constexpr (SHAPE == 1) {
        asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                     : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3]) : "d"(a[0]), "d"(a[1]), "d"(b[0]));
    } else if constexpr (SHAPE == 2) {
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
                     : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
    } else {
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, "
                     "{%12,%13,%14,%15}, {%0,%1,%2,%3};"
                     : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
                     : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]), "d"(b[0]),
                       "d"(b[1]), "d"(b[2]), "d"(b[3]));
    }
}

template <int SHAPE, int NACC>
__global__ void bench(int iters, long long *cycles, double *sink) {
    double a[8], b[4], c[NACC][4];
    for (int i = 0; i < 8; ++i) a[i] = 1.0 + 1e-9 * (threadIdx.x + i);
    for (int i = 0; i < 4; ++i) b[i] = 1.0 - 1e-9 * (threadIdx.x + i);
    for (int k = 0; k < NACC; ++k)
        for (int i = 0; i < 4; ++i) c[k][i] = 0.0;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < NACC; ++k) mma<SHAPE>(c[k], a, b);
    }
    const long long t1 = clock64();
    double s = 0.0;
    for (int k = 0; k < NACC; ++k)
        for (int i = 0; i < 4; ++i) s += c[k][i];
    if (s == 12345.678) sink[0] = s;
    if ((threadIdx.x & 31) == 0) cycles[blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)] = t1 - t0;
}

template <int SHAPE, int NACC>
void run(const char *name, int fma_per_inst) {
    long long *d_cycles;
    double *d_sink;
    cudaMalloc(&d_cycles, sizeof(long long) * 148 * 32);
    cudaMalloc(&d_sink, 8);
    const int iters = 2000;
    for (int warps : {1, 4, 8, 12, 16}) {
        bench<SHAPE, NACC><<<148, 32 * warps>>>(10, d_cycles, d_sink);
        bench<SHAPE, NACC><<<148, 32 * warps>>>(iters, d_cycles, d_sink);
        if (cudaDeviceSynchronize() != cudaSuccess) { printf("%s: launch failed\n", name); return; }
        long long h[148 * 32];
        cudaMemcpy(h, d_cycles, sizeof(long long) * 148 * warps, cudaMemcpyDeviceToHost);
        long long mx = 0;
        for (int i = 0; i < 148 * warps; ++i) mx = h[i] > mx ? h[i] : mx;
        const double per_inst = (double)mx / ((double)iters * NACC);
        printf("%-10s acc=%d warps/SM=%2d  cycles/inst/warp=%7.2f  FMA/clk/SM=%7.2f\n", name, NACC, warps, per_inst,
               (double)fma_per_inst * warps / per_inst);
    }
    cudaFree(d_cycles);
    cudaFree(d_sink);
}

int main() {
    run<0, 8>("m8n8k4", 256);
    run<0, 16>("m8n8k4", 256);
    run<1, 8>("m16n8k4", 512);
    run<2, 8>("m16n8k8", 1024);
    run<3, 4>("m16n8k16", 2048);
    run<3, 8>("m16n8k16", 2048);
    return 0;
}
