// mma_loop.cu -- what the Gram loop of irls_round_kernel can reach on its own: W warps per SM each run the
// k-step loop of one or two "halves" (A operands of two tile rows, B operands of four column tiles, all
// gathered from a staged slot in shared memory, v-scaling on the A side) with nothing else on the SM.
// Prints cycles per DMMA per sub-partition (16.7 = the pipe's peak) for several loop shapes.
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o mma_loop mma_loop.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

constexpr int ROWS = 32, PITCH = 268;

// HALVES: 1 = eight accumulators per k-step (one half), 2 = sixteen (two halves fused in one loop body)
// LOADS: operands from shared memory (1) or registers (0); SCALE: DMULs per half and k-step (0 or 2)
// POLL: warps beyond `mma_warps` spin on an mbarrier that never completes (what the producer, helper and
// eval warps of the real kernel do while they wait), until the MMA warps are done
template <int HALVES, int LOADS, int SCALE, int UNROLL>
__global__ void __launch_bounds__(512, 1) loop_kernel(int slots, long long *cycles, double *sink, int mma_warps) {
    extern __shared__ double sx[];
    __shared__ uint64_t bar;
    __shared__ volatile int done_flag;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&bar)), "r"(1));
        done_flag = 0;
    }
    for (int e = threadIdx.x; e < 3 * ROWS * PITCH + 3 * ROWS; e += blockDim.x) sx[e] = 1.0 + 1e-9 * e;
    __syncthreads();
    const int lane = threadIdx.x & 31, g = lane >> 2, k4 = lane & 3, warp = threadIdx.x >> 5;
    double acc[HALVES][2][4][2];
#pragma unroll
    for (int h = 0; h < HALVES; ++h)
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int n = 0; n < 4; ++n) acc[h][j][n][0] = acc[h][j][n][1] = 0.0;
    uint32_t offA[HALVES][2], offB[HALVES][4];
#pragma unroll
    for (int h = 0; h < HALVES; ++h) {
        // column lists with every residue mod 4 twice per eight entries, as the solve kernel orders them
        for (int j = 0; j < 2; ++j) offA[h][j] = (uint32_t)(((warp * 16 + h * 64 + j * 8 + g * 5) % 256) + k4 * PITCH) * 8u;
        for (int n = 0; n < 4; ++n) offB[h][n] = (uint32_t)(((warp * 8 + h * 32 + n * 8 + g * 5 + 3) % 256) + k4 * PITCH) * 8u;
    }
    const double *sv = sx + 3 * ROWS * PITCH;
    __syncthreads();
    if (warp >= mma_warps) {
        for (unsigned spins = 1;; ++spins) {
            uint32_t ok;
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok) : "r"((uint32_t)__cvta_generic_to_shared(&bar)), "r"(0u) : "memory");
            if (ok || done_flag >= mma_warps) break;
        }
        return;
    }
    const long long t0 = clock64();
    for (int it = 0; it < slots; ++it) {
        const char *sbase = reinterpret_cast<const char *>(sx + (size_t)(it % 3) * ROWS * PITCH);
        const double *vs = sv + (it % 3) * ROWS + k4;
#pragma unroll(UNROLL)
        for (int kk = 0; kk < ROWS / 4; ++kk) {
            const char *rowp = sbase + (size_t)kk * (PITCH * 32);
#pragma unroll
            for (int h = 0; h < HALVES; ++h) {
                double fb[4], fa[2];
                const double vv = LOADS ? vs[kk * 4] : 1.0 + 1e-9 * kk;
#pragma unroll
                for (int n = 0; n < 4; ++n) fb[n] = LOADS ? *reinterpret_cast<const double *>(rowp + offB[h][n]) : 1.0 - 1e-9 * (n + kk);
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    fa[j] = LOADS ? *reinterpret_cast<const double *>(rowp + offA[h][j]) : 1.0 + 1e-9 * (j + kk);
                    if (SCALE) fa[j] *= vv;
                }
#pragma unroll
                for (int j = 0; j < 2; ++j)
#pragma unroll
                    for (int n = 0; n < 4; ++n) dmma884(acc[h][j][n][0], acc[h][j][n][1], fa[j], fb[n]);
            }
        }
        __syncwarp();
    }
    const long long t1 = clock64();
    if (lane == 0) atomicAdd((int *)&done_flag, 1);
    double s = 0.0;
#pragma unroll
    for (int h = 0; h < HALVES; ++h)
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int n = 0; n < 4; ++n) s += acc[h][j][n][0] + acc[h][j][n][1];
    if (s == 12345.678) sink[0] = s;
    if (lane == 0) cycles[blockIdx.x * 16 + warp] = t1 - t0;
}

template <int HALVES, int LOADS, int SCALE, int UNROLL>
void run(const char *what, int pollers = 0) {
    long long *d_cycles;
    double *d_sink;
    cudaMalloc(&d_cycles, sizeof(long long) * 148 * 16);
    cudaMalloc(&d_sink, 8);
    const size_t smem = sizeof(double) * (3 * ROWS * PITCH + 3 * ROWS);
    cudaFuncSetAttribute(loop_kernel<HALVES, LOADS, SCALE, UNROLL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int slots = 400;
    for (int warps : {4, 8, 12, 16}) {
        if (warps + pollers > 16) continue;
        loop_kernel<HALVES, LOADS, SCALE, UNROLL><<<148, 32 * (warps + pollers), smem>>>(10, d_cycles, d_sink, warps);
        loop_kernel<HALVES, LOADS, SCALE, UNROLL><<<148, 32 * (warps + pollers), smem>>>(slots, d_cycles, d_sink, warps);
        if (cudaDeviceSynchronize() != cudaSuccess) { printf("%s: launch failed: %s\n", what, cudaGetErrorString(cudaGetLastError())); return; }
        long long h[148 * 16];
        cudaMemcpy(h, d_cycles, sizeof(h), cudaMemcpyDeviceToHost);
        long long mx = 0;
        for (int b = 0; b < 148; ++b)
            for (int w = 0; w < warps; ++w) mx = h[b * 16 + w] > mx ? h[b * 16 + w] : mx;
        const double dmma_per_sp = (double)slots * (ROWS / 4) * 8.0 * HALVES * warps / 4.0;
        printf("%-44s +%d pollers warps/SM=%2d  cycles per DMMA per sub-partition = %6.2f  (%.0f%% of the pipe)\n", what, pollers, warps, mx / dmma_per_sp,
               100.0 * 16.7 / (mx / dmma_per_sp));
    }
    cudaFree(d_cycles);
    cudaFree(d_sink);
}

int main() {
    run<1, 0, 0, 2>("1 half, registers, no scaling, unroll 2");
    run<1, 0, 2, 2>("1 half, registers, 2 DMUL, unroll 2");
    run<1, 1, 0, 2>("1 half, smem operands, no scaling, unroll 2");
    run<1, 1, 2, 2>("1 half, smem operands, 2 DMUL, unroll 2");
    run<1, 1, 2, 8>("1 half, smem operands, 2 DMUL, unroll 8");
    run<2, 1, 2, 1>("2 halves fused, smem operands, 4 DMUL, unroll 1");
    run<2, 1, 2, 2>("2 halves fused, smem operands, 4 DMUL, unroll 2");
    run<1, 1, 2, 2>("1 half, smem operands, 2 DMUL, unroll 2", 4);
    run<1, 1, 2, 2>("1 half, smem operands, 2 DMUL, unroll 2", 8);
    return 0;
}
