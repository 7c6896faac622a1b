// grid-wide barrier latency on B200: cooperative_groups grid.sync() against a flag-per-CTA barrier (release store of an epoch
// into the CTA's own 128-byte line, one warp polling all lines with acquire loads).  nvcc -arch=sm_100a -rdc=false -O3
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

__device__ __forceinline__ void stRelease(unsigned *p, unsigned v) { asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned ldAcquire(const unsigned *p) { unsigned v; asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }

__device__ __forceinline__ void flagBarrier(unsigned *flags, unsigned epoch) {
    __syncthreads();
    if (threadIdx.x < 32) {
        if (threadIdx.x == 0) stRelease(flags + blockIdx.x * 32, epoch);
        for (unsigned spin = 0; spin < (1u << 22); ++spin) { // bounded: a broken barrier must not hang the box
            unsigned m = 0xffffffffu;
#pragma unroll 4
            for (unsigned j = threadIdx.x; j < gridDim.x; j += 32) m = min(m, ldAcquire(flags + j * 32));
            if (__all_sync(0xffffffffu, m >= epoch)) break;
        }
    }
    __syncthreads();
}

__global__ void cgKernel(int n, unsigned long long *out, double *sink) {
    cg::grid_group grid = cg::this_grid();
    double acc = threadIdx.x;
    const unsigned long long t0 = clock64();
    for (int i = 0; i < n; ++i) { acc = acc * 1.0000001 + 1.0; grid.sync(); }
    if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = clock64() - t0;
    if (acc == -1.0) sink[0] = acc;
}
__global__ void flagKernel(int n, unsigned *flags, unsigned long long *out, double *sink) {
    double acc = threadIdx.x;
    const unsigned long long t0 = clock64();
    for (int i = 0; i < n; ++i) { acc = acc * 1.0000001 + 1.0; flagBarrier(flags, unsigned(i + 1)); }
    if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = clock64() - t0;
    if (acc == -1.0) sink[0] = acc;
}

int main() {
    int sms = 0, clk = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    unsigned long long *out; double *sink; unsigned *flags;
    cudaMalloc(&out, 8); cudaMalloc(&sink, 8); cudaMalloc(&flags, 4096 * 128);
    const int n = 2000;
    for (int per = 1; per <= 2; ++per) {
        for (int threads : {256, 512}) {
            int blocks = sms * per;
            void *a1[] = {(void *)&n, &out, &sink};
            cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
            float ms1 = 0, ms2 = 0;
            cudaLaunchCooperativeKernel((void *)cgKernel, dim3(blocks), dim3(threads), a1, 0, 0); cudaDeviceSynchronize();
            cudaEventRecord(e0); cudaLaunchCooperativeKernel((void *)cgKernel, dim3(blocks), dim3(threads), a1, 0, 0); cudaEventRecord(e1); cudaDeviceSynchronize();
            cudaEventElapsedTime(&ms1, e0, e1);
            void *a2[] = {(void *)&n, &flags, &out, &sink};
            cudaMemset(flags, 0, 4096 * 128);
            cudaLaunchCooperativeKernel((void *)flagKernel, dim3(blocks), dim3(threads), a2, 0, 0); cudaDeviceSynchronize();
            cudaMemset(flags, 0, 4096 * 128);
            cudaEventRecord(e0); cudaLaunchCooperativeKernel((void *)flagKernel, dim3(blocks), dim3(threads), a2, 0, 0); cudaEventRecord(e1);
            cudaError_t err = cudaDeviceSynchronize();
            cudaEventElapsedTime(&ms2, e0, e1);
            printf("%d CTAs x %d threads: grid.sync %.3f us per barrier, flag barrier %.3f us per barrier (%s)\n", blocks, threads, ms1 * 1e3 / n, ms2 * 1e3 / n, cudaGetErrorString(err));
        }
    }
    return 0;
}
