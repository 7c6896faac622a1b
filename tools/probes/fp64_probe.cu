// Microbenchmark: FP64 DFMA dependent-issue latency and throughput per SM sub-partition on B200, SHFL.64 and LDS.64 latency.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_probe fp64_probe.cu ; run on the GPU box.
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP>
__global__ void dfmaKernel(double *out, long long *cyc, int iters, double a, double b) {
    double x[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) x[i] = threadIdx.x * 1e-3 + i;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int i = 0; i < ILP; ++i) x[i] = fma(x[i], a, b);
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// DFMA (ILP 8) interleaved with MIX independent integer / FP32 instructions per DFMA: does a DFMA hold the issue port for the 2
// cycles its pipe needs (then time = 2 nDFMA + nOther), or can other instructions issue in its shadow (time = max)?
template <int MIX>
__global__ void mixKernel(double *out, long long *cyc, int iters, double a, double b, int ia, float fa) {
    double x[8]; int n[8]; float f[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { x[i] = threadIdx.x * 1e-3 + i; n[i] = threadIdx.x + i; f[i] = threadIdx.x * 0.5f + i; }
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                x[i] = fma(x[i], a, b);
#pragma unroll
                for (int m = 0; m < MIX; ++m) { if (m & 1) f[(i + m) & 7] = fmaf(f[(i + m) & 7], fa, 1.0f); else n[(i + m) & 7] = n[(i + m) & 7] * ia + 3; }
            }
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += x[i] + n[i] + f[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int MIX>
void runMix(int threads, double *out, long long *cyc) {
    const int iters = 1000;
    mixKernel<MIX><<<1, threads>>>(out, cyc, iters, 1.0000001, 1e-9, 3, 1.0001f);
    cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("MIX   warps/SM %2d  %d other instr per DFMA : %6.2f cycles per DFMA per warp\n", threads / 32, MIX, double(h) / (iters * 32.0));
}

__global__ void shflKernel(double *out, long long *cyc, int iters) {
    double x = threadIdx.x;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r) x = __shfl_down_sync(0xffffffffu, x, 1) + 1.0;
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = x;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

__global__ void ldsKernel(double *out, long long *cyc, int iters) {
    __shared__ double s[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) s[i] = (i * 7 + 3) % 1024;
    __syncthreads();
    double x = threadIdx.x;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r) x = s[int(x) & 1023];
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = x;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int ILP>
void runDfma(int threads, double *out, long long *cyc) {
    const int iters = 2000;
    dfmaKernel<ILP><<<1, threads>>>(out, cyc, iters, 1.0000001, 1e-9);
    cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    const double per = double(h) / (iters * 8.0 * ILP);
    printf("DFMA  warps/SM %2d (per SMSP %4.2f)  ILP %2d : %7.2f cycles per DFMA per warp  -> %6.2f warp-DFMA/clk/SM\n", threads / 32, threads / 128.0, ILP, per,
           (threads / 32) / per);
}

int main() {
    double *out; long long *cyc;
    cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 1 << 12);
    for (int threads : {32, 128, 256, 384, 512, 1024}) {
        runDfma<1>(threads, out, cyc); runDfma<2>(threads, out, cyc); runDfma<4>(threads, out, cyc); runDfma<8>(threads, out, cyc); runDfma<16>(threads, out, cyc);
    }
    for (int threads : {128, 384}) { runMix<0>(threads, out, cyc); runMix<1>(threads, out, cyc); runMix<2>(threads, out, cyc); runMix<4>(threads, out, cyc); }
    for (int threads : {32, 128, 384}) {
        shflKernel<<<1, threads>>>(out, cyc, 2000); cudaDeviceSynchronize();
        long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        printf("SHFL.64 + DADD dependent chain, %d warps: %.1f cycles per link\n", threads / 32, double(h) / 16000.0);
        ldsKernel<<<1, threads>>>(out, cyc, 2000); cudaDeviceSynchronize();
        cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        printf("LDS.64 (+ cvt) dependent chain, %d warps: %.1f cycles per link\n", threads / 32, double(h) / 16000.0);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
