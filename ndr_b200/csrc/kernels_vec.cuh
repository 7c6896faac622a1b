// ndr_b200 — PCG vector kernels with fused deterministic reductions (MG.hh:679-732), Dirichlet
// mask passes (MG.hh:364-398) and small helpers.
//
// Reductions: every block reduces its grid-stride share with warp shuffles in a fixed order
// and writes one partial; the last block to finish (ticket counter) adds the partials in index
// order.  The launch shape is fixed per vector length, so results are run-to-run reproducible.
#pragma once
#include "common.h"

namespace vfem {

constexpr int kRedBlock = 256;
constexpr int kRedMaxBlocks = 1184; // 148 SMs x 8 resident blocks

struct ReduceWork {
    Real *partials = nullptr;      // kRedMaxBlocks * 4
    unsigned int *ticket = nullptr; // zero-initialised, reset by the finishing block
};

inline int reduceBlocks(size_t n) {
    size_t b = (n + kRedBlock * 4 - 1) / (kRedBlock * 4);
    if (b < 1) b = 1;
    if (b > kRedMaxBlocks) b = kRedMaxBlocks;
    return int(b);
}

template <int K>
__device__ __forceinline__ void blockReduceAndFinish(Real (&v)[K], Real *partials, unsigned int *ticket, Real *out) {
    __shared__ Real sh[K][kRedBlock / 32];
    __shared__ bool last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < K; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
        if (lane == 0) sh[k][warp] = v[k];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < K; ++k) {
            Real s = 0.0;
            for (int w = 0; w < kRedBlock / 32; ++w) s += sh[k][w];
            partials[size_t(k) * kRedMaxBlocks + blockIdx.x] = s;
        }
        __threadfence();
        unsigned int t = atomicAdd(ticket, 1u);
        last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (last && threadIdx.x < 32) {
        __threadfence();
#pragma unroll
        for (int k = 0; k < K; ++k) {
            Real s = 0.0;
            for (int i = lane; i < (int)gridDim.x; i += 32) s += partials[size_t(k) * kRedMaxBlocks + i];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == 0) out[k] = s;
        }
        if (lane == 0) *ticket = 0u;
    }
}

// out[0] = a.b
static __global__ void __launch_bounds__(kRedBlock) dotKernel(const Real *__restrict__ a, const Real *__restrict__ b, size_t n, ReduceWork w, Real *out) {
    Real v[1] = {0.0};
    for (size_t i = size_t(blockIdx.x) * kRedBlock + threadIdx.x; i < n; i += size_t(gridDim.x) * kRedBlock) v[0] = fma(a[i], b[i], v[0]);
    blockReduceAndFinish<1>(v, w.partials, w.ticket, out);
}

// y += alpha x
static __global__ void __launch_bounds__(kRedBlock) axpyKernel(Real *__restrict__ y, const Real *__restrict__ x, Real alpha, size_t n) {
    for (size_t i = size_t(blockIdx.x) * kRedBlock + threadIdx.x; i < n; i += size_t(gridDim.x) * kRedBlock) y[i] = fma(alpha, x[i], y[i]);
}

// reduction scratch of one owner (scalars on the device + a pinned host mirror)
struct Reducer {
    DevBuf<Real> partials, scal;
    DevBuf<unsigned int> ticket;
    Real *h = nullptr;
    ReduceWork rw;
    void init(cudaStream_t s) {
        if (h) return;
        partials.alloc(size_t(kRedMaxBlocks) * 4); scal.alloc(8); ticket.alloc(1);
        scal.zero(s); ticket.zero(s);
        rw.partials = partials.p; rw.ticket = ticket.p;
        VFEM_CUDA(cudaMallocHost((void **)&h, 8 * sizeof(Real)));
    }
    Real fetch(int slot, cudaStream_t s) { // synchronous read-back of scal[slot]
        VFEM_CUDA(cudaMemcpyAsync(h + slot, scal.p + slot, sizeof(Real), cudaMemcpyDeviceToHost, s));
        VFEM_CUDA(cudaStreamSynchronize(s));
        return h[slot];
    }
    ~Reducer() { if (h) cudaFreeHost(h); }
};

// out[0] = sum(a)
static __global__ void __launch_bounds__(kRedBlock) sumKernel(const Real *__restrict__ a, size_t n, ReduceWork w, Real *out) {
    Real v[1] = {0.0};
    for (size_t i = size_t(blockIdx.x) * kRedBlock + threadIdx.x; i < n; i += size_t(gridDim.x) * kRedBlock) v[0] += a[i];
    blockReduceAndFinish<1>(v, w.partials, w.ticket, out);
}

// s = mask(s); out[0] = r.s          (MG.hh:713-716; dmask is per node, N components per node)
template <int N>
static __global__ void __launch_bounds__(kRedBlock) maskDotKernel(Real *__restrict__ s, const Real *__restrict__ r, const uint8_t *__restrict__ dmask,
                                                            size_t n, ReduceWork w, Real *out) {
    Real v[1] = {0.0};
    for (size_t i = size_t(blockIdx.x) * kRedBlock + threadIdx.x; i < n; i += size_t(gridDim.x) * kRedBlock) {
        Real si = s[i];
        if ((dmask[i / N] >> (i % N)) & 1u) { si = 0.0; s[i] = 0.0; }
        v[0] = fma(r[i], si, v[0]);
    }
    blockReduceAndFinish<1>(v, w.partials, w.ticket, out);
}

// d = s + (rMr / rMrOld) d   (first iteration: d = s), MG.hh:717-718
static __global__ void __launch_bounds__(kRedBlock) updateDirectionKernel(Real *__restrict__ d, const Real *__restrict__ s, size_t n,
                                                                    const Real *rMr, const Real *rMrOld, int first) {
    const Real beta = first ? 0.0 : (*rMr) / (*rMrOld);
    for (size_t i = size_t(blockIdx.x) * kRedBlock + threadIdx.x; i < n; i += size_t(gridDim.x) * kRedBlock)
        d[i] = first ? s[i] : s[i] + beta * d[i];
}

// alpha = rMr / dAd;  x += alpha d;  r -= alpha Ad;  out[0] = r.r    (MG.hh:721-724)
static __global__ void __launch_bounds__(kRedBlock) updateSolutionKernel(Real *__restrict__ x, Real *__restrict__ r, const Real *__restrict__ d,
                                                                   const Real *__restrict__ Ad, size_t n, const Real *rMr, const Real *dAd,
                                                                   ReduceWork w, Real *out) {
    const Real alpha = (*rMr) / (*dAd);
    Real v[1] = {0.0};
    for (size_t i = size_t(blockIdx.x) * kRedBlock + threadIdx.x; i < n; i += size_t(gridDim.x) * kRedBlock) {
        x[i] += alpha * d[i];
        Real ri = r[i] - alpha * Ad[i];
        r[i] = ri;
        v[0] = fma(ri, ri, v[0]);
    }
    blockReduceAndFinish<1>(v, w.partials, w.ticket, out);
}

// u[dof] = fixed ? (zero ? 0 : value) : u[dof]    (MG.hh:364-398)
template <int N>
static __global__ void __launch_bounds__(kRedBlock) enforceDirichletKernel(Real *__restrict__ u, const uint8_t *__restrict__ dmask,
                                                                     const Real *__restrict__ dval, size_t n, int zero) {
    for (size_t i = size_t(blockIdx.x) * kRedBlock + threadIdx.x; i < n; i += size_t(gridDim.x) * kRedBlock)
        if ((dmask[i / N] >> (i % N)) & 1u) u[i] = zero ? 0.0 : dval[i];
}

static __global__ void copyScalarKernel(Real *dst, const Real *src) { *dst = *src; }

// SIMP modulus per element (TPS.hh:725-727), evaluated once per density update instead of inside
// every apply/smooth as the reference does.
static __global__ void __launch_bounds__(kRedBlock) simpKernel(const Real *__restrict__ rho, Real *__restrict__ Ey, size_t ne, Real E0, Real Emin, Real gamma) {
    for (size_t i = size_t(blockIdx.x) * kRedBlock + threadIdx.x; i < ne; i += size_t(gridDim.x) * kRedBlock)
        Ey[i] = Emin + pow(rho[i], gamma) * (E0 - Emin);
}

} // namespace vfem
