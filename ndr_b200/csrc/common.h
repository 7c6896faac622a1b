// ndr_b200 — B200-native VoxelFEM elasticity solve (sm_100a).
// Shared host/device definitions: regular-grid descriptor, error handling, device buffers.
//
// Indexing contract (reference: VoxelFEM/NDVector.hh:279-352, TensorProductSimulator.hh:252-316,
// 998-1012): row-major with the FIRST axis slowest for nodes, elements and local nodes;
// vector fields are (numNodes, N) row-major, i.e. dof = N*node + component.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <vector>

namespace vfem {

typedef double Real;

struct CudaError : std::runtime_error { using std::runtime_error::runtime_error; };

#define VFEM_CUDA(call)                                                                         \
    do {                                                                                        \
        cudaError_t e_ = (call);                                                                \
        if (e_ != cudaSuccess)                                                                  \
            throw ::vfem::CudaError(std::string("CUDA error: ") + cudaGetErrorString(e_) +      \
                                    " at " + __FILE__ + ":" + std::to_string(__LINE__));        \
    } while (0)

#define VFEM_CHECK_LAUNCH() VFEM_CUDA(cudaGetLastError())

// Regular grid of Q1 elements, passed by value to kernels.
template <int N>
struct Grid {
    int E[N];        // elements per axis
    int n[N];        // nodes per axis (E+1)
    int ninc[N];     // flat node index increments (last axis fastest)
    int einc[N];     // flat element index increments
    int nn, ne;      // totals (the largest named grid, 512x256x256, has 33.9 M nodes)
    // Slab decomposition along storage axis 0: all indices above are LOCAL to the slab, which carries ghost
    // node planes below own0 and above own1; [own0, own1] (inclusive) are the node planes this rank owns, i.e. updates
    // in the smoother and counts in reductions.  A single-GPU grid owns everything (own0 = 0, own1 = n[0]-1).
    int own0, own1;
};

template <int N>
inline Grid<N> makeGrid(const int *E) {
    Grid<N> g;
    long long nn = 1, ne = 1;
    for (int d = 0; d < N; ++d) { g.E[d] = E[d]; g.n[d] = E[d] + 1; nn *= g.n[d]; ne *= g.E[d]; }
    if (nn * N >= (1ll << 31)) throw std::runtime_error("grid too large for 32-bit dof indices on one device");
    g.nn = int(nn); g.ne = int(ne);
    g.own0 = 0; g.own1 = g.n[0] - 1;
    g.ninc[N - 1] = 1; g.einc[N - 1] = 1;
    for (int d = N - 2; d >= 0; --d) { g.ninc[d] = g.ninc[d + 1] * g.n[d + 1]; g.einc[d] = g.einc[d + 1] * g.E[d + 1]; }
    return g;
}

// Owning device buffer (cudaMallocAsync-free: plain cudaMalloc, sized once per level).
template <class T>
struct DevBuf {
    T *p = nullptr;
    size_t n = 0;
    DevBuf() {}
    explicit DevBuf(size_t count) { alloc(count); }
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    DevBuf(DevBuf &&o) noexcept : p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
    DevBuf &operator=(DevBuf &&o) noexcept { if (this != &o) { release(); p = o.p; n = o.n; o.p = nullptr; o.n = 0; } return *this; }
    ~DevBuf() { release(); }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
    void alloc(size_t count) {
        if (count == n && p) return;
        release();
        if (count) VFEM_CUDA(cudaMalloc((void **)&p, count * sizeof(T)));
        n = count;
    }
    void zero(cudaStream_t s = 0) { if (n) VFEM_CUDA(cudaMemsetAsync(p, 0, n * sizeof(T), s)); }
    void upload(const T *h, size_t count, cudaStream_t s = 0) {
        alloc(count);
        if (count) VFEM_CUDA(cudaMemcpyAsync(p, h, count * sizeof(T), cudaMemcpyHostToDevice, s));
    }
    void upload(const std::vector<T> &h, cudaStream_t s = 0) { upload(h.data(), h.size(), s); }
    void download(T *h, cudaStream_t s = 0) const {
        if (n) VFEM_CUDA(cudaMemcpyAsync(h, p, n * sizeof(T), cudaMemcpyDeviceToHost, s));
        VFEM_CUDA(cudaStreamSynchronize(s));
    }
    std::vector<T> toHost(cudaStream_t s = 0) const { std::vector<T> h(n); download(h.data(), s); return h; }
    size_t bytes() const { return n * sizeof(T); }
};

// Element stiffness constants handed to kernels as a by-value parameter: kernel parameters
// live in the constant bank, so fully unrolled loops read K0 entries as immediate constant
// operands of DFMA (no register or shared-memory traffic).
template <int N>
struct K0Param {
    static constexpr int NPE = 1 << N, KS = N * NPE;
    Real k[KS * KS];
};

constexpr int ipow3(int n) { return n == 0 ? 1 : 3 * ipow3(n - 1); }

// assembled coarse-level operator, node-major, one matrix row after the other: entry (a,b) (i = a*N+b) of the block
// coupling node n to neighbour slot k sits at ((n*N + a)*3^N + k)*N + b, so row a of node n is 3^N*N contiguous doubles
// that line up with the neighbour values u[n+delta(k)][b]
template <int N>
__host__ __device__ __forceinline__ size_t stencilIndex(int n, int k, int i) { return ((size_t(n) * N + i / N) * ipow3(N) + k) * N + (i % N); }

enum MemSpace { MS_HOST = 0, MS_DEVICE = 1 };

// Count of kernels launched by this library (bench.py's "gpu_launches").
extern unsigned long long g_launchCount;
inline void countLaunch(unsigned long long k = 1) { g_launchCount += k; }

} // namespace vfem
