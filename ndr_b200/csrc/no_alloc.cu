// ndr_b200 — vfem_dot_no_alloc: a dot product whose reduction scratch is a file-scope __device__ array, for callers that
// must not trigger cudaMalloc (code under CUDA-graph capture, serving loops, harnesses that forbid allocation).
// No counterpart in the reference.
//
// The library's own reductions (kernels_vec.cuh: dotKernel + Reducer) keep their per-block partials and the ticket counter
// in cudaMalloc'ed buffers owned by a solver object.  This variant gives every block a FIXED share of kNoAllocShare
// elements, so its scratch need is ceil(n / kNoAllocShare) partials, and keeps them in g_noAllocScratch below:
// kNoAllocMaxPartials doubles + one ticket line = 64.25 KB that every GPU which loads the library holds until exit.
// That bounds n at kNoAllocMaxPartials * kNoAllocShare = 33 554 432 entries (256 MB per vector; the dofs of a
// 256x128x128 grid are 12.8 M); longer vectors are refused.
//
// ALL calls share the one array and it keeps the last call's contents: the ticket is cleared before every launch, the
// partials are overwritten before they are read.  Two calls that overlap on the GPU RACE on it.  Calls are issued on the
// legacy default stream, which orders them one after the other within a process; calls from several processes each have
// their own copy of the array.  Do not issue them from per-thread default streams or move them to concurrent streams.
#include "../../include/vfem_b200.h"
#include "common.h"

namespace vfem {
void setLastError(const std::string &msg); // capi.cu
}

namespace {
using vfem::Real;

constexpr int kNoAllocBlock = 256;                 // threads per block
constexpr int kNoAllocItems = 16;                  // elements per thread
constexpr size_t kNoAllocShare = size_t(kNoAllocBlock) * kNoAllocItems; // 4096 elements per block
constexpr size_t kNoAllocMaxPartials = 8192;       // 64 KB of partials
constexpr size_t kNoAllocMaxN = kNoAllocMaxPartials * kNoAllocShare;
constexpr size_t kNoAllocTicketSlot = kNoAllocMaxPartials; // the ticket lives in the 256-byte line after the partials
constexpr size_t kNoAllocScratchDoubles = kNoAllocMaxPartials + 32;

// In host code this name is a host-side stand-in for the symbol, NOT a device address: the kernel is handed the address
// cudaGetSymbolAddress returns (scratchOfCurrentDevice), once per device.
static __device__ __align__(256) Real g_noAllocScratch[kNoAllocScratchDoubles];

// out[0] = a.b — block i reduces elements [i * 4096, (i + 1) * 4096) in a fixed order, the last block to finish adds the
// partials in index order (deterministic for a given n).
__global__ void __launch_bounds__(kNoAllocBlock) dotFixedShareKernel(const Real *__restrict__ a, const Real *__restrict__ b, size_t n,
                                                                    Real *scratch, Real *__restrict__ out) {
    __shared__ Real sh[kNoAllocBlock / 32];
    __shared__ bool last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t base = size_t(blockIdx.x) * kNoAllocShare;
    Real v = 0.0;
#pragma unroll
    for (int k = 0; k < kNoAllocItems; ++k) {
        const size_t i = base + size_t(k) * kNoAllocBlock + threadIdx.x;
        if (i < n) v = fma(a[i], b[i], v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) sh[warp] = v;
    __syncthreads();
    unsigned int *ticket = reinterpret_cast<unsigned int *>(scratch + kNoAllocTicketSlot);
    if (threadIdx.x == 0) {
        Real s = 0.0;
        for (int w = 0; w < kNoAllocBlock / 32; ++w) s += sh[w];
        scratch[blockIdx.x] = s;
        __threadfence();
        last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (last && threadIdx.x < 32) {
        __threadfence();
        Real s = 0.0;
        for (unsigned i = lane; i < gridDim.x; i += 32) s += __ldcg(scratch + i);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) { out[0] = s; *ticket = 0u; }
    }
}

constexpr int kMaxDevices = 64;
Real *g_scratchAddr[kMaxDevices] = {};

// device address of g_noAllocScratch on the CURRENT device, looked up once per device (no allocation: the array is part
// of the module image)
Real *scratchOfCurrentDevice() {
    int dev = 0;
    VFEM_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) throw std::runtime_error("vfem_dot_no_alloc: device index " + std::to_string(dev) + " out of range");
    if (!g_scratchAddr[dev]) {
        Real *p = nullptr;
        VFEM_CUDA(cudaGetSymbolAddress((void **)&p, g_noAllocScratch));
        if (!p) throw std::runtime_error("vfem_dot_no_alloc: cudaGetSymbolAddress returned a null address");
        g_scratchAddr[dev] = p;
    }
    return g_scratchAddr[dev];
}
} // namespace

extern "C" {

size_t vfem_dot_no_alloc_max_n(void) { return kNoAllocMaxN; }

int vfem_dot_no_alloc(const double *a_device, const double *b_device, size_t n, double *result_device) {
    try {
        if (n > kNoAllocMaxN)
            throw std::runtime_error("vfem_dot_no_alloc: n = " + std::to_string(n) + " exceeds the static scratch (at most " +
                                     std::to_string(kNoAllocMaxN) + " entries = " + std::to_string(kNoAllocMaxPartials) + " partials of " +
                                     std::to_string(kNoAllocShare) + " elements); use vfem_mg_dot");
        if (!result_device || (n > 0 && (!a_device || !b_device))) throw std::runtime_error("vfem_dot_no_alloc: null array");
        if (n == 0) { VFEM_CUDA(cudaMemsetAsync(result_device, 0, sizeof(double), 0)); return 0; }
        Real *scratch = scratchOfCurrentDevice();
        // the array keeps whatever the last call (or an aborted one) left: the ticket must start at zero
        VFEM_CUDA(cudaMemsetAsync(scratch + kNoAllocTicketSlot, 0, 32 * sizeof(Real), 0));
        const unsigned blocks = unsigned((n + kNoAllocShare - 1) / kNoAllocShare);
        dotFixedShareKernel<<<blocks, kNoAllocBlock, 0, 0>>>(a_device, b_device, n, scratch, result_device);
        VFEM_CHECK_LAUNCH(); vfem::countLaunch();
    } catch (const std::exception &e) { vfem::setLastError(e.what()); return 1; } catch (...) { vfem::setLastError("unknown error"); return 1; }
    return 0;
}

} // extern "C"
