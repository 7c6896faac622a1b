// ndr_b200 — transports of the slab solve (see comm.h).
#include "comm.h"
#include <chrono>
#include <cstring>
#include <dlfcn.h>
#include <nccl.h> // types and enums only: the entry points are resolved with dlsym

namespace vfem {

// ---- NCCL ----------------------------------------------------------------------------------------------
namespace {
struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};

NcclApi &nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        // RTLD_NOLOAD first: share the copy torch.distributed has already mapped into this process
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char *n : names) { api.lib = dlopen(n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL); if (api.lib) break; }
        if (!api.lib) for (const char *n : names) { api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (api.lib) break; }
        if (!api.lib) return;
#define VFEM_NCCL_SYM(name) api.name = reinterpret_cast<decltype(api.name)>(dlsym(api.lib, "nccl" #name))
        VFEM_NCCL_SYM(GetUniqueId); VFEM_NCCL_SYM(CommInitRank); VFEM_NCCL_SYM(CommDestroy); VFEM_NCCL_SYM(Send); VFEM_NCCL_SYM(Recv);
        VFEM_NCCL_SYM(AllReduce); VFEM_NCCL_SYM(AllGather); VFEM_NCCL_SYM(GroupStart); VFEM_NCCL_SYM(GroupEnd); VFEM_NCCL_SYM(GetErrorString);
#undef VFEM_NCCL_SYM
    });
    if (!api.lib || !api.Send || !api.CommInitRank)
        throw std::runtime_error("NCCL is not available (libnccl.so.2 could not be loaded): the multi-GPU slab solve needs it");
    return api;
}

#define VFEM_NCCL(call)                                                                                                  \
    do {                                                                                                                 \
        ncclResult_t r_ = (call);                                                                                        \
        if (r_ != ncclSuccess) throw std::runtime_error(std::string("NCCL error: ") + nccl().GetErrorString(r_) + " at " + __FILE__ + ":" + std::to_string(__LINE__)); \
    } while (0)

// (Round 1's peer-memory mailbox shift - data + flag + acknowledgement over CUDA IPC - measured slower than ncclSend/ncclRecv
// for these 100 KB planes, profiles/r01m_peer_halo.txt, and is gone; the numbers stay in the profile.)
struct NcclComm : SlabComm {
    ncclComm_t comm = nullptr;
    bool graphs = false;
    ~NcclComm() override { if (comm) nccl().CommDestroy(comm); }
    void shift(const Real *sendUp, Real *recvBelow, const Real *sendDown, Real *recvAbove, size_t count, bool up, bool down,
               cudaStream_t s) override {
        NcclApi &a = nccl();
        VFEM_NCCL(a.GroupStart());
        if (up) {
            if (rank + 1 < world) VFEM_NCCL(a.Send(sendUp, count, ncclDouble, rank + 1, comm, s));
            if (rank > 0) VFEM_NCCL(a.Recv(recvBelow, count, ncclDouble, rank - 1, comm, s));
        }
        if (down) {
            if (rank > 0) VFEM_NCCL(a.Send(sendDown, count, ncclDouble, rank - 1, comm, s));
            if (rank + 1 < world) VFEM_NCCL(a.Recv(recvAbove, count, ncclDouble, rank + 1, comm, s));
        }
        VFEM_NCCL(a.GroupEnd());
    }
    void allreduceSum(Real *buf, size_t n, cudaStream_t s) override { VFEM_NCCL(nccl().AllReduce(buf, buf, n, ncclDouble, ncclSum, comm, s)); }
    void allreduceSumBytes(uint8_t *buf, size_t n, cudaStream_t s) override { VFEM_NCCL(nccl().AllReduce(buf, buf, n, ncclUint8, ncclSum, comm, s)); }
    bool graphSafe() const override { return graphs; }
    bool sideStreamSafe() const override { return true; }
};
} // namespace

void newNcclUniqueId(uint8_t *out128) {
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    VFEM_NCCL(nccl().GetUniqueId(&id));
    memcpy(out128, &id, 128);
}

std::shared_ptr<SlabComm> makeNcclComm(const uint8_t *uniqueId128, int rank, int world) {
    auto c = std::make_shared<NcclComm>();
    c->rank = rank; c->world = world;
    c->graphs = getenv("VFEM_SLAB_NO_GRAPH") == nullptr; // ncclSend / ncclRecv / ncclAllReduce are graph-capturable (NCCL >= 2.9): the slab preconditioner replays as one graph
    ncclUniqueId id;
    memcpy(&id, uniqueId128, 128);
    VFEM_NCCL(nccl().CommInitRank(&c->comm, world, id, rank));
    return c;
}

// ---- threads sharing one device and one stream ---------------------------------------------------------------
namespace {
constexpr int kMaxThreadRanks = 16;
struct PtrTable { const void *p[kMaxThreadRanks]; };

template <class T>
__global__ void sumRanksKernel(PtrTable t, int world, size_t n, T *out) {
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) {
        T s = static_cast<const T *>(t.p[0])[i];
        for (int r = 1; r < world; ++r) s += static_cast<const T *>(t.p[r])[i]; // fixed rank order: deterministic
        out[i] = s;
    }
}

struct ThreadShared {
    std::mutex m;
    std::condition_variable cv;
    int world = 1, arrived = 0;
    unsigned long long generation = 0;
    const void *a[kMaxThreadRanks] = {}, *b[kMaxThreadRanks] = {};
    void barrier() {
        std::unique_lock<std::mutex> lk(m);
        const unsigned long long gen = generation;
        if (++arrived == world) { arrived = 0; ++generation; cv.notify_all(); return; }
        if (!cv.wait_for(lk, std::chrono::seconds(30), [&] { return generation != gen; }))
            throw std::runtime_error("ThreadComm: a rank did not reach the collective within 30 s (ranks must make the same calls)");
    }
};

struct ThreadComm : SlabComm {
    std::shared_ptr<ThreadShared> sh;
    DevBuf<uint8_t> scratch;
    void shift(const Real *sendUp, Real *recvBelow, const Real *sendDown, Real *recvAbove, size_t count, bool up, bool down,
               cudaStream_t s) override {
        sh->a[rank] = sendUp; sh->b[rank] = sendDown;
        sh->barrier(); // every rank has enqueued the kernels that produce its planes (one shared in-order stream)
        if (up && rank > 0) VFEM_CUDA(cudaMemcpyAsync(recvBelow, sh->a[rank - 1], count * sizeof(Real), cudaMemcpyDeviceToDevice, s));
        if (down && rank + 1 < world) VFEM_CUDA(cudaMemcpyAsync(recvAbove, sh->b[rank + 1], count * sizeof(Real), cudaMemcpyDeviceToDevice, s));
        sh->barrier(); // ... and the copies are enqueued before anybody overwrites a source plane
    }
    template <class T>
    void allreduce(T *buf, size_t n, cudaStream_t s) {
        if (scratch.n < n * sizeof(T)) scratch.alloc(n * sizeof(T));
        sh->a[rank] = buf;
        sh->barrier();
        PtrTable t;
        for (int r = 0; r < world; ++r) t.p[r] = sh->a[r];
        const unsigned blocks = unsigned(std::min<size_t>((n + 255) / 256, 1184));
        sumRanksKernel<T><<<blocks, 256, 0, s>>>(t, world, n, reinterpret_cast<T *>(scratch.p));
        VFEM_CHECK_LAUNCH();
        sh->barrier(); // all sums are enqueued before any input is overwritten
        VFEM_CUDA(cudaMemcpyAsync(buf, scratch.p, n * sizeof(T), cudaMemcpyDeviceToDevice, s));
        sh->barrier();
    }
    void allreduceSum(Real *buf, size_t n, cudaStream_t s) override { allreduce<Real>(buf, n, s); }
    void allreduceSumBytes(uint8_t *buf, size_t n, cudaStream_t s) override { allreduce<uint8_t>(buf, n, s); }
    bool graphSafe() const override { return false; } // replaying per-rank graphs would serialise whole cycles
    bool sideStreamSafe() const override { return false; } // the host barriers order work on the ONE shared stream only
};
} // namespace

std::vector<std::shared_ptr<SlabComm>> makeThreadComms(int world) {
    if (world < 1 || world > kMaxThreadRanks) throw std::runtime_error("ThreadComm supports 1.." + std::to_string(kMaxThreadRanks) + " ranks");
    auto sh = std::make_shared<ThreadShared>();
    sh->world = world;
    std::vector<std::shared_ptr<SlabComm>> out;
    for (int r = 0; r < world; ++r) {
        auto c = std::make_shared<ThreadComm>();
        c->rank = r; c->world = world; c->sh = sh;
        out.push_back(c);
    }
    return out;
}

} // namespace vfem
