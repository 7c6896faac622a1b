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

// ---- halo shift through NVLink peer memory ---------------------------------------------------------------------------
// OPT-IN (VFEM_PEER_HALO=1): measured on 2 B200 it is SLOWER than NCCL's LL protocol for these 100 KB planes (about 20 us
// against 12 us per shift inside a GS sweep, profiles/r01m_peer_halo.txt), because the data + flag + acknowledgement chain
// pays three system-scope fences over NVLink; it stays as the base for fusing the push into the producing kernel.
// Every rank owns a MAILBOX in its HBM, mapped into its two neighbours with CUDA IPC: the sender's blocks store the plane straight into the neighbour's mailbox over NVLink, fence,
// and the last one raises the neighbour's sequence flag; the receiver's blocks (same launch on that rank) wait for the
// flag, move the plane from the mailbox to the ghost plane and acknowledge.  Two slots per direction, so a sender only
// ever waits for the acknowledgement of the message before the previous one.  Flags sit on their own 128-byte lines and
// have exactly one writer each.
enum PeerFlag { SEQ_FROM_BELOW, SEQ_FROM_ABOVE, ACK_FROM_BELOW, ACK_FROM_ABOVE, DONE_SEND_UP, DONE_RECV_UP, DONE_SEND_DOWN, DONE_RECV_DOWN, PEER_FLAGS };
constexpr size_t kFlagStride = 128;
constexpr int kPeerThreads = 256, kPeerMaxBlocks = 8;

struct PeerShiftArgs {
    const Real *sendUp, *sendDown;
    Real *recvBelow, *recvAbove;
    size_t count, cap;
    unsigned char *self, *above, *below;   // mailboxes; above / below are IPC mappings (nullptr at the ends of the chain)
    unsigned long long seqUp, seqDown;     // message numbers of this call, 0 = direction not taken
    unsigned long long target[4];          // cumulative block count of each role after this call
    int nblk;
    int *err;
};

__device__ inline unsigned long long *peerFlag(unsigned char *box, int f) { return reinterpret_cast<unsigned long long *>(box + size_t(f) * kFlagStride); }
__device__ inline Real *peerData(unsigned char *box, int fromAbove, unsigned long long seq, size_t cap) {
    return reinterpret_cast<Real *>(box + size_t(PEER_FLAGS) * kFlagStride) + (size_t(fromAbove) * 2 + (seq & 1ull)) * cap;
}
__device__ inline bool peerWait(unsigned long long *flag, unsigned long long want, int *err) {
    unsigned long long t0, t, v;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flag) : "memory");
        if (v >= want) return true;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t - t0 > 20000000000ull) { *err = 1; return false; } // 20 s: a neighbour never came - report instead of hanging the GPU
    }
}
__device__ inline void peerSignal(unsigned long long *flag, unsigned long long v) {
    __threadfence_system();
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag), "l"(v) : "memory");
}

__global__ void __launch_bounds__(kPeerThreads) peerShiftKernel(PeerShiftArgs a) {
    const int role = blockIdx.x / a.nblk, b = blockIdx.x % a.nblk; // 0 send up, 1 receive from below, 2 send down, 3 receive from above
    const bool upward = role < 2, sender = (role & 1) == 0;
    const unsigned long long seq = upward ? a.seqUp : a.seqDown;
    unsigned char *peer = sender ? (upward ? a.above : a.below) : (upward ? a.below : a.above);
    if (!seq || !peer) return;
    __shared__ int ok;
    if (sender) {
        const Real *src = upward ? a.sendUp : a.sendDown;
        if (threadIdx.x == 0) ok = peerWait(peerFlag(a.self, upward ? ACK_FROM_ABOVE : ACK_FROM_BELOW), seq >= 2 ? seq - 2 : 0, a.err);
        __syncthreads();
        if (ok) {
            Real *dst = peerData(peer, upward ? 0 : 1, seq, a.cap);
            for (size_t i = size_t(b) * kPeerThreads + threadIdx.x; i < a.count; i += size_t(a.nblk) * kPeerThreads) dst[i] = src[i];
        }
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0 && atomicAdd(peerFlag(a.self, upward ? DONE_SEND_UP : DONE_SEND_DOWN), 1ull) + 1 == a.target[role])
            peerSignal(peerFlag(peer, upward ? SEQ_FROM_BELOW : SEQ_FROM_ABOVE), seq);
    } else {
        Real *dst = upward ? a.recvBelow : a.recvAbove;
        if (threadIdx.x == 0) ok = peerWait(peerFlag(a.self, upward ? SEQ_FROM_BELOW : SEQ_FROM_ABOVE), seq, a.err);
        __syncthreads();
        if (ok) {
            const Real *src = peerData(a.self, upward ? 0 : 1, seq, a.cap);
            for (size_t i = size_t(b) * kPeerThreads + threadIdx.x; i < a.count; i += size_t(a.nblk) * kPeerThreads) dst[i] = __ldcg(src + i);
        }
        __syncthreads();
        if (threadIdx.x == 0 && atomicAdd(peerFlag(a.self, upward ? DONE_RECV_UP : DONE_RECV_DOWN), 1ull) + 1 == a.target[role])
            peerSignal(peerFlag(peer, upward ? ACK_FROM_ABOVE : ACK_FROM_BELOW), seq);
    }
}

struct NcclComm : SlabComm {
    ncclComm_t comm = nullptr;
    bool graphs = false;
    // peer-memory shift state
    unsigned char *box = nullptr, *boxAbove = nullptr, *boxBelow = nullptr;
    size_t cap = 0;
    unsigned long long seqUp = 0, seqDown = 0, blocksDone[4] = {0, 0, 0, 0};
    int *errHost = nullptr, *errDev = nullptr;
    bool peer = false;
    ~NcclComm() override {
        if (boxAbove) cudaIpcCloseMemHandle(boxAbove);
        if (boxBelow) cudaIpcCloseMemHandle(boxBelow);
        if (box) cudaFree(box);
        if (errHost) cudaFreeHost(errHost);
        if (comm) nccl().CommDestroy(comm);
    }
    void preparePeer(size_t maxCount, cudaStream_t s) override {
        if (world < 2 || peer || !getenv("VFEM_PEER_HALO")) return;
        NcclApi &a = nccl();
        cap = (maxCount + 15) / 16 * 16;
        const size_t bytes = size_t(PEER_FLAGS) * kFlagStride + 4 * cap * sizeof(Real);
        int bad = 0;
        cudaIpcMemHandle_t mine;
        memset(&mine, 0, sizeof mine);
        if (cudaMalloc((void **)&box, bytes) != cudaSuccess) { box = nullptr; bad = 1; }
        if (!bad && (cudaMemsetAsync(box, 0, bytes, s) != cudaSuccess || cudaIpcGetMemHandle(&mine, box) != cudaSuccess)) bad = 1;
        if (cudaHostAlloc((void **)&errHost, sizeof(int), cudaHostAllocMapped) != cudaSuccess || cudaHostGetDevicePointer((void **)&errDev, errHost, 0) != cudaSuccess) bad = 1;
        else *errHost = 0;
        (void)cudaGetLastError();
        static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handles are 64 bytes");
        DevBuf<uint8_t> all(size_t(world) * 64);
        VFEM_CUDA(cudaMemcpyAsync(all.p + size_t(rank) * 64, &mine, 64, cudaMemcpyHostToDevice, s));
        VFEM_NCCL(a.AllGather(all.p + size_t(rank) * 64, all.p, 64, ncclUint8, comm, s)); // in place; orders every rank's memset before any neighbour's first store
        std::vector<uint8_t> handles = all.toHost(s);
        auto open = [&](int r, unsigned char **out) {
            cudaIpcMemHandle_t h;
            memcpy(&h, handles.data() + size_t(r) * 64, 64);
            if (cudaIpcOpenMemHandle((void **)out, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { *out = nullptr; bad = 1; (void)cudaGetLastError(); }
        };
        if (!bad && rank + 1 < world) open(rank + 1, &boxAbove);
        if (!bad && rank > 0) open(rank - 1, &boxBelow);
        DevBuf<Real> flag(1);
        const Real mineBad = bad;
        VFEM_CUDA(cudaMemcpyAsync(flag.p, &mineBad, sizeof(Real), cudaMemcpyHostToDevice, s));
        allreduceSum(flag.p, 1, s); // either every rank shifts through peer memory or none does
        Real anyBad = 0;
        flag.download(&anyBad, s);
        peer = anyBad == 0;
        if (!peer && rank == 0) fprintf(stderr, "ndr_b200: peer-memory halo shift unavailable (CUDA IPC failed on %d rank(s)); using ncclSend/ncclRecv\n", int(anyBad));
    }
    void peerShift(const Real *sendUp, Real *recvBelow, const Real *sendDown, Real *recvAbove, size_t count, bool up, bool down, cudaStream_t s) {
        if (*errHost) throw std::runtime_error("peer-memory halo shift: a neighbouring rank did not answer within 20 s");
        PeerShiftArgs a;
        a.sendUp = sendUp; a.sendDown = sendDown; a.recvBelow = recvBelow; a.recvAbove = recvAbove;
        a.count = count; a.cap = cap; a.self = box; a.above = boxAbove; a.below = boxBelow;
        a.seqUp = up ? ++seqUp : 0; a.seqDown = down ? ++seqDown : 0;
        a.nblk = int(std::min<size_t>(kPeerMaxBlocks, std::max<size_t>(1, count / 2048)));
        const bool active[4] = {up && boxAbove, up && boxBelow, down && boxBelow, down && boxAbove};
        for (int r = 0; r < 4; ++r) { if (active[r]) blocksDone[r] += a.nblk; a.target[r] = blocksDone[r]; }
        a.err = errDev;
        peerShiftKernel<<<4 * a.nblk, kPeerThreads, 0, s>>>(a);
        VFEM_CHECK_LAUNCH();
    }
    void shift(const Real *sendUp, Real *recvBelow, const Real *sendDown, Real *recvAbove, size_t count, bool up, bool down,
               cudaStream_t s) override {
        if (peer && count <= cap) { peerShift(sendUp, recvBelow, sendDown, recvAbove, count, up, down, s); return; }
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
    bool graphSafe() const override { return graphs && !peer; } // the peer shift numbers its messages on the host
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
