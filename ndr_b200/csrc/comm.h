// ndr_b200 — communication layer of the slab-decomposed solve.
//
// The grid is cut into slabs along storage axis 0; rank r talks to r-1 ("below") and r+1 ("above") only, plus
// scalar / coarsest-level all-reduces.  Two transports behind one interface:
//   * NcclComm   — one process per GPU; ncclSend/ncclRecv/ncclAllReduce enqueued on the solver's stream (NVLink 5 /
//                  NVSwitch).  libnccl is resolved with dlopen so that the copy torch already loaded is shared and the
//                  library keeps loading on boxes without NCCL.
//   * ThreadComm — W ranks as W host threads of ONE process sharing ONE device and ONE stream: an exchange is a host
//                  barrier (every rank has enqueued its producer kernels) + device-to-device copies + a second barrier.
//                  It exists so that the slab algorithm is parity-tested on a single GPU.
#pragma once
#include "common.h"
#include <condition_variable>
#include <memory>
#include <mutex>

namespace vfem {

struct SlabComm {
    int rank = 0, world = 1;
    virtual ~SlabComm() {}
    // up:   send `sendUp`   to rank+1, receive `recvBelow` from rank-1
    // down: send `sendDown` to rank-1, receive `recvAbove` from rank+1          (count doubles each, device pointers)
    virtual void shift(const Real *sendUp, Real *recvBelow, const Real *sendDown, Real *recvAbove, size_t count, bool up, bool down,
                       cudaStream_t s) = 0;
    virtual void allreduceSum(Real *buf, size_t n, cudaStream_t s) = 0;
    virtual void allreduceSumBytes(uint8_t *buf, size_t n, cudaStream_t s) = 0;
    virtual bool graphSafe() const = 0; // may the calls above be captured into a CUDA graph?
    virtual bool sideStreamSafe() const = 0; // may a shift be enqueued on another stream than the solver's (overlap with compute)?
};

std::shared_ptr<SlabComm> makeNcclComm(const uint8_t *uniqueId128, int rank, int world);
void newNcclUniqueId(uint8_t *out128);
std::vector<std::shared_ptr<SlabComm>> makeThreadComms(int world);

} // namespace vfem
