// ndr_b200 — mbarrier + bulk asynchronous copy primitives (sm_90+ PTX; on sm_100a cp.async.bulk compiles to UBLKCP, the
// TMA unit's non-tensor path).  Tensor-map TMA (cp.async.bulk.tensor) does not fit the fields of this library: a row of an
// (nn, 3) FP64 field starts at 24 * node bytes and the row pitch 24 * nz is not a multiple of 16 for odd nz (nz = E + 1), so
// the global strides a CUtensorMap demands do not exist.  Row-wise bulk copies with a per-row 8-byte parity shift do.
#pragma once
#include "common.h"

namespace vfem {

__device__ __forceinline__ unsigned smemAddr(const void *p) { return unsigned(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbarInit(unsigned long long *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smemAddr(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbarArriveExpectTx(unsigned long long *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbarWait(unsigned long long *bar, unsigned phase) {
    unsigned done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(smemAddr(bar)), "r"(phase) : "memory");
}
// global -> shared bulk copy (16-byte aligned addresses, size a multiple of 16), completion bytes counted on `bar`
__device__ __forceinline__ void bulkLoad(void *dst, const void *src, unsigned bytes, unsigned long long *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smemAddr(dst)), "l"(src), "r"(bytes), "r"(smemAddr(bar)) : "memory");
}

__device__ __forceinline__ void mbarArrive(unsigned long long *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smemAddr(bar)) : "memory");
}

} // namespace vfem
