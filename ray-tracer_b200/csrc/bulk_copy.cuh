// Bulk global -> shared copies on the Blackwell copy engine (cp.async.bulk, the 1-D form of TMA: SASS UBLKCP) completed through an
// mbarrier.  The wavefront kernels stage their read-only scene tables (BVH pairs, primitive heads, frame-rect list, materials) once
// per block; issuing that as ONE asynchronous bulk copy lets the block's threads request their first rays from HBM while the
// tables stream in, instead of running a scattered LDG + STS loop and a barrier first.
//
// Protocol (one phase, parity 0):   thread 0: mbarrier.init(1) ; fence ; __syncthreads ;
//                                   thread 0: arrive.expect_tx(total bytes) ; cp.async.bulk ... ;   all: (other work) ; try_wait.parity(0)
// Sources and sizes must be 16-byte aligned / multiples of 16 (every table here is float4-grained).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rtb {

__device__ __forceinline__ void mbar_init(uint32_t mbar, uint32_t arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(arrivals) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");  // make the initialised barrier visible to the async proxy
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t mbar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst_smem, const void* src_gmem, uint32_t bytes, uint32_t mbar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_smem), "l"(src_gmem), "r"(bytes),
                 "r"(mbar)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t mbar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(mbar),
        "r"(parity)
        : "memory");
}

// Up to three tables into consecutive shared-memory regions, one barrier.  Called by every thread; returns after the copies are
// ISSUED (thread 0) — the caller does independent work, then mbar_wait(mbar, 0).  The __syncthreads orders the barrier's
// initialisation before any thread's wait.
struct BulkPart {
    const void* src;
    uint32_t dst, bytes;
};
__device__ __forceinline__ void bulk_stage(uint32_t mbar, const BulkPart p0, const BulkPart p1 = BulkPart{nullptr, 0u, 0u}, const BulkPart p2 = BulkPart{nullptr, 0u, 0u}) {
    if (threadIdx.x == 0) mbar_init(mbar, 1u);
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_expect_tx(mbar, p0.bytes + p1.bytes + p2.bytes);
        if (p0.bytes) bulk_g2s(p0.dst, p0.src, p0.bytes, mbar);
        if (p1.bytes) bulk_g2s(p1.dst, p1.src, p1.bytes, mbar);
        if (p2.bytes) bulk_g2s(p2.dst, p2.src, p2.bytes, mbar);
    }
}

}  // namespace rtb
