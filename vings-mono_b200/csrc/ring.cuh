// Shared-memory ring of 32-surfel record batches used by both blend kernels (sm_100a).
//
// The sorted surfels' 128-byte records are gathered with one cp.async.bulk (TMA engine, 1-D copy) per surfel; a
// stage is handed to the consumer warps by an mbarrier armed with the expected byte count (expect_tx / complete_tx)
// and handed back through a shared counter: the LAST warp to finish a batch refills its stage with the batch
// BB_STAGES ahead.  No CTA-wide barrier inside the walk.
#pragma once
#include "common.cuh"

__device__ __forceinline__ uint32_t bb_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bb_mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bb_smem(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void bb_mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bb_smem(bar)) : "memory");
}
__device__ __forceinline__ void bb_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bb_smem(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bb_mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "BB_WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra BB_DONE_%=;\n"
        "bra BB_WAIT_%=;\n"
        "BB_DONE_%=:\n"
        "}\n" ::"r"(bb_smem(bar)), "r"(parity) : "memory");
}
// 1-D bulk copy global -> shared through the TMA engine; completion is signalled on `bar` in bytes
__device__ __forceinline__ void bb_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(bb_smem(dst)), "l"(src), "r"(bytes), "r"(bb_smem(bar)) : "memory");
}

// non-blocking probe of a phase (for waits that must also watch an exit condition)
__device__ __forceinline__ bool bb_mbar_try(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(ok) : "r"(bb_smem(bar)), "r"(parity) : "memory");
    return ok != 0;
}
