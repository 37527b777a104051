// Shared-memory ring of 32-surfel record batches used by both blend kernels (sm_100a).
//
// A stage is handed to the consumer warps by an mbarrier and handed back through a shared counter: the LAST warp to
// finish a batch refills its stage with the batch BB_STAGES ahead.  No CTA-wide barrier inside the walk.
//
// Two gathers of the sorted surfels' 128-byte records, same ring, same mbarriers (bb_gather below):
//   * SB_RING_LDGSTS=1 (default): eight 16-byte cp.async (LDGSTS) per lane -- a warp instruction moves 4 whole records,
//     8 of them a batch -- whose completion arrives on the stage's mbarrier (cp.async.mbarrier.arrive.noinc; the
//     barrier expects the 32 lanes of the filling warp);
//   * SB_RING_LDGSTS=0: one cp.async.bulk (TMA engine, 1-D copy) per record, completion counted in bytes
//     (expect_tx / complete_tx).  This was the gather until the forward stopped being issue-bound on the 1080p-class
//     configs: the TMA engine serves ~one request per 40 cycles and SM whatever its size, which at 128 B per request
//     caps a CTA-per-tile walk at 25 M records in 3.4 ms (measured: C5 forward 3.45 ms with half of all stall samples
//     on the stage wait, profiles/r02_after2_C5_blend_forward_lines.txt).  A/B on one box (profiles/
//     r02_ab_variants.txt, block r02o), cp.async vs TMA: forward C2 0.473 vs 0.493 ms, C4 2.06 vs 2.63, C5 2.69 vs
//     3.47; backward (whose fill already skips 4 of 5 records there) 1.140 vs 1.140, 3.25 vs 3.16, 4.15 vs 4.03.
#pragma once
#include "common.cuh"

#ifndef SB_RING_LDGSTS
#define SB_RING_LDGSTS 1
#endif

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

#if SB_RING_LDGSTS
#define BB_RING_ARRIVALS 32   // every lane of the filling warp arrives (asynchronously) once per fill
__device__ __forceinline__ void bb_ldgsts16(void* dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(bb_smem(dst)), "l"(src) : "memory");
}
// the arrival is triggered when all cp.async this thread issued before it have landed (at once if there are none)
__device__ __forceinline__ void bb_ldgsts_arrive(uint64_t* bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bb_smem(bar)) : "memory");
}
#else
#define BB_RING_ARRIVALS 1    // lane 0 arrives with the byte count the bulk copies will complete
#endif

// no copy of this thread may still be in flight when its CTA retires (a tile can stop with fills outstanding)
__device__ __forceinline__ void bb_ring_drain() {
#if SB_RING_LDGSTS
    asm volatile("cp.async.wait_all;" ::: "memory");
#endif
}

// Gather the records of one batch into a ring stage; called by one whole warp.  `id` = the surfel id of record `lane`
// (any value where the record is not wanted), bit r of `want` = record r is gathered, PITCH = float4s per ring slot.
template <int PITCH>
__device__ __forceinline__ void bb_gather(float4 (*dst)[PITCH], const float4* __restrict__ rec, uint32_t id, uint32_t want,
                                          uint64_t* bar, int lane) {
#if SB_RING_LDGSTS
    const int sub = lane >> 3, q = lane & 7;   // lanes 8s..8s+7 move the 8 quads of one record: 128 contiguous bytes
#pragma unroll
    for (int k = 0; k < 8; k++) {
        if (((want >> (4 * k)) & 0xFu) == 0u) continue;   // (warp-uniform)
        const int r = 4 * k + sub;
        const uint32_t rid = __shfl_sync(0xffffffffu, id, r);
        if ((want >> r) & 1u) bb_ldgsts16(&dst[r][q], rec + (size_t)rid * SB_REC_F4 + q);
    }
    __threadfence_block();   // the filling warp's plain stores to the stage (ids, masks) precede its arrivals
    bb_ldgsts_arrive(bar);
#else
    if (lane == 0) {
        if (want) bb_mbar_expect_tx(bar, (uint32_t)__popc(want) * (SB_REC_F4 * 16));
        else bb_mbar_arrive(bar);   // nothing to wait for: the phase completes at once
    }
    __syncwarp();
    if ((want >> lane) & 1u) bb_bulk_g2s(&dst[lane][0], rec + (size_t)id * SB_REC_F4, SB_REC_F4 * 16, bar);
#endif
}
