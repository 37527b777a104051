// Cross-GPU hand-off between the kernels of the multi-GPU exchange (dp_exchange.cu, surfel_update.cu): monotonic
// step counters in peer-mapped memory, written with system-scope release stores by the last CTA of a producing kernel
// and awaited with system-scope acquire loads by every CTA of the consuming kernel.
#pragma once
#include "common.cuh"

#define SB_DP_MAX_RANKS 16

struct SbDpSync {
    uint32_t* peer_flags[SB_DP_MAX_RANKS];   // rank q's flag block u32[2][SB_DP_MAX_RANKS]: [channel][source rank]
    uint32_t* counter;                       // local u32[2]: finished CTAs of the signalling kernel, per channel
    uint32_t* error;                         // local u32: a wait gave up (peer gone)
    unsigned long long* wait_ns;             // local u64[2]: time CTA 0 spent waiting, per channel (diagnostics)
    uint32_t step;                           // number of this exchange step (increasing, wraps)
    int rank, n_ranks;                       // n_ranks <= 1: nothing to synchronise
};

// End of a producing kernel (ALL threads of every CTA call it, none may have returned): the last CTA to finish tells
// every peer that this rank's rows of step `step` are complete.
__device__ __forceinline__ void sb_dp_signal(const SbDpSync& y, int ch) {
    if (y.n_ranks <= 1) return;
    __shared__ bool s_last;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();   // cumulative over the CTA's stores (ordered before it by the barrier)
        s_last = atomicAdd(y.counter + ch, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    if (threadIdx.x == 0) y.counter[ch] = 0;   // for the next launch (stream ordered behind this kernel)
    if ((int)threadIdx.x < y.n_ranks && (int)threadIdx.x != y.rank) {
        __threadfence_system();
        uint32_t* f = y.peer_flags[threadIdx.x] + ch * SB_DP_MAX_RANKS + y.rank;
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f), "r"(y.step) : "memory");
    }
}

// Start of a consuming kernel (ALL threads call it): wait until every peer has signalled step `step` on channel ch.
__device__ __forceinline__ void sb_dp_wait(const SbDpSync& y, int ch) {
    if (y.n_ranks <= 1) return;
    unsigned long long w0 = 0;
    if (blockIdx.x == 0 && threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(w0));
    if ((int)threadIdx.x < y.n_ranks && (int)threadIdx.x != y.rank) {
        const uint32_t* f = y.peer_flags[y.rank] + ch * SB_DP_MAX_RANKS + threadIdx.x;
        unsigned long long t0 = 0;
        for (uint32_t spins = 0;; spins++) {
            uint32_t v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
            if ((int)(v - y.step) >= 0) break;
            if ((spins & 1023u) == 1023u) {   // a peer that never arrives must not hang the GPU: give up after 20 s
                unsigned long long t;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                if (t0 == 0) t0 = t;
                else if (t - t0 > 20000000000ull) { atomicExch(y.error, 1u); break; }
            }
        }
    }
    __syncthreads();
    if (blockIdx.x == 0 && threadIdx.x == 0) {   // how long the first CTA waited for the slowest peer (rank skew)
        unsigned long long w1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(w1));
        y.wait_ns[ch] += w1 - w0;
    }
}

