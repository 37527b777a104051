// K2b: own onesweep LSD radix sort of (u64 tile|depth key, u32 surfel index) pairs.
//
// Replaces cub::DeviceRadixSort::SortPairs at DSR/cuda_rasterizer/rasterizer_impl.cu:410-418: a
// *stable* least-significant-digit sort over bits [0, end_bit), 8 bits per pass.  Structure:
//   1. one histogram kernel reads the keys once and counts every pass's 256 digits,
//   2. a tiny kernel turns each histogram into exclusive digit bases,
//   3. one "onesweep" kernel per pass: a CTA takes a 4096-pair tile (ticket order), ranks its keys
//      with warp match-any multisplit, resolves each digit's global offset by decoupled look-back
//      over the preceding tiles' status words (no separate scan pass, keys are read once and written
//      once per pass), reorders the tile in shared memory and writes digit runs coalesced.
// Stability: items keep warp-striped index order inside a tile, warps are prefix-summed in order,
// tiles are chained in ticket order == index order.
#include "common.cuh"

#define ST_FLAG_AGG 0x40000000u
#define ST_FLAG_INC 0x80000000u
#define ST_VAL_MASK 0x3fffffffu

__global__ void __launch_bounds__(256)
sb_sort_hist_kernel(const uint64_t* __restrict__ keys, int64_t n, int passes, int end_bit,
                    uint32_t* __restrict__ hist) {
    __shared__ uint32_t sh[SB_MAX_PASSES * SB_RADIX];
    for (int i = threadIdx.x; i < passes * SB_RADIX; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x; i0 < n; i0 += (int64_t)gridDim.x * blockDim.x) {
        int64_t i = i0 + threadIdx.x;
        bool valid = i < n;
        uint64_t key = valid ? keys[i] : 0;
        for (int p = 0; p < passes; p++) {
            int bits = min(SB_RADIX_BITS, end_bit - p * SB_RADIX_BITS);
            uint32_t d = (uint32_t)(key >> (p * SB_RADIX_BITS)) & ((1u << bits) - 1u);
            // warp-aggregated shared atomics: neighbouring pairs share their tile digits
            unsigned peers = __match_any_sync(0xffffffffu, valid ? d : 0xffffffffu);
            if (valid && lane == __ffs(peers) - 1) atomicAdd(&sh[p * SB_RADIX + d], __popc(peers));
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * SB_RADIX; i += blockDim.x)
        if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

// one CTA of 256 threads per pass: hist[p][0..256) -> exclusive prefix in place
__global__ void __launch_bounds__(256) sb_sort_scan_kernel(uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_w[8];
    uint32_t* h = hist + blockIdx.x * SB_RADIX;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t v = h[threadIdx.x], incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_w[wid] = incl;
    __syncthreads();
    uint32_t add = 0;
    for (int w = 0; w < wid; w++) add += s_w[w];
    h[threadIdx.x] = add + incl - v;
}

struct SortSmem {
    uint64_t keys[SB_SORT_TILE];
    uint32_t vals[SB_SORT_TILE];
    uint32_t whist[SB_SORT_THREADS / 32][SB_RADIX];
    uint32_t lbase[SB_RADIX];
    uint32_t delta[SB_RADIX];
    uint32_t wsum[SB_SORT_THREADS / 32];
    uint32_t tile;
};

__global__ void __launch_bounds__(SB_SORT_THREADS)
sb_onesweep_kernel(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                   uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int64_t n, int shift,
                   uint32_t mask, const uint32_t* __restrict__ digit_base, uint32_t* __restrict__ ticket,
                   uint32_t* __restrict__ status) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SortSmem& sm = *reinterpret_cast<SortSmem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    constexpr int NW = SB_SORT_THREADS / 32;

    if (tid == 0) sm.tile = atomicAdd(ticket, 1u);
    for (int i = tid; i < NW * SB_RADIX; i += SB_SORT_THREADS) (&sm.whist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = sm.tile;
    const int64_t base = (int64_t)tile * SB_SORT_TILE;
    const int64_t wbase = base + (int64_t)wid * (32 * SB_SORT_ITEMS);

    uint64_t key[SB_SORT_ITEMS];
    uint32_t val[SB_SORT_ITEMS];
    uint32_t rank[SB_SORT_ITEMS];
#pragma unroll
    for (int i = 0; i < SB_SORT_ITEMS; i++) {
        int64_t idx = wbase + i * 32 + lane;
        bool valid = idx < n;
        key[i] = valid ? keys_in[idx] : ~0ull;
        val[i] = valid ? vals_in[idx] : 0u;
    }
    // ---- rank inside the warp (multisplit by match-any) ----
    const unsigned lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < SB_SORT_ITEMS; i++) {
        bool valid = (wbase + i * 32 + lane) < n;
        uint32_t d = (uint32_t)(key[i] >> shift) & mask;
        unsigned peers = __match_any_sync(0xffffffffu, valid ? d : 0xffffffffu);
        int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (valid && lane == leader) {
            old = sm.whist[wid][d];
            sm.whist[wid][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        rank[i] = old + __popc(peers & lt_mask);
        __syncwarp();
    }
    __syncthreads();
    // ---- per digit: exclusive prefix over warps, tile count, look-back ----
    {
        const int d = tid;  // SB_SORT_THREADS == SB_RADIX
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < NW; w++) {
            uint32_t t = sm.whist[w][d];
            sm.whist[w][d] = run;
            run += t;
        }
        volatile uint32_t* st = status;
        uint32_t prefix = 0;
        if (tile == 0) {
            st[d] = ST_FLAG_INC | run;
        } else {
            st[(size_t)tile * SB_RADIX + d] = ST_FLAG_AGG | run;
            int64_t t = (int64_t)tile - 1;
            while (true) {
                uint32_t w = st[(size_t)t * SB_RADIX + d];
                if ((w & (ST_FLAG_AGG | ST_FLAG_INC)) == 0) continue;
                prefix += w & ST_VAL_MASK;
                if (w & ST_FLAG_INC) break;
                t--;
            }
            st[(size_t)tile * SB_RADIX + d] = ST_FLAG_INC | (prefix + run);
        }
        // block exclusive scan of `run` over the 256 digits -> local base
        uint32_t incl = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) sm.wsum[wid] = incl;
        __syncthreads();
        uint32_t add = 0;
        for (int w = 0; w < wid; w++) add += sm.wsum[w];
        uint32_t lb = add + incl - run;
        sm.lbase[d] = lb;
        sm.delta[d] = digit_base[d] + prefix - lb;
    }
    __syncthreads();
    // ---- reorder the tile in shared memory ----
#pragma unroll
    for (int i = 0; i < SB_SORT_ITEMS; i++) {
        bool valid = (wbase + i * 32 + lane) < n;
        if (valid) {
            uint32_t d = (uint32_t)(key[i] >> shift) & mask;
            uint32_t pos = sm.lbase[d] + sm.whist[wid][d] + rank[i];
            sm.keys[pos] = key[i];
            sm.vals[pos] = val[i];
        }
    }
    __syncthreads();
    // ---- coalesced write-out of the digit runs ----
    int64_t remain = n - base;
    const int count = remain < SB_SORT_TILE ? (int)remain : SB_SORT_TILE;
    for (int j = tid; j < count; j += SB_SORT_THREADS) {
        uint64_t k = sm.keys[j];
        uint32_t d = (uint32_t)(k >> shift) & mask;
        uint32_t dst = (uint32_t)j + sm.delta[d];
        keys_out[dst] = k;
        vals_out[dst] = sm.vals[j];
    }
}

// Sorts n pairs on bits [0, end_bit).  keys_in/vals_in are consumed (ping-pong with keys_tmp/vals_tmp);
// the final pass writes values to vals_final; *keys_sorted receives the buffer holding the sorted keys.
int sb_launch_sort_pairs(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_tmp, uint32_t* vals_tmp,
                         uint64_t** keys_sorted, uint32_t* vals_final, int64_t n, int end_bit, char* bin_base,
                         const SbBinLayout& L, cudaStream_t s) {
    static bool attr_set = false;
    if (!attr_set) {
        SB_CUDA_CHECK(cudaFuncSetAttribute(sb_onesweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)sizeof(SortSmem)));
        attr_set = true;
    }
    *keys_sorted = const_cast<uint64_t*>(keys_in);
    if (n <= 0) return 0;
    if (n >= (int64_t)ST_VAL_MASK) { sb_set_error("sort: %lld pairs exceed the 2^30 look-back word", (long long)n); return 1; }
    const int passes = (end_bit + SB_RADIX_BITS - 1) / SB_RADIX_BITS;
    if (passes < 1 || passes > SB_MAX_PASSES) { sb_set_error("sort: bad end_bit %d", end_bit); return 1; }
    const int64_t ntiles = (n + SB_SORT_TILE - 1) / SB_SORT_TILE;
    uint32_t* hist = (uint32_t*)(bin_base + L.hist);
    uint32_t* tickets = (uint32_t*)(bin_base + L.tickets);
    uint32_t* status = (uint32_t*)(bin_base + L.status);
    size_t zero_bytes = (L.status - L.zero_begin) + (size_t)passes * ntiles * SB_RADIX * 4;
    SB_CUDA_CHECK(cudaMemsetAsync(bin_base + L.zero_begin, 0, zero_bytes, s));
    int hgrid = (int)((n + 256 * 16 - 1) / (256 * 16));
    if (hgrid > 148 * 8) hgrid = 148 * 8;
    sb_sort_hist_kernel<<<hgrid, 256, 0, s>>>(keys_in, n, passes, end_bit, hist);
    sb_sort_scan_kernel<<<passes, 256, 0, s>>>(hist);
    const uint64_t* kin = keys_in;
    const uint32_t* vin = vals_in;
    uint64_t* kbuf[2] = {keys_tmp, const_cast<uint64_t*>(keys_in)};
    uint32_t* vbuf[2] = {vals_tmp, const_cast<uint32_t*>(vals_in)};
    for (int p = 0; p < passes; p++) {
        int bits = end_bit - p * SB_RADIX_BITS;
        if (bits > SB_RADIX_BITS) bits = SB_RADIX_BITS;
        uint64_t* kout = kbuf[p & 1];
        uint32_t* vout = (p == passes - 1) ? vals_final : vbuf[p & 1];
        sb_onesweep_kernel<<<(unsigned)ntiles, SB_SORT_THREADS, sizeof(SortSmem), s>>>(
            kin, vin, kout, vout, n, p * SB_RADIX_BITS, (1u << bits) - 1u, hist + p * SB_RADIX, tickets + p,
            status + (size_t)p * ntiles * SB_RADIX);
        kin = kout;
        vin = vout;
    }
    sb_count_launch(2 + passes);
    SB_CUDA_CHECK(cudaGetLastError());
    *keys_sorted = const_cast<uint64_t*>(kin);
    return 0;
}
