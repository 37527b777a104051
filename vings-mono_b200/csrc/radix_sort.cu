// K2b: own onesweep LSD radix sort of (u32 key, u32 value) pairs.
//
// Replaces cub::DeviceRadixSort::SortPairs at DSR/cuda_rasterizer/rasterizer_impl.cu:410-418.  The reference
// sorts R (tile << 32 | depth) keys of 64 bits -- six or seven 8-bit passes over every (tile, surfel) pair.  A
// stable LSD sort on that key is the same permutation as
//   (1) a stable sort of the SURFELS by their 32-bit depth word (four passes over V visible surfels, V ~ R / 8),
//   (2) emitting the pairs in that surfel order (binning.cu), and
//   (3) a stable sort of the pairs by tile id alone (two passes of <= 8 bits for up to 65 536 tiles),
// because equal (tile, depth) keys keep emission order, which is surfel order, in both.  Both sorts run on the
// kernel below.  Structure of one sort:
//   1. digit bases: a histogram kernel + scan (surfel sort) or the tile-count scan of binning.cu (pair sort),
//   2. one "onesweep" kernel per pass: a CTA takes a tile of 256 x ITEMS pairs (ticket order), ranks its keys
//      with warp match-any multisplit, resolves each digit's global offset by decoupled look-back over the
//      preceding tiles' status words (no separate scan pass, keys are read once and written once per pass),
//      reorders the tile in shared memory and writes digit runs coalesced.
// Stability: items keep warp-striped index order inside a tile, warps are prefix-summed in order, tiles are
// chained in ticket order == index order.
// The first pass of the surfel sort reads the keys in surfel order with implicit values (the index) and DROPS
// the invisible surfels (key 0xFFFFFFFF): a stable compaction for free; later passes take their length from the
// device-side count.
#include "common.cuh"

#define ST_FLAG_AGG 0x40000000u
#define ST_FLAG_INC 0x80000000u
#define ST_VAL_MASK 0x3fffffffu

#define HIST_THREADS 512
#ifndef SB_LOOKBACK
#define SB_LOOKBACK 8
#endif

// hist[p][d] += number of valid keys whose pass-p digit is d
__global__ void __launch_bounds__(HIST_THREADS)
sb_sort_hist_kernel(const uint32_t* __restrict__ keys, int64_t n, int passes, int width, int end_bit, int drop_invalid,
                    uint32_t* __restrict__ hist) {
    __shared__ uint32_t sh[SB_MAX_PASSES * SB_RADIX];
    for (int i = threadIdx.x; i < passes * SB_RADIX; i += HIST_THREADS) sh[i] = 0;
    __syncthreads();
    for (int64_t i0 = (int64_t)blockIdx.x * HIST_THREADS; i0 < n; i0 += (int64_t)gridDim.x * HIST_THREADS) {
        const int64_t i = i0 + threadIdx.x;
        const uint32_t key = i < n ? keys[i] : SB_KEY_INVALID;
        const bool valid = i < n && !(drop_invalid && key == SB_KEY_INVALID);
        if (!valid) continue;
        for (int p = 0; p < passes; p++) {
            const int bits = min(width, end_bit - p * width);
            atomicAdd(&sh[p * SB_RADIX + ((key >> (p * width)) & ((1u << bits) - 1u))], 1u);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * SB_RADIX; i += HIST_THREADS)
        if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

// one CTA of 256 threads per pass: hist[p][0..256) -> exclusive prefix in place; pass 0 also stores the total
__global__ void __launch_bounds__(256) sb_sort_scan_kernel(uint32_t* __restrict__ hist, uint32_t* __restrict__ n_valid) {
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
    if (blockIdx.x == 0 && threadIdx.x == 255 && n_valid) *n_valid = add + incl;
}

template <int ITEMS>
struct SortSmem {
    uint32_t keys[SB_SORT_THREADS * ITEMS];
    uint32_t vals[SB_SORT_THREADS * ITEMS];
    uint32_t whist[SB_SORT_THREADS / 32][SB_RADIX];
    uint32_t lbase[SB_RADIX];
    uint32_t delta[SB_RADIX];
    uint32_t wsum[SB_SORT_THREADS / 32];
    uint32_t tile;
    uint32_t count;
};

// vals_in == nullptr : first pass of the surfel sort (value = index, keys equal to SB_KEY_INVALID are dropped)
// keys_out == nullptr: last pass (only the values are wanted)
// n_dev != nullptr   : the live length is *n_dev (<= n)
template <int ITEMS>
__global__ void __launch_bounds__(SB_SORT_THREADS)
sb_onesweep_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                   uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int64_t n,
                   const uint32_t* __restrict__ n_dev, int shift, int nbits,
                   const uint32_t* __restrict__ digit_base, uint32_t* __restrict__ ticket,
                   uint32_t* __restrict__ status) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SortSmem<ITEMS>& sm = *reinterpret_cast<SortSmem<ITEMS>*>(smem_raw);
    constexpr int TILE = SB_SORT_THREADS * ITEMS;
    constexpr int NW = SB_SORT_THREADS / 32;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t mask = (1u << nbits) - 1u;

    if (tid == 0) sm.tile = atomicAdd(ticket, 1u);
    for (int i = tid; i < NW * SB_RADIX; i += SB_SORT_THREADS) (&sm.whist[0][0])[i] = 0;
    if (n_dev) n = min(n, (int64_t)__ldg(n_dev));
    __syncthreads();
    const uint32_t tile = sm.tile;
    const int64_t base = (int64_t)tile * TILE;
    if (base >= n) return;  // whole CTA; nobody looks back at a tile past the end
    const int64_t wbase = base + (int64_t)wid * (32 * ITEMS);
    const bool first = vals_in == nullptr;

    uint32_t key[ITEMS];
    uint32_t val[ITEMS];
    uint32_t rank[ITEMS];
    uint32_t vmask = 0;  // bit i: item i is live
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const int64_t idx = wbase + i * 32 + lane;
        const bool in = idx < n;
        key[i] = in ? keys_in[idx] : SB_KEY_INVALID;
        val[i] = first ? (uint32_t)idx : (in ? vals_in[idx] : 0u);
        if (in && !(first && key[i] == SB_KEY_INVALID)) vmask |= 1u << i;
    }
    // ---- rank inside the warp (multisplit by match-any) ----
    const unsigned lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const bool valid = (vmask >> i) & 1u;
        const uint32_t d = (key[i] >> shift) & mask;
        // (a ballot per digit bit instead of match.any was measured 3x slower per pass on B200)
        const unsigned peers = __match_any_sync(0xffffffffu, valid ? d : 0xffffffffu);
        const int leader = __ffs(peers) - 1;
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
            // look back SB_LOOKBACK predecessors per round trip: with every tile of a short input resident at once
            // the walk is a serial chain of L2 latencies, one hop per load otherwise
            int64_t t = (int64_t)tile - 1;
            bool done = false;
            {   // steady state: the predecessor already holds its inclusive prefix
                const uint32_t w0 = st[(size_t)t * SB_RADIX + d];
                if (w0 & ST_FLAG_INC) { prefix = w0 & ST_VAL_MASK; done = true; }
            }
            while (!done) {
                uint32_t w[SB_LOOKBACK];
#pragma unroll
                for (int k = 0; k < SB_LOOKBACK; k++)
                    w[k] = (t - k >= 0) ? st[(size_t)(t - k) * SB_RADIX + d] : ST_FLAG_INC;  // virtual tile -1
                int used = 0;
                bool stop = false;
#pragma unroll
                for (int k = 0; k < SB_LOOKBACK; k++) {
                    if (!stop) {
                        if ((w[k] & (ST_FLAG_AGG | ST_FLAG_INC)) == 0) {
                            stop = true;  // not published yet: retry from here
                        } else {
                            prefix += w[k] & ST_VAL_MASK;
                            used++;
                            if (w[k] & ST_FLAG_INC) { stop = true; done = true; }
                        }
                    }
                }
                t -= used;
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
        const uint32_t lb = add + incl - run;
        sm.lbase[d] = lb;
        sm.delta[d] = digit_base[d] + prefix - lb;
        if (tid == SB_SORT_THREADS - 1) sm.count = add + incl;
    }
    __syncthreads();
    // ---- reorder the tile in shared memory ----
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        if ((vmask >> i) & 1u) {
            const uint32_t d = (key[i] >> shift) & mask;
            const uint32_t pos = sm.lbase[d] + sm.whist[wid][d] + rank[i];
            sm.keys[pos] = key[i];
            sm.vals[pos] = val[i];
        }
    }
    __syncthreads();
    // ---- coalesced write-out of the digit runs ----
    const int count = (int)sm.count;
    for (int j = tid; j < count; j += SB_SORT_THREADS) {
        const uint32_t k = sm.keys[j];
        const uint32_t d = (k >> shift) & mask;
        const uint32_t dst = (uint32_t)j + sm.delta[d];
        if (keys_out) keys_out[dst] = k;
        vals_out[dst] = sm.vals[j];
    }
}

// digit width of a sort over `end_bit` key bits: as few passes as 8-bit digits allow, bits spread evenly
SbSortPlan sb_sort_plan(int end_bit) {
    SbSortPlan pl;
    pl.end_bit = end_bit < 1 ? 1 : end_bit;
    pl.passes = (pl.end_bit + SB_RADIX_BITS - 1) / SB_RADIX_BITS;
    pl.width = (pl.end_bit + pl.passes - 1) / pl.passes;
    return pl;
}

template <int ITEMS>
static int sb_run_passes(const SbSortPlan& pl, const uint32_t* keys_in, const uint32_t* vals_in, uint32_t* const kbuf[2],
                         uint32_t* const vbuf[2], uint32_t* vals_final, int64_t n, const uint32_t* n_dev,
                         const uint32_t* digit_base, uint32_t* tickets, uint32_t* status, cudaStream_t s) {
    static bool attr_set[SB_MAX_DEVICES] = {};  // function attributes are per device
    const int dev = sb_current_device();
    if (!attr_set[dev]) {
        SB_CUDA_CHECK(cudaFuncSetAttribute(sb_onesweep_kernel<ITEMS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)sizeof(SortSmem<ITEMS>)));
        attr_set[dev] = true;
    }
    const int64_t ntiles = (n + SB_SORT_THREADS * ITEMS - 1) / (SB_SORT_THREADS * ITEMS);
    const uint32_t* kin = keys_in;
    const uint32_t* vin = vals_in;
    for (int p = 0; p < pl.passes; p++) {
        int bits = pl.end_bit - p * pl.width;
        if (bits > pl.width) bits = pl.width;
        const bool last = p == pl.passes - 1;
        uint32_t* kout = last ? nullptr : kbuf[p & 1];
        uint32_t* vout = last ? vals_final : vbuf[p & 1];
        sb_onesweep_kernel<ITEMS><<<(unsigned)ntiles, SB_SORT_THREADS, sizeof(SortSmem<ITEMS>), s>>>(
            kin, vin, kout, vout, n, (p == 0 && vals_in == nullptr) ? nullptr : n_dev, p * pl.width, bits,
            digit_base + p * SB_RADIX, tickets + p, status + (size_t)p * ntiles * SB_RADIX);
        kin = kout;
        vin = vout;
    }
    sb_count_launch(pl.passes);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}

// Digit histograms of every pass + exclusive scan -> digit_base[passes][256]; *n_valid = number of live keys.
int sb_launch_sort_hist(const uint32_t* keys, int64_t n, const SbSortPlan& pl, int drop_invalid, uint32_t* digit_base,
                        uint32_t* n_valid, cudaStream_t s) {
    if (n <= 0) return 0;
    int grid = (int)((n + HIST_THREADS * 8 - 1) / (HIST_THREADS * 8));
    if (grid > 148 * 2) grid = 148 * 2;
    sb_sort_hist_kernel<<<grid, HIST_THREADS, 0, s>>>(keys, n, pl.passes, pl.width, pl.end_bit, drop_invalid, digit_base);
    sb_sort_scan_kernel<<<pl.passes, 256, 0, s>>>(digit_base, n_valid);
    sb_count_launch(2);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}

// Runs the passes of one sort.  keys_in/vals_in feed pass 0 (vals_in == nullptr: values are the indices and invalid
// keys are dropped); passes ping-pong through kbuf/vbuf[0], [1], [0], ...; the last pass writes only values, to
// vals_final.  digit_base [passes][256] must hold the exclusive digit offsets; tickets [passes] and status
// [passes * ntiles * 256] must be zero.  small != 0 selects 2048-pair tiles (more CTAs for short inputs).
int sb_launch_sort_passes(const SbSortPlan& pl, int small, const uint32_t* keys_in, const uint32_t* vals_in,
                          uint32_t* const kbuf[2], uint32_t* const vbuf[2], uint32_t* vals_final, int64_t n,
                          const uint32_t* n_dev, const uint32_t* digit_base, uint32_t* tickets, uint32_t* status,
                          cudaStream_t s) {
    if (n <= 0) return 0;
    if (n >= (int64_t)ST_VAL_MASK) { sb_set_error("sort: %lld pairs exceed the 2^30 look-back word", (long long)n); return 1; }
    if (pl.passes < 1 || pl.passes > SB_MAX_PASSES) { sb_set_error("sort: bad end_bit %d", pl.end_bit); return 1; }
    if (small)
        return sb_run_passes<SB_SORT_ITEMS_SMALL>(pl, keys_in, vals_in, kbuf, vbuf, vals_final, n, n_dev, digit_base,
                                                  tickets, status, s);
    return sb_run_passes<SB_SORT_ITEMS>(pl, keys_in, vals_in, kbuf, vbuf, vals_final, n, n_dev, digit_base, tickets,
                                        status, s);
}
