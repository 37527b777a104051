// K2a/K2c: (tile, surfel) pair emission in depth order, tile ranges and tile launch order.
//
// Replaces duplicateWithKeys (DSR/cuda_rasterizer/rasterizer_impl.cu:123-164), the cub::DeviceScan over
// tiles_touched (:387) and identifyTileRanges (:171-196).  The reference emits key = (tile_id << 32) |
// float_bits(view depth), value = surfel index in surfel order and sorts the 64-bit keys.  Here the visible
// surfels are already sorted by depth (stable, radix_sort.cu), so the pairs are emitted in (depth, surfel id)
// order with the tile id alone as key: a stable sort by tile id then yields exactly the reference's order, and
// the tile ranges are the exclusive scan of the per-tile pair counts, which the emission accumulates.
//
//   sb_pair_offsets_kernel  inclusive prefix sum of tiles_touched over the depth-ordered surfels (single pass,
//                           decoupled look-back), plus, for every 1024-pair chunk of the output, the surfel that
//                           owns the chunk's first pair.
//   sb_duplicate_kernel     persistent CTAs take 1024-pair chunks: the chunk's surfels (offset, id, rect) are
//                           staged in shared memory, every thread locates the owner of its pairs by binary search
//                           and writes (tile, id); work is balanced by PAIRS, not by surfels - in depth order the
//                           first CTA's 256 surfels are the nearest ones and can cover the whole screen each
//                           (430 k pairs at C2 against a mean of 2.8 k), which is also why the reference's
//                           one-thread-per-surfel loop has a long tail.  Tile counts go to a shared-memory
//                           histogram flushed once per CTA.
#include "common.cuh"

#define OFF_THREADS 256
#define OFF_ITEMS 4
#define OFF_TILE (OFF_THREADS * OFF_ITEMS)
// look-back status word: [63:62] flag (1 = aggregate, 2 = inclusive prefix), [31:0] value
#define SCAN_FLAG_AGG (1ull << 62)
#define SCAN_FLAG_INC (2ull << 62)

__global__ void __launch_bounds__(OFF_THREADS)
sb_pair_offsets_kernel(const uint32_t* __restrict__ sorted_ids, const uint32_t* __restrict__ n_visible,
                       const uint32_t* __restrict__ tiles_touched, uint32_t* __restrict__ ticket,
                       unsigned long long* __restrict__ scan_state, uint32_t* __restrict__ offsets,
                       uint32_t* __restrict__ chunk_first, uint32_t max_chunks) {
    __shared__ uint32_t s_warp[OFF_THREADS / 32];
    __shared__ uint32_t s_blk, s_prefix;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    // ticket = logical block index; guarantees every predecessor has started (look-back progress)
    if (threadIdx.x == 0) s_blk = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t blk = s_blk;
    const uint32_t V = __ldg(n_visible);
    if (blk * OFF_TILE >= V) return;  // whole CTA
    // thread t owns the 4 consecutive surfels blk*1024 + 4t .. +3 of the depth order
    const uint32_t i0 = blk * OFF_TILE + threadIdx.x * OFF_ITEMS;
    uint32_t cnt[OFF_ITEMS];
    uint32_t tsum = 0;
    if (i0 + OFF_ITEMS <= V) {
        const uint4 ids = *reinterpret_cast<const uint4*>(sorted_ids + i0);
        cnt[0] = tiles_touched[ids.x]; cnt[1] = tiles_touched[ids.y];
        cnt[2] = tiles_touched[ids.z]; cnt[3] = tiles_touched[ids.w];
    } else {
#pragma unroll
        for (int k = 0; k < OFF_ITEMS; k++) cnt[k] = (i0 + k < V) ? tiles_touched[sorted_ids[i0 + k]] : 0;
    }
#pragma unroll
    for (int k = 0; k < OFF_ITEMS; k++) tsum += cnt[k];
    uint32_t incl = tsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) s_warp[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        uint32_t w = (lane < OFF_THREADS / 32) ? s_warp[lane] : 0;
#pragma unroll
        for (int o = 1; o < OFF_THREADS / 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += v;
        }
        if (lane < OFF_THREADS / 32) s_warp[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const uint32_t block_total = s_warp[OFF_THREADS / 32 - 1];
    const uint32_t warp_excl = wid ? s_warp[wid - 1] : 0;

    // ---- decoupled look-back across CTAs (one warp) ----
    if (wid == 0) {
        volatile unsigned long long* st = scan_state;
        if (lane == 0) st[blk] = (blk == 0 ? SCAN_FLAG_INC : SCAN_FLAG_AGG) | block_total;
        uint32_t prefix = 0;
        if (blk > 0) {
            int look = (int)blk - 1 - lane;  // each lane inspects one predecessor per round
            while (true) {
                unsigned long long w = 0;
                if (look >= 0) {
                    do { w = st[look]; } while ((w >> 62) == 0);
                } else {
                    w = SCAN_FLAG_INC;  // virtual predecessor before block 0
                }
                const unsigned inc_mask = __ballot_sync(0xffffffffu, (w >> 62) == 2);
                uint32_t val = (uint32_t)w;
                if (inc_mask) {
                    const int first = __ffs(inc_mask) - 1;  // nearest predecessor holding an inclusive prefix
                    if (lane > first) val = 0;
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
                prefix += val;
                if (inc_mask) break;
                look -= 32;
            }
            if (lane == 0) {
                __threadfence();
                st[blk] = SCAN_FLAG_INC | (unsigned long long)(prefix + block_total);
            }
        }
        if (lane == 0) s_prefix = prefix;
    }
    __syncthreads();
    uint32_t run = s_prefix + warp_excl + incl - tsum;  // exclusive offset of this thread's first surfel
#pragma unroll
    for (int k = 0; k < OFF_ITEMS; k++) {
        if (i0 + k < V) {
            const uint32_t e = run;
            run += cnt[k];
            offsets[i0 + k] = run;
            // this surfel owns the first pair of every chunk that starts inside [e, run)
            for (uint32_t c = (e + SB_EMIT_CHUNK - 1) / SB_EMIT_CHUNK; c * SB_EMIT_CHUNK < run && c < max_chunks; c++)
                chunk_first[c] = i0 + k;
        }
    }
}

#define DUP_THREADS 256
static_assert(SB_EMIT_CHUNK / DUP_THREADS == 4, "the duplicate kernel emits 4 pairs per thread");
#define DUP_WINDOW (SB_EMIT_CHUNK + DUP_THREADS)  // a chunk of n pairs starts in at most n surfels

// dynamic shared memory: uint32 s_hist[hist_words] (tile counts of this CTA; 0 words = count with global REDs)
__global__ void __launch_bounds__(DUP_THREADS)
sb_duplicate_kernel(uint32_t grid_x, int64_t R, const unsigned long long* __restrict__ R_dev,
                    const uint32_t* __restrict__ sorted_ids,
                    const uint32_t* __restrict__ n_visible, const ushort4* __restrict__ rects,
                    const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ chunk_first,
                    uint32_t* __restrict__ tile_counts, uint32_t* __restrict__ keys, uint32_t* __restrict__ vals,
                    int hist_words) {
    extern __shared__ uint32_t s_hist[];
    __shared__ uint32_t s_incl[DUP_WINDOW], s_id[DUP_WINDOW], s_rmin[DUP_WINDOW], s_rw[DUP_WINDOW];
    const int tid = threadIdx.x;
    const uint32_t V = __ldg(n_visible);
    for (int t = tid; t < hist_words; t += DUP_THREADS) s_hist[t] = 0;
    if (R_dev) R = min(R, (int64_t)__ldg(R_dev));  // R is the scratch capacity: emit what fits (caller re-runs)
    const uint32_t nchunks = (uint32_t)((R + SB_EMIT_CHUNK - 1) / SB_EMIT_CHUNK);
    for (uint32_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
        const uint32_t begin = c * SB_EMIT_CHUNK;
        const uint32_t end = (uint32_t)min((int64_t)begin + SB_EMIT_CHUNK, R);
        const uint32_t s0 = __ldg(chunk_first + c);
        const uint32_t base_excl = s0 ? __ldg(offsets + s0 - 1) : 0u;
        __syncthreads();  // previous chunk's window is no longer read; s_hist zeroing is visible
        // ---- stage the chunk's surfels: s0, s0+1, ... up to the one whose inclusive offset reaches `end` ----
        int n = 0;
        while (true) {
            const uint32_t i = s0 + n + tid;
            uint32_t o = 0xffffffffu;
            if (i < V) {
                o = __ldg(offsets + i);
                const uint32_t id = __ldg(sorted_ids + i);
                const ushort4 r = __ldg(rects + id);
                s_id[n + tid] = id;
                s_rmin[n + tid] = (uint32_t)r.x | ((uint32_t)r.y << 16);
                s_rw[n + tid] = (uint32_t)r.z - (uint32_t)r.x;
            }
            s_incl[n + tid] = o;
            n += DUP_THREADS;
            // the last surfel loaded so far (thread 255's) decides whether the window is complete
            if (__syncthreads_or(tid == DUP_THREADS - 1 && o >= end) || n >= DUP_WINDOW) break;
        }
        // ---- every thread takes 4 consecutive pairs: one binary search (first k with s_incl[k] > j) for the first,
        // the others step forward from it; the 4 (tile, id) results leave as one 128-bit store each ----
        {
            constexpr int PER = SB_EMIT_CHUNK / DUP_THREADS;
            const uint32_t j0 = begin + PER * tid;
            uint32_t tk[PER], tv[PER];
            if (j0 < end) {
                int lo = 0, len = n;
                while (len > 0) {
                    const int half = len >> 1;
                    if (s_incl[lo + half] <= j0) { lo += half + 1; len -= half + 1; } else { len = half; }
                }
                uint32_t excl = lo ? s_incl[lo - 1] : base_excl;
                uint32_t incl = s_incl[lo], rw = s_rw[lo], rmin = s_rmin[lo], id = s_id[lo];
#pragma unroll
                for (int m = 0; m < PER; m++) {
                    const uint32_t j = j0 + m;
                    tk[m] = 0; tv[m] = 0;
                    if (j < end) {
                        while (j >= incl) {  // next surfel (every staged surfel owns at least one pair)
                            lo++;
                            excl = incl;
                            incl = s_incl[lo]; rw = s_rw[lo]; rmin = s_rmin[lo]; id = s_id[lo];
                        }
                        const uint32_t t = j - excl;
                        const uint32_t ty = t / rw, tx = t - ty * rw;
                        const uint32_t tile = ((rmin >> 16) + ty) * grid_x + (rmin & 0xffffu) + tx;
                        tk[m] = tile; tv[m] = id;
                        if (hist_words) atomicAdd(&s_hist[tile], 1u);
                        else atomicAdd(tile_counts + tile, 1u);
                    }
                }
                if (j0 + PER <= end) {  // chunk starts are multiples of 1024: 16-byte aligned
                    *reinterpret_cast<uint4*>(keys + j0) = make_uint4(tk[0], tk[1], tk[2], tk[3]);
                    *reinterpret_cast<uint4*>(vals + j0) = make_uint4(tv[0], tv[1], tv[2], tv[3]);
                } else {
#pragma unroll
                    for (int m = 0; m < PER; m++)
                        if (j0 + m < end) { keys[j0 + m] = tk[m]; vals[j0 + m] = tv[m]; }
                }
            }
        }
    }
    __syncthreads();
    for (int t = tid; t < hist_words; t += DUP_THREADS)
        if (s_hist[t]) atomicAdd(tile_counts + t, s_hist[t]);
}

// R_is_capacity: R is the size of the caller's scratch and the live count is read from the geometry workspace
int sb_launch_duplicate(int P, int W, int H, int64_t R, int R_is_capacity, const char* geom, uint32_t* chunk_first,
                        uint32_t* tile_counts, uint32_t* keys, uint32_t* vals, cudaStream_t s) {
    if (P == 0 || R <= 0) return 0;
    SbGeomLayout L = sb_geom_layout(P);
    const uint32_t grid_x = (W + SB_TILE_X - 1) / SB_TILE_X;
    const int tiles = (int)grid_x * ((H + SB_TILE_Y - 1) / SB_TILE_Y);
    const uint32_t* sorted_ids = (const uint32_t*)(geom + L.sort_v[1]);
    const uint32_t* n_visible = (const uint32_t*)(geom + L.counts + 8);
    uint32_t* offsets = (uint32_t*)(geom + L.sort_k[0]);  // free once the surfel sort is done
    const int oblocks = (P + OFF_TILE - 1) / OFF_TILE;    // CTAs past the V visible surfels exit at once
    sb_pair_offsets_kernel<<<oblocks, OFF_THREADS, 0, s>>>(sorted_ids, n_visible, (const uint32_t*)(geom + L.tiles),
                                                           (uint32_t*)(geom + L.tickets) + 4,
                                                           (unsigned long long*)(geom + L.dup_state), offsets,
                                                           chunk_first,
                                                           (uint32_t)((R + SB_EMIT_CHUNK - 1) / SB_EMIT_CHUNK));
    // shared-memory tile histogram when it fits beside the window (up to ~40 k tiles), else global REDs
    const int max_dyn = 160 * 1024;
    static bool attr_set[SB_MAX_DEVICES] = {};  // function attributes are per device
    const int dev = sb_current_device();
    if (!attr_set[dev]) {
        SB_CUDA_CHECK(cudaFuncSetAttribute(sb_duplicate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_dyn));
        attr_set[dev] = true;
    }
    const int hist_words = (size_t)tiles * 4 <= (size_t)max_dyn ? tiles : 0;
    const int64_t nchunks = (R + SB_EMIT_CHUNK - 1) / SB_EMIT_CHUNK;
    const int per_sm = hist_words * 4 > 64 * 1024 ? 1 : (hist_words * 4 > 24 * 1024 ? 2 : 4);
    const int grid = (int)(nchunks < 148 * per_sm ? nchunks : 148 * per_sm);
    sb_duplicate_kernel<<<grid, DUP_THREADS, (size_t)hist_words * 4, s>>>(
        grid_x, R, R_is_capacity ? (const unsigned long long*)(geom + L.counts) : nullptr, sorted_ids, n_visible, (const ushort4*)(geom + L.rect), offsets, chunk_first, tile_counts, keys, vals,
        hist_words);
    sb_count_launch(2);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}

// One CTA over the per-tile pair counts: tile ranges (exclusive scan in tile order, identifyTileRanges), the
// exclusive digit bases of the pair sort's passes (a pass-p digit's count is the sum over the tiles carrying it),
// and the longest-list-first launch order of the tiles for the two blend kernels: a CTA's run time is
// proportional to its tile's list length, lists differ by 100x across an image, and blockIdx order is dispatch
// order, so the long tiles start first and the short ones fill the tail (LPT scheduling; 1024-bin counting sort
// by scaled length, order inside a bin is arbitrary).
#define ORD_THREADS 1024
__global__ void __launch_bounds__(ORD_THREADS)
sb_tile_scan_kernel(int tiles, int passes, int width, int end_bit, const uint32_t* __restrict__ counts,
                    uint2* __restrict__ ranges, uint32_t* __restrict__ order, uint32_t* __restrict__ digit_base) {
    __shared__ uint32_t s_bin[ORD_THREADS];
    __shared__ uint32_t s_dig[SB_MAX_PASSES * SB_RADIX];
    __shared__ uint32_t s_w[32];
    __shared__ uint32_t s_max, s_carry;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) { s_max = 0; s_carry = 0; }
    s_bin[tid] = 0;
    for (int i = tid; i < passes * SB_RADIX; i += ORD_THREADS) s_dig[i] = 0;
    __syncthreads();
    // ---- ranges, digit counts, longest list ----
    uint32_t mx = 0;
    for (int t0 = 0; t0 < tiles; t0 += ORD_THREADS) {
        const int t = t0 + tid;
        const uint32_t v = t < tiles ? counts[t] : 0;
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += u;
        }
        if (lane == 31) s_w[wid] = incl;
        __syncthreads();
        uint32_t add = s_carry;
        for (int w = 0; w < wid; w++) add += s_w[w];
        if (t < tiles) {
            ranges[t] = v ? make_uint2(add + incl - v, add + incl) : make_uint2(0u, 0u);  // untouched tile: (0,0)
            if (v) {
                for (int p = 0; p < passes; p++) {
                    const int bits = min(width, end_bit - p * width);
                    atomicAdd(&s_dig[p * SB_RADIX + (((uint32_t)t >> (p * width)) & ((1u << bits) - 1u))], v);
                }
            }
        }
        mx = max(mx, v);
        __syncthreads();
        if (tid == ORD_THREADS - 1) s_carry = add + incl;
        __syncthreads();
    }
    mx = __reduce_max_sync(0xffffffffu, mx);
    if (lane == 0 && mx) atomicMax(&s_max, mx);
    __syncthreads();
    // ---- exclusive scan of each pass's 256 digit counts (warp p*8+k owns digits 32k..32k+31 of pass p) ----
    if (tid < passes * SB_RADIX) {
        const uint32_t v = s_dig[tid];
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += u;
        }
        if (lane == 31) s_w[wid] = incl;
        __syncwarp();
        // the 8 warps of a pass are consecutive; wait for all of them below
        s_dig[tid] = incl - v;
    }
    __syncthreads();
    if (digit_base && tid < passes * SB_RADIX) {  // no pair sort (and no scratch) when R == 0
        uint32_t add = 0;
        for (int w = wid & ~7; w < wid; w++) add += s_w[w];
        digit_base[tid] = s_dig[tid] + add;
    }
    // ---- launch order ----
    const uint64_t top = s_max ? s_max : 1;
    // bin 0 = longest
    for (int t = tid; t < tiles; t += ORD_THREADS) {
        const uint32_t n = counts[t];
        atomicAdd(&s_bin[(ORD_THREADS - 1) - (uint32_t)(((uint64_t)n * (ORD_THREADS - 1)) / top)], 1u);
    }
    __syncthreads();
    // exclusive scan of the 1024 bins (Hillis-Steele over a copy in registers)
    uint32_t v = s_bin[tid];
    const uint32_t own = v;
    for (int o = 1; o < ORD_THREADS; o <<= 1) {
        const uint32_t add = tid >= o ? s_bin[tid - o] : 0;
        __syncthreads();
        v += add;
        s_bin[tid] = v;
        __syncthreads();
    }
    s_bin[tid] = v - own;
    __syncthreads();
    for (int t = tid; t < tiles; t += ORD_THREADS) {
        const uint32_t n = counts[t];
        const uint32_t pos = atomicAdd(&s_bin[(ORD_THREADS - 1) - (uint32_t)(((uint64_t)n * (ORD_THREADS - 1)) / top)], 1u);
        order[pos] = (uint32_t)t;
    }
}

int sb_launch_tile_scan(int tiles, const SbSortPlan& pl, const uint32_t* tile_counts, uint2* ranges, uint32_t* order,
                        uint32_t* digit_base, cudaStream_t s) {
    sb_tile_scan_kernel<<<1, ORD_THREADS, 0, s>>>(tiles, pl.passes, pl.width, pl.end_bit, tile_counts, ranges, order,
                                                  digit_base);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}
