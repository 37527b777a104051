// K2a/K2c: tile-key duplication and tile-range identification.
//
// Replaces duplicateWithKeys (DSR/cuda_rasterizer/rasterizer_impl.cu:123-164) and identifyTileRanges
// (:171-196).  Key = (tile_id << 32) | float_bits(view depth), value = surfel index; a surfel's pairs
// are emitted row-major over its tile rect starting at offsets[i] - tiles_touched[i], exactly the
// reference's emission order (that order is what makes the stable sort's output well defined).
//
// The reference lets one thread emit all tiles of its surfel serially.  Here a warp flattens the
// pair lists of its 32 surfels: lane j of each round owns flattened pair j, finds the owning surfel
// with a 5-step shuffle search over the warp's inclusive counts, and stores - consecutive lanes hit
// consecutive addresses, and a surfel covering thousands of tiles is spread over the whole warp.
#include "common.cuh"

#define DUP_THREADS 256

__global__ void __launch_bounds__(DUP_THREADS)
sb_duplicate_kernel(int P, uint32_t grid_x, const float* __restrict__ depths, const ushort4* __restrict__ rects,
                    const uint32_t* __restrict__ tiles_touched, const uint32_t* __restrict__ offsets,
                    uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const int lane = threadIdx.x & 31;
    const int warp_global = (blockIdx.x * DUP_THREADS + threadIdx.x) >> 5;
    const int first = warp_global * 32;
    if (first >= P) return;
    const int idx = first + lane;
    uint32_t cnt = 0, incl_g = 0, depth_bits = 0, rmin = 0, rw = 1;
    if (idx < P) {
        cnt = tiles_touched[idx];
        incl_g = offsets[idx];
        if (cnt) {
            depth_bits = __float_as_uint(depths[idx]);
            ushort4 r = rects[idx];
            rmin = (uint32_t)r.x | ((uint32_t)r.y << 16);
            rw = (uint32_t)r.z - (uint32_t)r.x;
        }
    }
    // global start of this warp's pair segment = exclusive offset of its first surfel
    const uint32_t seg_begin = __shfl_sync(0xffffffffu, incl_g - cnt, 0);
    uint32_t incl = incl_g - seg_begin;  // inclusive count within the warp
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31 < (P - first - 1) ? 31 : (P - first - 1));
    if (idx >= P) incl = total;  // keep the search array monotone past the end of the cloud
    for (uint32_t j = lane; j < ((total + 31) & ~31u); j += 32) {
        // smallest lane s with incl[s] > j
        int s = 0;
#pragma unroll
        for (int step = 16; step > 0; step >>= 1) {
            uint32_t probe = __shfl_sync(0xffffffffu, incl, s + step - 1);
            if (probe <= j) s += step;
        }
        uint32_t s_incl = __shfl_sync(0xffffffffu, incl, s);
        uint32_t s_cnt = __shfl_sync(0xffffffffu, cnt, s);
        uint32_t s_depth = __shfl_sync(0xffffffffu, depth_bits, s);
        uint32_t s_rmin = __shfl_sync(0xffffffffu, rmin, s);
        uint32_t s_rw = __shfl_sync(0xffffffffu, rw, s);
        if (j < total) {
            uint32_t t = j - (s_incl - s_cnt);
            uint32_t ty = t / s_rw, tx = t - ty * s_rw;
            uint32_t x = (s_rmin & 0xffffu) + tx, y = (s_rmin >> 16) + ty;
            uint64_t key = (uint64_t)(y * grid_x + x);
            key <<= 32;
            key |= s_depth;
            keys[seg_begin + j] = key;
            vals[seg_begin + j] = (uint32_t)(first + s);
        }
    }
}

__global__ void sb_tile_ranges_kernel(int64_t R, const uint64_t* __restrict__ keys, uint2* __restrict__ ranges) {
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= R) return;
    uint32_t cur = (uint32_t)(keys[idx] >> 32);
    if (idx == 0) {
        ranges[cur].x = 0;
    } else {
        uint32_t prev = (uint32_t)(keys[idx - 1] >> 32);
        if (cur != prev) {
            ranges[prev].y = (uint32_t)idx;
            ranges[cur].x = (uint32_t)idx;
        }
    }
    if (idx == R - 1) ranges[cur].y = (uint32_t)R;
}

int sb_launch_duplicate(int P, int W, int H, const char* geom, uint64_t* keys, uint32_t* vals, cudaStream_t s) {
    if (P == 0) return 0;
    SbGeomLayout L = sb_geom_layout(P);
    uint32_t grid_x = (W + SB_TILE_X - 1) / SB_TILE_X;
    int nwarps = (P + 31) / 32;
    int nblocks = (nwarps + DUP_THREADS / 32 - 1) / (DUP_THREADS / 32);
    sb_duplicate_kernel<<<nblocks, DUP_THREADS, 0, s>>>(P, grid_x, (const float*)(geom + L.depth),
                                                        (const ushort4*)(geom + L.rect),
                                                        (const uint32_t*)(geom + L.tiles),
                                                        (const uint32_t*)(geom + L.offsets), keys, vals);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}

// Longest-list-first launch order of the tiles for the two blend kernels: a CTA's run time is proportional to its
// tile's list length, lists differ by 100x across an image, and blockIdx order is dispatch order, so the long
// tiles start first and the short ones fill the tail (LPT scheduling).  One CTA: max, 1024-bin counting sort by
// scaled length (order inside a bin is arbitrary), scatter.
#define ORD_THREADS 1024
__global__ void __launch_bounds__(ORD_THREADS) sb_tile_order_kernel(int tiles, const uint2* __restrict__ ranges,
                                                                     uint32_t* __restrict__ order) {
    __shared__ uint32_t s_bin[ORD_THREADS];
    __shared__ uint32_t s_max;
    const int tid = threadIdx.x;
    if (tid == 0) s_max = 0;
    s_bin[tid] = 0;
    __syncthreads();
    uint32_t mx = 0;
    for (int t = tid; t < tiles; t += ORD_THREADS) mx = max(mx, ranges[t].y - ranges[t].x);
    mx = __reduce_max_sync(0xffffffffu, mx);
    if ((tid & 31) == 0 && mx) atomicMax(&s_max, mx);
    __syncthreads();
    const uint64_t top = s_max ? s_max : 1;
    // bin 0 = longest
    for (int t = tid; t < tiles; t += ORD_THREADS) {
        const uint32_t n = ranges[t].y - ranges[t].x;
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
        const uint32_t n = ranges[t].y - ranges[t].x;
        const uint32_t pos = atomicAdd(&s_bin[(ORD_THREADS - 1) - (uint32_t)(((uint64_t)n * (ORD_THREADS - 1)) / top)], 1u);
        order[pos] = (uint32_t)t;
    }
}

int sb_launch_tile_ranges(int64_t R, int tiles, const uint64_t* keys_sorted, uint2* ranges, uint32_t* order,
                          cudaStream_t s) {
    SB_CUDA_CHECK(cudaMemsetAsync(ranges, 0, (size_t)tiles * sizeof(uint2), s));
    if (R > 0) {
        int64_t nblocks = (R + 255) / 256;
        sb_tile_ranges_kernel<<<(unsigned)nblocks, 256, 0, s>>>(R, keys_sorted, ranges);
        sb_count_launch(1);
    }
    sb_tile_order_kernel<<<1, ORD_THREADS, 0, s>>>(tiles, ranges, order);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}
