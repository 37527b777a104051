// K1: per-surfel forward preprocess fused with the tiles_touched prefix sum, plus mark_visible.
//
// Replaces DSR/cuda_rasterizer/forward.cu:160-249 (preprocessCUDA), the cub::DeviceScan call at
// rasterizer_impl.cu:387 and checkFrustum (rasterizer_impl.cu:106-118).  One thread per surfel,
// 256 surfels per CTA.  The [P,3] inputs are staged through shared memory with 128-bit coalesced
// loads; rotations are natural float4s.  Each CTA publishes its tile-count aggregate and resolves
// its exclusive prefix by decoupled look-back, so offsets are written by the same kernel.
#include "surfel_math.cuh"
#include "surfel_math_pinned.cuh"

#define PRE_THREADS 256

// look-back status word: [63:62] flag (1 = aggregate, 2 = inclusive prefix), [31:0] value
#define SCAN_FLAG_AGG (1ull << 62)
#define SCAN_FLAG_INC (2ull << 62)

// Cooperative 128-bit load of `n3` consecutive floats starting at src (16-B aligned) into smem.
__device__ __forceinline__ void sb_stage_f3(float* dst, const float* __restrict__ src, int nfloats, bool aligned) {
    if (aligned) {
        int nvec = nfloats >> 2;
        const float4* s4 = reinterpret_cast<const float4*>(src);
        float4* d4 = reinterpret_cast<float4*>(dst);
        for (int i = threadIdx.x; i < nvec; i += PRE_THREADS) d4[i] = __ldg(s4 + i);
        for (int i = (nvec << 2) + threadIdx.x; i < nfloats; i += PRE_THREADS) dst[i] = __ldg(src + i);
    } else {
        for (int i = threadIdx.x; i < nfloats; i += PRE_THREADS) dst[i] = __ldg(src + i);
    }
}

__global__ void __launch_bounds__(PRE_THREADS)
sb_preprocess_kernel(int P, int W, int H, const float* __restrict__ means3D, const float2* __restrict__ scales,
                     const float4* __restrict__ rotations, const float* __restrict__ opacities,
                     const float* __restrict__ colors, const float* __restrict__ viewmatrix,
                     const float* __restrict__ projmatrix, float4* __restrict__ rec, float* __restrict__ depths,
                     ushort4* __restrict__ rects, uint32_t* __restrict__ tiles_touched,
                     uint32_t* __restrict__ offsets, unsigned long long* __restrict__ scan_state,
                     int* __restrict__ radii, int nblocks) {
    __shared__ __align__(16) float s_mean[PRE_THREADS * 3];
    __shared__ __align__(16) float s_col[PRE_THREADS * 3];
    __shared__ float s_vm[16], s_pm[16];
    __shared__ uint32_t s_warp[PRE_THREADS / 32];
    __shared__ uint32_t s_blk, s_prefix;

    // ticket = logical block index; guarantees every predecessor has started (look-back progress)
    unsigned long long* ticket = scan_state + nblocks;
    if (threadIdx.x == 0) s_blk = (uint32_t)atomicAdd(ticket, 1ull);
    if (threadIdx.x < 16) s_vm[threadIdx.x] = viewmatrix[threadIdx.x];
    else if (threadIdx.x < 32) s_pm[threadIdx.x - 16] = projmatrix[threadIdx.x - 16];
    __syncthreads();
    const uint32_t blk = s_blk;
    const int base = blk * PRE_THREADS;
    const int n_here = min(PRE_THREADS, P - base);
    const bool al = ((((uintptr_t)(means3D + (size_t)base * 3)) | ((uintptr_t)(colors + (size_t)base * 3))) & 15) == 0;
    sb_stage_f3(s_mean, means3D + (size_t)base * 3, n_here * 3, al);
    sb_stage_f3(s_col, colors + (size_t)base * 3, n_here * 3, al);
    __syncthreads();

    const int idx = base + threadIdx.x;
    uint32_t touched = 0;
    if (idx < P) {
        int radius_i = 0;
        float3 p_orig = {s_mean[3 * threadIdx.x], s_mean[3 * threadIdx.x + 1], s_mean[3 * threadIdx.x + 2]};
        // near cull only (auxiliary.h:197-222); p_hom/p_proj of the reference are dead code
        // operation order of the oracle build (FMUL on the m6*y term, auxiliary.h:91-99 as nvcc contracts it)
        float pz = __fadd_rn(__fmaf_rn(p_orig.z, s_vm[10], __fmaf_rn(p_orig.x, s_vm[2], __fmul_rn(p_orig.y, s_vm[6]))),
                             s_vm[14]);
        if (!(pz <= 0.2f)) {
            float Tm[9];
            float3 normal;
            const float4 rot = __ldg(rotations + idx);
            const float2 scl = __ldg(scales + idx);
            sb_transmat_pinned(rot.x, rot.y, rot.z, rot.w, scl.x, scl.y, p_orig.x, p_orig.y, p_orig.z, s_pm, s_vm, W, H,
                               Tm, normal);
            float2 point_image;
            float radius;
            if (sb_aabb_pinned(Tm, point_image, radius)) {
                uint2 rmin, rmax;
                uint2 grid = {(unsigned)((W + SB_TILE_X - 1) / SB_TILE_X), (unsigned)((H + SB_TILE_Y - 1) / SB_TILE_Y)};
                sb_rect_pinned(point_image, (int)radius, grid, rmin, rmax);
                uint32_t area = (rmax.x - rmin.x) * (rmax.y - rmin.y);
                if (area != 0) {
                    touched = area;
                    radius_i = (int)radius;
                    depths[idx] = pz;
                    rects[idx] = make_ushort4((unsigned short)rmin.x, (unsigned short)rmin.y,
                                              (unsigned short)rmax.x, (unsigned short)rmax.y);
                    const float opa = __ldg(opacities + idx);
                    const float3 Tu = {Tm[0], Tm[1], Tm[2]};
                    const float3 Tv = {Tm[3], Tm[4], Tm[5]};
                    const float3 Tw = {Tm[6], Tm[7], Tm[8]};
                    SbCull cull = sb_cull_data(Tu, Tv, Tw, point_image, opa);
                    float4* r = rec + (size_t)idx * SB_REC_F4;
                    r[0] = make_float4(cull.A.x, cull.A.y, cull.A.z, point_image.x);
                    r[1] = make_float4(cull.B.x, cull.B.y, cull.B.z, point_image.y);
                    r[2] = make_float4(cull.C.x, cull.C.y, cull.C.z, cull.tau);
                    r[3] = make_float4(Tu.x, Tu.y, Tu.z, opa);
                    r[4] = make_float4(Tv.x, Tv.y, Tv.z, s_col[3 * threadIdx.x]);
                    r[5] = make_float4(Tw.x, Tw.y, Tw.z, s_col[3 * threadIdx.x + 1]);
                    r[6] = make_float4(normal.x, normal.y, normal.z, s_col[3 * threadIdx.x + 2]);
                    r[7] = cull.bbox;
                }
            }
        }
        radii[idx] = radius_i;
        tiles_touched[idx] = touched;
    }

    // ---- block inclusive scan of `touched` ----
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t incl = touched;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) s_warp[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        uint32_t w = (lane < PRE_THREADS / 32) ? s_warp[lane] : 0;
#pragma unroll
        for (int o = 1; o < PRE_THREADS / 32; o <<= 1) {
            uint32_t v = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += v;
        }
        if (lane < PRE_THREADS / 32) s_warp[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const uint32_t block_total = s_warp[PRE_THREADS / 32 - 1];
    const uint32_t warp_excl = wid ? s_warp[wid - 1] : 0;

    // ---- decoupled look-back across CTAs (one warp) ----
    if (wid == 0) {
        volatile unsigned long long* st = scan_state;
        if (lane == 0) {
            unsigned long long word = (blk == 0 ? SCAN_FLAG_INC : SCAN_FLAG_AGG) | block_total;
            st[blk] = word;
        }
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
                unsigned inc_mask = __ballot_sync(0xffffffffu, (w >> 62) == 2);
                uint32_t val = (uint32_t)w;
                if (inc_mask) {
                    int first = __ffs(inc_mask) - 1;  // nearest predecessor holding an inclusive prefix
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
        if (lane == 0) {
            s_prefix = prefix;
            if (blk == (uint32_t)nblocks - 1) scan_state[nblocks + 1] = (unsigned long long)prefix + block_total;  // R
        }
    }
    __syncthreads();
    if (idx < P) offsets[idx] = s_prefix + warp_excl + incl;
}

__global__ void sb_mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ vm,
                                       bool* __restrict__ present) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    float3 p = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
    float pz = vm[2] * p.x + vm[6] * p.y + vm[10] * p.z + vm[14];
    present[idx] = !(pz <= 0.2f);
}

int sb_launch_preprocess(int P, int W, int H, const float* means3D, const float* scales, const float* rotations,
                         const float* opacities, const float* colors, const float* viewmatrix,
                         const float* projmatrix, char* geom, int* radii, cudaStream_t s) {
    SbGeomLayout L = sb_geom_layout(P);
    int nblocks = (P + PRE_THREADS - 1) / PRE_THREADS;
    SB_CUDA_CHECK(cudaMemsetAsync(geom + L.scan_state, 0, (size_t)(nblocks + 2) * 8, s));
    if (P == 0) return 0;
    sb_preprocess_kernel<<<nblocks, PRE_THREADS, 0, s>>>(
        P, W, H, means3D, (const float2*)scales, (const float4*)rotations, opacities, colors, viewmatrix, projmatrix,
        (float4*)(geom + L.rec), (float*)(geom + L.depth), (ushort4*)(geom + L.rect), (uint32_t*)(geom + L.tiles),
        (uint32_t*)(geom + L.offsets), (unsigned long long*)(geom + L.scan_state), radii, nblocks);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}

int sb_launch_mark_visible(int P, const float* means3D, const float* viewmatrix, bool* present, cudaStream_t s) {
    if (P == 0) return 0;
    sb_mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, viewmatrix, present);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}
