// K1: per-surfel forward preprocess, the depth-ordered surfel list, plus mark_visible.
//
// Replaces DSR/cuda_rasterizer/forward.cu:160-249 (preprocessCUDA) and checkFrustum (rasterizer_impl.cu:106-118).
// One thread per surfel, 256 surfels per CTA.  The [P,3] inputs are staged through shared memory with 128-bit
// coalesced loads; rotations are natural float4s.  Besides the render record the kernel writes the surfel's
// sort key (its depth word, SB_KEY_INVALID when it touches no tile) and adds its tile count to R; the
// reference's prefix sum over tiles_touched (cub::DeviceScan, rasterizer_impl.cu:387) moves into the duplicate
// kernel because pairs are emitted in depth order here (radix_sort.cu).
#include "surfel_math.cuh"
#include "surfel_math_pinned.cuh"

#define PRE_THREADS 256

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

#ifndef PRE_MIN_CTAS
#define PRE_MIN_CTAS 6
#endif
// PRECOMP: the caller supplies the 3x3 splat-to-pixel transform of every surfel (`transMat_precomp` of the
// reference, forward.cu:204-219) in place of scales + rotations; the normal is then (0,0,1).
template <bool PRECOMP>
__global__ void __launch_bounds__(PRE_THREADS, PRE_MIN_CTAS)
sb_preprocess_kernel(int P, int W, int H, const float* __restrict__ means3D, const float2* __restrict__ scales,
                     const float4* __restrict__ rotations, const float* __restrict__ transmat_precomp,
                     const float* __restrict__ opacities,
                     const float* __restrict__ colors, const float* __restrict__ viewmatrix,
                     const float* __restrict__ projmatrix, float4* __restrict__ rec, float* __restrict__ depths,
                     ushort4* __restrict__ rects, uint32_t* __restrict__ tiles_touched,
                     uint32_t* __restrict__ sort_keys, unsigned long long* __restrict__ total_pairs,
                     int* __restrict__ radii) {
    __shared__ __align__(16) float s_mean[PRE_THREADS * 3];
    __shared__ __align__(16) float s_col[PRE_THREADS * 3];
    __shared__ float s_vm[16], s_pm[16];
    __shared__ uint32_t s_warp[PRE_THREADS / 32];

    if (threadIdx.x < 16) s_vm[threadIdx.x] = viewmatrix[threadIdx.x];
    else if (threadIdx.x < 32) s_pm[threadIdx.x - 16] = projmatrix[threadIdx.x - 16];
    const int base = blockIdx.x * PRE_THREADS;
    const int n_here = min(PRE_THREADS, P - base);
    const bool al = ((((uintptr_t)(means3D + (size_t)base * 3)) | ((uintptr_t)(colors + (size_t)base * 3))) & 15) == 0;
    sb_stage_f3(s_mean, means3D + (size_t)base * 3, n_here * 3, al);
    sb_stage_f3(s_col, colors + (size_t)base * 3, n_here * 3, al);
    __syncthreads();

    const int idx = base + threadIdx.x;
    uint32_t touched = 0;
    if (idx < P) {
        int radius_i = 0;
        uint32_t sort_key = SB_KEY_INVALID;
        float3 p_orig = {s_mean[3 * threadIdx.x], s_mean[3 * threadIdx.x + 1], s_mean[3 * threadIdx.x + 2]};
        // near cull only (auxiliary.h:197-222); p_hom/p_proj of the reference are dead code
        // operation order of the oracle build (FMUL on the m6*y term, auxiliary.h:91-99 as nvcc contracts it)
        float pz = __fadd_rn(__fmaf_rn(p_orig.z, s_vm[10], __fmaf_rn(p_orig.x, s_vm[2], __fmul_rn(p_orig.y, s_vm[6]))),
                             s_vm[14]);
        if (!(pz <= 0.2f)) {
            float Tm[9];
            float3 normal;
            if (PRECOMP) {
#pragma unroll
                for (int k = 0; k < 9; k++) Tm[k] = __ldg(transmat_precomp + 9 * (size_t)idx + k);
                normal = make_float3(0.0f, 0.0f, 1.0f);
            } else {
                const float4 rot = __ldg(rotations + idx);
                const float2 scl = __ldg(scales + idx);
                sb_transmat_pinned(rot.x, rot.y, rot.z, rot.w, scl.x, scl.y, p_orig.x, p_orig.y, p_orig.z, s_pm, s_vm,
                                   W, H, Tm, normal);
            }
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
                    sort_key = __float_as_uint(pz);
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
        sort_keys[idx] = sort_key;
    }

    // ---- R += this CTA's pairs (one atomic per CTA) ----
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint32_t wsum = __reduce_add_sync(0xffffffffu, touched);
    if (lane == 0) s_warp[wid] = wsum;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
#pragma unroll
        for (int w = 0; w < PRE_THREADS / 32; w++) t += s_warp[w];
        if (t) atomicAdd(total_pairs, t);
    }
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
                         const float* transmat_precomp, const float* opacities, const float* colors,
                         const float* viewmatrix, const float* projmatrix, char* geom, int* radii, cudaStream_t s) {
    SbGeomLayout L = sb_geom_layout(P);
    int nblocks = (P + PRE_THREADS - 1) / PRE_THREADS;
    SB_CUDA_CHECK(cudaMemsetAsync(geom + L.state, 0, L.state_bytes, s));
    if (P == 0) return 0;
    auto kern = transmat_precomp ? sb_preprocess_kernel<true> : sb_preprocess_kernel<false>;
    kern<<<nblocks, PRE_THREADS, 0, s>>>(
        P, W, H, means3D, (const float2*)scales, (const float4*)rotations, transmat_precomp, opacities, colors,
        viewmatrix, projmatrix, (float4*)(geom + L.rec), (float*)(geom + L.depth), (ushort4*)(geom + L.rect),
        (uint32_t*)(geom + L.tiles), (uint32_t*)(geom + L.sort_k[1]), (unsigned long long*)(geom + L.counts), radii);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}

// Stable sort of the visible surfels by depth word: geom.sort_v[1][0..V) = their ids, V -> geom.counts[1].
int sb_launch_surfel_sort(int P, char* geom, cudaStream_t s) {
    if (P == 0) return 0;
    SbGeomLayout L = sb_geom_layout(P);
    const SbSortPlan pl = sb_sort_plan(SB_DEPTH_BITS);
    uint32_t* kbuf[2] = {(uint32_t*)(geom + L.sort_k[0]), (uint32_t*)(geom + L.sort_k[1])};
    uint32_t* vbuf[2] = {(uint32_t*)(geom + L.sort_v[0]), (uint32_t*)(geom + L.sort_v[1])};
    uint32_t* hist = (uint32_t*)(geom + L.hist);
    uint32_t* n_valid = (uint32_t*)(geom + L.counts + 8);
    int rc = sb_launch_sort_hist(kbuf[1], P, pl, 1, hist, n_valid, s);
    if (rc) return rc;
    return sb_launch_sort_passes(pl, 1, kbuf[1], nullptr, kbuf, vbuf, vbuf[1], P, n_valid, hist,
                                 (uint32_t*)(geom + L.tickets), (uint32_t*)(geom + L.status), s);
}

int sb_launch_mark_visible(int P, const float* means3D, const float* viewmatrix, bool* present, cudaStream_t s) {
    if (P == 0) return 0;
    sb_mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, viewmatrix, present);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}
