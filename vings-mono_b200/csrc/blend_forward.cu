// K3: per-tile front-to-back alpha blending of colour, depth, alpha, normal, median depth, distortion.
//
// Replaces renderCUDA (DSR/cuda_rasterizer/forward.cu:254-504).  One CTA per 16x16 tile (the tile
// size is part of the key contract); a warp owns an 8x4-pixel block of it, one thread per pixel.
// The kernel is instruction-issue bound, not HBM bound (profiles/): ~90 % of the (pixel, surfel)
// pairs a tile list implies contribute nothing, so the design goal is to reject those pairs with as
// few issued instructions as possible while running the reference's exact arithmetic on the rest:
//   1. the sorted surfel *records* (128 B = one line, packed once by the preprocess kernel) are gathered by 16-byte
//      cp.async (four whole records per warp instruction) into an 8-stage shared-memory ring of 32-surfel batches
//      handed over by mbarriers (ring.cuh; the TMA bulk-copy gather it replaced is kept behind SB_RING_LDGSTS=0), so
//      the gathers run up to 8 batches ahead of the blend and the warps of a CTA never meet at a barrier inside the
//      walk (the reference does 4 scattered array reads plus a per-pixel global colour read per pair, behind two CTA
//      barriers per 256 surfels);
//   2. warp-level cull: each lane tests ONE surfel of the batch against the warp's 8x4 block -- its conservative
//      pixel box, then an interval bound of the lane-level test over the whole block (sb_block_far) -- a ballot
//      compacts the survivors, and only those are visited (1 test per 32 pairs);
//   3. lane-level cull: a division-free test (6 FFMA for the affine ray-splat cross product + a conic
//      compare) rejects pixels outside the surfel's alpha >= 1/255 footprint (surfel_math.cuh); the OR of the
//      lanes' pass bits is stored as the (batch, warp) hit mask the backward walks;
//   4. pairs that pass run the reference's expression order, so the gates (alpha < 1/255,
//      T < 1e-4, T > 0.5, rho3d <= rho2d, depth < near) flip on the same pairs;
//   5. no per-32-splat checkpoints are written (the backward re-walks the tile instead), removing the
//      reference's 13 KiB-per-bucket SampleState traffic; termination is tracked per warp and per CTA.
#include "pair_math.cuh"
#include "ring.cuh"

#define FB_THREADS 256
#define FB_WARPS 8
#define FB_BATCH 32
#define FB_STAGES 8
#ifndef FB_POLL_NS
#define FB_POLL_NS 200   // measured (C5 forward -3 %, C2 neutral): see the wait loop
#endif
// Warp-level interval cull (pair_math.cuh::sb_block_far) behind the box test: the lane that owns a surfel of the batch
// also bounds the lane-level test over the warp's whole 8x4 block.  On the CPU model of the walk (tools/
// forward_walk_model.py, profiles/r02_forward_walk_model.txt) it removes 69 % (C2), 29 % (C3), 94 % (C4) and 96 % (C5)
// of the box-passing (warp, surfel) visits -- all of which phase 1 would reject pixel by pixel: on the 1080p-class
// configs half of all pairs belong to < 1 % of the surfels, the ones that cross the camera plane (unbounded boxes,
// every tile).  Measured (profiles/r02_ab_variants.txt, block r02m): C2 0.574 -> 0.489 ms, C4 5.48 -> 2.61, C5 7.66 -> 3.45.
#ifndef FB_BLOCK_FAR
#define FB_BLOCK_FAR 1
#endif
// float4s per ring slot: the 8 of the record + 1 of padding, so that the per-lane record reads of the warp-level cull
// (lane = surfel: row stride 144 B) spread over all bank groups instead of hitting one 32 ways.  With the 128-byte
// pitch the same cull measures 0.542 (C2) and 6.71 ms (C5): the conflicts cost more than the cull saves.
#ifndef FB_REC_PITCH
#define FB_REC_PITCH 9
#endif

// (a min-CTAs bound of 4/5/6 -> 56/48/40 registers was measured 2-5 % slower than the unconstrained 58)
__global__ void __launch_bounds__(FB_THREADS)
sb_blend_forward_kernel(int W, int H, int grid_x, const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order,
                        const uint32_t* __restrict__ point_list, const float4* __restrict__ rec,
                        const float* __restrict__ bg, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                        float* __restrict__ out_color, float* __restrict__ out_others, uint32_t* __restrict__ hit_masks) {
    __shared__ __align__(128) float4 s_rec[FB_STAGES][FB_BATCH][FB_REC_PITCH];
    __shared__ __align__(8) uint64_t s_full[FB_STAGES];
    __shared__ uint32_t s_cnt[FB_STAGES];
    __shared__ uint32_t s_donemask;  // warps whose 32 pixels have all terminated
    __shared__ int s_limit;          // first batch that is NOT loaded (set once every warp is done)

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
#ifdef SB_DEBUG_NO_LPT
    const uint32_t tile = blockIdx.x;
#else
    const uint32_t tile = tile_order[blockIdx.x];  // longest lists first (binning.cu)
#endif
    const uint32_t tx = tile % grid_x, ty = tile / grid_x;
    // warp w owns the 8x4 block at (w&1, w>>1); lane l the pixel (l&7, l>>3) of it
    const uint32_t wx0 = tx * SB_TILE_X + (wid & 1) * 8, wy0 = ty * SB_TILE_Y + (wid >> 1) * 4;
    const uint32_t px = wx0 + (lane & 7), py = wy0 + (lane >> 3);
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const uint32_t pix_id = (uint32_t)W * py + px;
    const float2 pixf = {(float)px, (float)py};
    const float bx0 = (float)wx0, bx1 = (float)(wx0 + 7), by0 = (float)wy0, by1 = (float)(wy0 + 3);
    const float bcx = (float)wx0 + 3.5f, bcy = (float)wy0 + 1.5f;   // block centre (sb_block_far)
    bool done = !inside;

    const uint2 range = ranges[tile];
    const int n = (int)(range.y - range.x);
    const int nbatches = (n + FB_BATCH - 1) / FB_BATCH;
    // hit-mask words of this tile's batches (sb_hit_slot, common.cuh): one word per (batch, warp) = the surfels of the
    // batch some live pixel of the warp's block may blend.  The backward walks exactly those (blend_backward.cu).
    uint32_t* my_masks = hit_masks + ((size_t)sb_hit_slot(range.x, tile) * FB_WARPS + wid);

    if (tid < FB_STAGES) {
        bb_mbar_init(&s_full[tid], BB_RING_ARRIVALS);
        s_cnt[tid] = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid == 0) {
        s_donemask = 0;
        s_limit = nbatches;
    }
    __syncthreads();
    // gather the records of batch b into ring stage b % FB_STAGES (called by one whole warp)
    auto fill = [&](int b) {
        const int st = b % FB_STAGES;
        const int cnt = min(FB_BATCH, n - b * FB_BATCH);
        uint32_t id = 0;
        if (lane < cnt) id = __ldg(point_list + range.x + b * FB_BATCH + lane);
        bb_gather<FB_REC_PITCH>(s_rec[st], rec, id, cnt < FB_BATCH ? (1u << cnt) - 1u : 0xffffffffu, &s_full[st], lane);
    };
    for (int b0 = wid; b0 < FB_STAGES && b0 < nbatches; b0 += FB_WARPS) fill(b0);  // prime the ring

    float T = 1.0f;
    uint32_t last_contributor = 0;
    float C[3] = {0.f, 0.f, 0.f};
    float N[3] = {0.f, 0.f, 0.f};
    float D = 0.f, M1 = 0.f, M2 = 0.f, distortion = 0.f, median_depth = 0.f;
    float median_contributor = -1.f;

    bool wdone = __all_sync(0xffffffffu, done);
    if (wdone && lane == 0) atomicOr(&s_donemask, 1u << wid);
    for (int b = 0; b < nbatches; b++) {
        const int st = b % FB_STAGES;
        // wait for the batch -- or learn that every pixel of the tile has terminated (forward.cu:348-350: the
        // reference votes with a CTA barrier per 256 surfels; here the last warp to finish a batch stops the
        // refills once all eight warps are done, and nobody waits for a batch that will not come)
        bool stop = false;
        while (!bb_mbar_try(&s_full[st], (uint32_t)(b / FB_STAGES) & 1u)) {
            if (b >= *(volatile int*)&s_limit) { stop = true; break; }
#if FB_POLL_NS > 0
            __nanosleep(FB_POLL_NS);   // a warp that is ahead of its CTA should not spend issue slots on probing
#endif
        }
        if (stop) break;
        const int cnt = min(FB_BATCH, n - b * FB_BATCH);
        const float4(*sq)[FB_REC_PITCH] = s_rec[st];
        if (!wdone) {
            const int c = 0;
            // ---- warp-level cull: lane tests surfel c+lane's box against the warp's pixel block ----
            // (measured and dropped, round 2, tools/time_stages.py A/B on C2-C5: shrinking the block to the bounding box of
            // its still-live pixels is neutral; a four-corner "footprint contains the whole block" test that skips phase 1
            // costs 12-14 % at C4/C5 -- the synthetic clouds' randomly oriented surfels project to thin ellipses that
            // rarely contain an 8x4 block; bounds along the two diagonals as well (x+y, x-y: conservative, 16 % fewer
            // box-passing blocks on CPU) cost 2-3 % at C4/C5 -- the eight int16 compares per surfel outweigh them)
            bool ov = false;
            if (c + lane < cnt) {
                const float4 bb = sq[c + lane][SB_REC_BBOX];
                ov = bb.x <= bx1 && bb.y >= bx0 && bb.z <= by1 && bb.w >= by0;
#if FB_BLOCK_FAR
                if (ov) ov = !sb_block_far(sq[c + lane][0], sq[c + lane][1], sq[c + lane][2], bcx, bcy);
#endif
#ifdef SB_BLOCK_CULL  // opt-in: measured neutral
                if (ov) ov = !sb_block_culled(sq[c + lane][0], sq[c + lane][1], sq[c + lane][2], bx0, bx1, by0, by1);
#endif
            }
            unsigned m = __ballot_sync(0xffffffffu, ov);
            // ---- phase 1: lane-level cull (division free) of every surviving surfel -> one pass bit per
            // (lane, surfel); two surfels per trip and no vote / branch on the arithmetic, so the LDS -> FFMA
            // chains overlap ----
            uint32_t pm = 0;
            while (m) {
                const int j0 = c + __ffs(m) - 1;
                m &= m - 1;
                const int j1 = m ? c + __ffs(m) - 1 : j0;
                m &= m - 1;
                const float4 a0 = sq[j0][0], a1 = sq[j0][1], a2 = sq[j0][2];
                const float4 e0 = sq[j1][0], e1 = sq[j1][1], e2 = sq[j1][2];
                const float dx0 = a0.w - pixf.x, dy0 = a1.w - pixf.y, dx1 = e0.w - pixf.x, dy1 = e1.w - pixf.y;
                const float r0 = dx0 * dx0 + dy0 * dy0, r1 = dx1 * dx1 + dy1 * dy1;
                const float qx0 = a2.x + a0.x * dx0 + a1.x * dy0, qx1 = e2.x + e0.x * dx1 + e1.x * dy1;
                const float qy0 = a2.y + a0.y * dx0 + a1.y * dy0, qy1 = e2.y + e0.y * dx1 + e1.y * dy1;
                const float qz0 = a2.z + a0.z * dx0 + a1.z * dy0, qz1 = e2.z + e0.z * dx1 + e1.z * dy1;
                const float E0 = (qx0 * qx0 + qy0 * qy0) - a2.w * (qz0 * qz0);
                const float E1 = (qx1 * qx1 + qy1 * qy1) - e2.w * (qz1 * qz1);
                const bool p0 = !(E0 > 0.0f && (r0 + r0) > a2.w);
                const bool p1 = !(E1 > 0.0f && (r1 + r1) > e2.w);
                pm |= ((uint32_t)p0 << (j0 - c)) | ((uint32_t)p1 << (j1 - c));
            }
            if (done) pm = 0;
            // ---- phase 2: surfels some live lane of the warp passes ----
            uint32_t am = __reduce_or_sync(0xffffffffu, pm);
#ifndef FB_NO_MASK_STORE   // (timing experiments only: the backward needs the masks)
            if (lane == 0) my_masks[(size_t)b * FB_WARPS] = am;
#endif
            while (am) {
                const int jj = __ffs(am) - 1;
                const int j = c + jj;
                am &= am - 1;
                if (done || !((pm >> jj) & 1u)) continue;
                const float2 d = {sq[j][0].w - pixf.x, sq[j][1].w - pixf.y};  // == the reference's d (forward.cu:413)
                // ---- the reference's arithmetic (forward.cu:401-468) ----
                const float4 q3 = sq[j][SB_REC_TU], q4 = sq[j][SB_REC_TV], q5 = sq[j][SB_REC_TW];
                const float3 Tu = {q3.x, q3.y, q3.z}, Tv = {q4.x, q4.y, q4.z}, Tw = {q5.x, q5.y, q5.z};
                const SbPair g = sb_pair(pixf, Tu, Tv, Tw);
                if (g.p.z == 0.0f) continue;
                const float rho2d = sb_rho2d_forward(d);
                const float rho = fminf(g.rho3d, rho2d);
                const float depth = sb_pair_depth(g, rho2d, Tw.z);
                if (depth < SB_NEAR) continue;
                const float power = __fmul_rn(rho, -0.5f);
                if (power > 0.0f) continue;
                const float alpha = fminf(0.99f, __fmul_rn(q3.w, expf(power)));
                if (alpha < 1.0f / 255.0f) continue;
                const float test_T = __fmul_rn(T, __fadd_rn(-alpha, 1.0f));
                if (test_T < 0.0001f) {
                    done = true;
                    continue;
                }
                const float w = __fmul_rn(T, alpha);
                const float4 q6 = sq[j][SB_REC_N];
                const uint32_t contributor = (uint32_t)(b * FB_BATCH + j + 1);
                // distortion, depth, moments (forward.cu:444-449), in the oracle build's operation order
                const float A = __fadd_rn(-T, 1.0f);
                const float mm = sb_ndc_depth(depth);
                const float m2 = __fmul_rn(mm, mm);
                const float e = __fmaf_rn(-M1, __fadd_rn(mm, mm), __fmaf_rn(A, m2, M2));
                distortion = __fmaf_rn(w, e, distortion);
                D = __fmaf_rn(depth, w, D);
                M1 = __fmaf_rn(w, mm, M1);
                M2 = __fmaf_rn(w, m2, M2);
                if (T > 0.5f) {
                    median_depth = depth;
                    median_contributor = contributor;
                }
                N[0] = __fmaf_rn(q6.x, w, N[0]);
                N[1] = __fmaf_rn(q6.y, w, N[1]);
                N[2] = __fmaf_rn(q6.z, w, N[2]);
                C[0] = __fmaf_rn(q4.w, w, C[0]);
                C[1] = __fmaf_rn(q5.w, w, C[1]);
                C[2] = __fmaf_rn(q6.w, w, C[2]);
                T = test_T;
                last_contributor = contributor;
            }
            wdone = __all_sync(0xffffffffu, done);
            if (wdone && lane == 0) atomicOr(&s_donemask, 1u << wid);
        }
        // hand the stage back: the last of the 8 warps to get here refills it with the batch FB_STAGES ahead
        __syncwarp();
        uint32_t old = 0;
        if (lane == 0) {
            __threadfence_block();
            old = atomicAdd(&s_cnt[st], 1u);
        }
        old = __shfl_sync(0xffffffffu, old, 0);
        if (old == FB_WARPS - 1) {
            if (lane == 0) s_cnt[st] = 0;
            __threadfence_block();
            if (b + FB_STAGES < nbatches) {
                if (*(volatile uint32_t*)&s_donemask == (1u << FB_WARPS) - 1u) {
                    if (lane == 0) atomicMin(&s_limit, b + FB_STAGES);
                } else {
                    fill(b + FB_STAGES);
                }
            }
        }
    }

    bb_ring_drain();
    if (inside) {
        const size_t HW = (size_t)H * W;
        final_T[pix_id] = T;
        final_T[pix_id + HW] = M1;
        final_T[pix_id + 2 * HW] = M2;
        n_contrib[pix_id] = last_contributor;
        n_contrib[pix_id + HW] = (uint32_t)median_contributor;
        out_color[pix_id] = C[0] + T * bg[0];
        out_color[pix_id + HW] = C[1] + T * bg[1];
        out_color[pix_id + 2 * HW] = C[2] + T * bg[2];
        out_others[pix_id] = D;
        out_others[pix_id + HW] = 1 - T;
        out_others[pix_id + 2 * HW] = N[0];
        out_others[pix_id + 3 * HW] = N[1];
        out_others[pix_id + 4 * HW] = N[2];
        out_others[pix_id + 5 * HW] = median_depth;
        out_others[pix_id + 6 * HW] = distortion;
    }
}

int sb_launch_blend_forward(int W, int H, const float* bg, const char* geom, const uint32_t* point_list,
                            char* image, float* out_color, float* out_others, uint32_t* hit_masks, cudaStream_t s) {
    SbImageLayout IL = sb_image_layout(W, H);
    const int grid_x = (W + SB_TILE_X - 1) / SB_TILE_X, grid_y = (H + SB_TILE_Y - 1) / SB_TILE_Y;
    // geom layout: the record array is first and its offset does not depend on P
    sb_blend_forward_kernel<<<grid_x * grid_y, FB_THREADS, 0, s>>>(
        W, H, grid_x, (const uint2*)(image + IL.ranges), (const uint32_t*)(image + IL.order), point_list, (const float4*)geom, bg,
        (float*)(image + IL.final_T), (uint32_t*)(image + IL.n_contrib), out_color, out_others, hit_masks);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}
