// K3: per-tile front-to-back alpha blending of colour, depth, alpha, normal, median depth, distortion.
//
// Replaces renderCUDA (DSR/cuda_rasterizer/forward.cu:254-504).  One CTA per 16x16 tile (the tile
// size is part of the key contract), one thread per pixel.  Differences from the reference:
//   * the sorted surfel *records* (80 B, packed once by the preprocess kernel) are gathered with
//     cp.async 16-B copies into a double-buffered shared-memory ring, so the gather of batch b+1
//     overlaps the blend of batch b; the reference does 4 scattered array reads + a per-pixel global
//     colour read per (pixel, surfel);
//   * no per-32-splat checkpoints are written (the backward re-walks the tile front to back instead),
//     removing the reference's 13 KiB-per-bucket SampleState traffic entirely;
//   * termination is tracked per warp (ballot) as well as per CTA.
// The per-(pixel, surfel) arithmetic keeps the reference's expression order so thresholds
// (alpha < 1/255, T < 1e-4, T > 0.5, rho3d <= rho2d, depth < near) flip on the same pairs.
#include "common.cuh"
#include <cuda_pipeline.h>

#define FB_THREADS 256
#define FB_BATCH 128

__global__ void __launch_bounds__(FB_THREADS)
sb_blend_forward_kernel(int W, int H, int grid_x, const uint2* __restrict__ ranges,
                        const uint32_t* __restrict__ point_list, const float4* __restrict__ rec,
                        const float* __restrict__ bg, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                        float* __restrict__ out_color, float* __restrict__ out_others) {
    __shared__ __align__(16) float4 s_rec[2][FB_BATCH][SB_REC_F4];

    const int tid = threadIdx.x;
    const uint32_t tile = blockIdx.x;
    const uint32_t tx = tile % grid_x, ty = tile / grid_x;
    const uint32_t px = tx * SB_TILE_X + (tid & 15), py = ty * SB_TILE_Y + (tid >> 4);
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const uint32_t pix_id = (uint32_t)W * py + px;
    const float2 pixf = {(float)px, (float)py};
    bool done = !inside;

    const uint2 range = ranges[tile];
    const int n = (int)(range.y - range.x);
    const int nbatches = (n + FB_BATCH - 1) / FB_BATCH;

    auto issue = [&](int b) {
        const int cnt = min(FB_BATCH, n - b * FB_BATCH);
        const uint32_t* ids = point_list + range.x + b * FB_BATCH;
        for (int e = tid; e < cnt * SB_REC_F4; e += FB_THREADS) {
            int sp = e / SB_REC_F4, q = e - sp * SB_REC_F4;
            uint32_t id = __ldg(ids + sp);
            __pipeline_memcpy_async(&s_rec[b & 1][sp][q], rec + (size_t)id * SB_REC_F4 + q, 16);
        }
        __pipeline_commit();
    };

    float T = 1.0f;
    uint32_t contributor = 0, last_contributor = 0;
    float C[3] = {0.f, 0.f, 0.f};
    float N[3] = {0.f, 0.f, 0.f};
    float D = 0.f, M1 = 0.f, M2 = 0.f, distortion = 0.f, median_depth = 0.f;
    float median_contributor = -1.f;

    if (nbatches > 0) issue(0);
    for (int b = 0; b < nbatches; b++) {
        if (b + 1 < nbatches) {
            issue(b + 1);
            __pipeline_wait_prior(1);
        } else {
            __pipeline_wait_prior(0);
        }
        // barrier doubles as the CTA-wide termination vote (forward.cu:348-350)
        if (__syncthreads_count(done) == FB_THREADS) break;
        const int cnt = min(FB_BATCH, n - b * FB_BATCH);
        if (!__all_sync(0xffffffffu, done)) {
            const float4(*sr)[SB_REC_F4] = s_rec[b & 1];
            for (int j = 0; j < cnt && !done; j++) {
                contributor++;
                const float4 q0 = sr[j][0], q1 = sr[j][1], q2 = sr[j][2];
                const float3 Tu = {q0.x, q0.y, q0.z}, Tv = {q1.x, q1.y, q1.z}, Tw = {q2.x, q2.y, q2.z};
                const float2 xy = {q0.w, q1.w};
                // two homogeneous planes through the pixel, intersected with the splat (forward.cu:405-412)
                float3 k = {pixf.x * Tw.x - Tu.x, pixf.x * Tw.y - Tu.y, pixf.x * Tw.z - Tu.z};
                float3 l = {pixf.y * Tw.x - Tv.x, pixf.y * Tw.y - Tv.y, pixf.y * Tw.z - Tv.z};
                float3 p = {k.y * l.z - k.z * l.y, k.z * l.x - k.x * l.z, k.x * l.y - k.y * l.x};
                if (p.z == 0.0f) continue;
                float2 s = {p.x / p.z, p.y / p.z};
                float rho3d = (s.x * s.x + s.y * s.y);
                float2 d = {xy.x - pixf.x, xy.y - pixf.y};
                float rho2d = SB_FILTER_INV_SQUARE * (d.x * d.x + d.y * d.y);
                float rho = fminf(rho3d, rho2d);
                float depth = (rho3d <= rho2d) ? (s.x * Tw.x + s.y * Tw.y) + Tw.z : Tw.z;
                if (depth < SB_NEAR) continue;
                float opa = q2.w;
                float power = -0.5f * rho;
                if (power > 0.0f) continue;
                float alpha = fminf(0.99f, opa * expf(power));
                if (alpha < 1.0f / 255.0f) continue;
                float test_T = T * (1 - alpha);
                if (test_T < 0.0001f) {
                    done = true;
                    continue;
                }
                float w = alpha * T;
                const float4 q3 = sr[j][3], q4 = sr[j][4];
                // distortion, depth, moments (forward.cu:444-449)
                float A = 1 - T;
                float m = SB_FAR / (SB_FAR - SB_NEAR) * (1 - SB_NEAR / depth);
                distortion += (m * m * A + M2 - 2 * m * M1) * w;
                D += depth * w;
                M1 += m * w;
                M2 += m * m * w;
                if (T > 0.5) {
                    median_depth = depth;
                    median_contributor = contributor;
                }
                N[0] += q3.x * w;
                N[1] += q3.y * w;
                N[2] += q3.z * w;
                C[0] += q3.w * w;
                C[1] += q4.x * w;
                C[2] += q4.y * w;
                T = test_T;
                last_contributor = contributor;
            }
            // a pixel that left the loop early still has to count the rest of the batch
            // (contributor is only used while !done, so no fix-up is needed)
        }
        __syncthreads();  // everyone finished reading buffer b&1 before batch b+2 lands in it
    }
    __pipeline_wait_prior(0);

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
                            char* image, float* out_color, float* out_others, cudaStream_t s) {
    SbImageLayout IL = sb_image_layout(W, H);
    const int grid_x = (W + SB_TILE_X - 1) / SB_TILE_X, grid_y = (H + SB_TILE_Y - 1) / SB_TILE_Y;
    // geom layout: the record array is first and its offset does not depend on P
    sb_blend_forward_kernel<<<grid_x * grid_y, FB_THREADS, 0, s>>>(
        W, H, grid_x, (const uint2*)(image + IL.ranges), point_list, (const float4*)geom, bg,
        (float*)(image + IL.final_T), (uint32_t*)(image + IL.n_contrib), out_color, out_others);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}
