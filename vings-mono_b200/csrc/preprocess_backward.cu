// K4b: backward of the per-surfel preprocess, plus the masked Adam step (next row N1).
//
// Restates preprocessCUDA/compute_transmat_aabb of the reference's backward
// (DSR/cuda_rasterizer/backward.cu:849-1045) with its quirks: W,H re-derived from focal*tan (Q8), the
// low-pass dL/dmean2D added twice before it feeds the centre terms (Q2), the one-term quotient rule
// (Q10), the un-normalised quaternion vjp (Q11, auxiliary.h:248-292) and the densification heuristic
// that overwrites dL/dmean2D (Q9).  Unlike the reference, which needs 120 B/surfel of zero-filled
// outputs (rasterize_points.cu:205-215), this kernel writes every dense output row itself (zeros for
// culled surfels), so the only memset left is the 80 B/surfel atomic accumulation record.
#include "prebwd_math.cuh"

// PRECOMP (transMat_precomp, backward.cu:875-883,950-964): T is the caller's; only dL/dtransMat (with the centre terms
// folded in when the low-pass gradient is non-zero), colours, opacity, scores and the densification heuristic
// are produced -- means3D / scales / rotations get zero rows.
template <bool PRECOMP>
__global__ void __launch_bounds__(256)
sb_preprocess_backward_kernel(int P, int Wq, int Hq, const float* __restrict__ means3D,
                              const float2* __restrict__ scales, const float4* __restrict__ rotations,
                              const float* __restrict__ transmat_precomp,
                              const float* __restrict__ viewmatrix, const float* __restrict__ projmatrix,
                              const int* __restrict__ radii, const float4* __restrict__ rec,
                              const float* __restrict__ grad_acc, float* __restrict__ dL_dmeans3D,
                              float* __restrict__ dL_dmeans2D, float* __restrict__ dL_dcolors,
                              float* __restrict__ dL_dopacity, float2* __restrict__ dL_dscales,
                              float4* __restrict__ dL_drotations, float* __restrict__ dL_dtransMat,
                              float2* __restrict__ scores) {
    __shared__ float s_vm[16], s_pm[16];
    if (threadIdx.x < 16) s_vm[threadIdx.x] = viewmatrix[threadIdx.x];
    else if (threadIdx.x < 32) s_pm[threadIdx.x - 16] = projmatrix[threadIdx.x - 16];
    __syncthreads();
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;

    float3 o_mean3D = {0.f, 0.f, 0.f}, o_color = {0.f, 0.f, 0.f};
    float2 o_mean2D = {0.f, 0.f}, o_scale = {0.f, 0.f}, o_score = {0.f, 0.f};
    float4 o_rot = {0.f, 0.f, 0.f, 0.f};
    float o_opacity = 0.f;
    float a[SB_GRAD_F];
#pragma unroll
    for (int i = 0; i < SB_GRAD_F; i++) a[i] = 0.f;

    if (radii[idx] > 0) {
        const float4* ga = reinterpret_cast<const float4*>(grad_acc + (size_t)idx * SB_GRAD_F);
#pragma unroll
        for (int i = 0; i < SB_GRAD_F / 4; i++) {
            float4 v = ga[i];
            a[4 * i] = v.x; a[4 * i + 1] = v.y; a[4 * i + 2] = v.z; a[4 * i + 3] = v.w;
        }
        sb_prebwd_surfel<PRECOMP>(idx, Wq, Hq, means3D, scales, rotations, transmat_precomp, s_vm, s_pm, a, o_mean3D,
                                  o_scale, o_rot);
        // densification heuristic replaces the screen-space gradient (Q9, :1041-1044); uses the raw
        // accumulated dL/dTu.z, dL/dTv.z and the *forward* Tw.z
        const float depth = rec[(size_t)idx * SB_REC_F4 + SB_REC_TW].z;
        o_mean2D.x = a[2] * depth * 0.5 * float(Wq);
        o_mean2D.y = a[5] * depth * 0.5 * float(Hq);
        o_color = make_float3(a[15], a[16], a[17]);
        o_opacity = a[14];
        o_score = make_float2(a[18], a[19]);
    }
    dL_dmeans3D[3 * idx] = o_mean3D.x; dL_dmeans3D[3 * idx + 1] = o_mean3D.y; dL_dmeans3D[3 * idx + 2] = o_mean3D.z;
    dL_dmeans2D[3 * idx] = o_mean2D.x; dL_dmeans2D[3 * idx + 1] = o_mean2D.y; dL_dmeans2D[3 * idx + 2] = 0.f;
    dL_dcolors[3 * idx] = o_color.x; dL_dcolors[3 * idx + 1] = o_color.y; dL_dcolors[3 * idx + 2] = o_color.z;
    dL_dopacity[idx] = o_opacity;
    dL_dscales[idx] = o_scale;
    dL_drotations[idx] = o_rot;
    scores[idx] = o_score;
    if (dL_dtransMat) {
#pragma unroll
        for (int i = 0; i < 9; i++) dL_dtransMat[9 * idx + i] = a[i];
    }
}

// adamUpdateCUDA (DSR/cuda_rasterizer/adam.cu:9-38): visibility-masked Adam without bias correction.
__global__ void sb_adam_kernel(float* __restrict__ param, const float* __restrict__ grad, float* __restrict__ exp_avg,
                               float* __restrict__ exp_avg_sq, const bool* __restrict__ visible, float lr, float b1,
                               float b2, float eps, uint32_t N, uint32_t M) {
    const uint32_t p_idx = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t g_idx = p_idx / M;
    if (g_idx >= N) return;
    if (visible[g_idx]) {
        float g = grad[p_idx];
        float m = exp_avg[p_idx];
        float v = exp_avg_sq[p_idx];
        float p = param[p_idx];
        sb_adam_scalar(p, m, v, g, lr, b1, b2, eps);
        param[p_idx] = p;
        exp_avg[p_idx] = m;
        exp_avg_sq[p_idx] = v;
    }
}

int sb_launch_preprocess_backward(int P, int W, int H, float tanfovx, float tanfovy, const float* means3D,
                                  const float* scales, const float* rotations, const float* transmat_precomp,
                                  const float* viewmatrix,
                                  const float* projmatrix, const int* radii, const char* geom, const float* grad_acc,
                                  float* dL_dmeans3D, float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity,
                                  float* dL_dscales, float* dL_drotations, float* dL_dtransMat, float* scores,
                                  cudaStream_t s) {
    if (P == 0) return 0;
    // Q8: the reference re-derives the image size from focal*tan in fp32 (backward.cu:1022-1023,
    // rasterizer_impl.cu:527-528); it can truncate to W-1 / H-1
    const float focal_y = H / (2.0f * tanfovy);
    const float focal_x = W / (2.0f * tanfovx);
    volatile float fx2 = focal_x * tanfovx;
    volatile float fy2 = focal_y * tanfovy;
    const int Wq = int(fx2 * 2);
    const int Hq = int(fy2 * 2);
    auto kern = transmat_precomp ? sb_preprocess_backward_kernel<true> : sb_preprocess_backward_kernel<false>;
    kern<<<(P + 255) / 256, 256, 0, s>>>(
        P, Wq, Hq, means3D, (const float2*)scales, (const float4*)rotations, transmat_precomp, viewmatrix, projmatrix, radii,
        (const float4*)geom, grad_acc, dL_dmeans3D, dL_dmeans2D, dL_dcolors, dL_dopacity, (float2*)dL_dscales,
        (float4*)dL_drotations, dL_dtransMat, (float2*)scores);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}

int sb_launch_adam(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, const bool* visible,
                   float lr, float b1, float b2, float eps, uint32_t N, uint32_t M, cudaStream_t s) {
    const uint32_t cnt = N * M;
    if (cnt == 0) return 0;
    sb_adam_kernel<<<(cnt + 255) / 256, 256, 0, s>>>(param, grad, exp_avg, exp_avg_sq, visible, lr, b1, b2, eps, N, M);
    sb_count_launch(1);
    SB_CUDA_CHECK(cudaGetLastError());
    return 0;
}
