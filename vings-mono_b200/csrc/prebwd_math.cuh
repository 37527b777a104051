// Per-surfel backward of the preprocess: the arithmetic of preprocessCUDA / compute_transmat_aabb in the reference's
// backward (DSR/cuda_rasterizer/backward.cu:849-1045), shared by the stand-alone kernel (preprocess_backward.cu, the
// drop-in path's dense outputs) and the fused per-surfel pass of the mapping step (surfel_update.cu).
#pragma once
#include "surfel_math.cuh"

__device__ __forceinline__ float4 sb_quat_to_rotmat_vjp(float4 quat, const SbMat3& v_R) {
    float s = rsqrtf(quat.w * quat.w + quat.x * quat.x + quat.y * quat.y + quat.z * quat.z);
    float w = quat.x * s;
    float x = quat.y * s;
    float y = quat.z * s;
    float z = quat.w * s;
    float4 v;
    v.x = 2.f * (x * (v_R.c[1][2] - v_R.c[2][1]) + y * (v_R.c[2][0] - v_R.c[0][2]) + z * (v_R.c[0][1] - v_R.c[1][0]));
    v.y = 2.f * (-2.f * x * (v_R.c[1][1] + v_R.c[2][2]) + y * (v_R.c[0][1] + v_R.c[1][0]) +
                 z * (v_R.c[0][2] + v_R.c[2][0]) + w * (v_R.c[1][2] - v_R.c[2][1]));
    v.z = 2.f * (x * (v_R.c[0][1] + v_R.c[1][0]) - 2.f * y * (v_R.c[0][0] + v_R.c[2][2]) +
                 z * (v_R.c[1][2] + v_R.c[2][1]) + w * (v_R.c[2][0] - v_R.c[0][2]));
    v.w = 2.f * (x * (v_R.c[0][2] + v_R.c[2][0]) + y * (v_R.c[1][2] + v_R.c[2][1]) -
                 2.f * z * (v_R.c[0][0] + v_R.c[1][1]) + w * (v_R.c[0][1] - v_R.c[1][0]));
    return v;
}

// a[SB_GRAD_F]: the surfel's accumulated record (in); for PRECOMP a[0..8] returns the final dL/dtransMat.
// Outputs: dL/dmeans3D, dL/dscales, dL/drotations w.r.t. the rasterizer's (activated) inputs.
template <bool PRECOMP>
__device__ __forceinline__ void sb_prebwd_surfel(int idx, int Wq, int Hq, const float* __restrict__ means3D,
                                                 const float2* __restrict__ scales, const float4* __restrict__ rotations,
                                                 const float* __restrict__ transmat_precomp, const float* s_vm,
                                                 const float* s_pm, float* a, float3& o_mean3D, float2& o_scale,
                                                 float4& o_rot) {
    float4 rot = {1.f, 0.f, 0.f, 0.f};
    float2 scale = {0.f, 0.f};
    SbMat3 R, T;
    SbMat3x4 Pm;
    float3 normal = {0.f, 0.f, 0.f};
    if (PRECOMP) {
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int r = 0; r < 3; r++) T.c[c][r] = transmat_precomp[9 * (size_t)idx + 3 * c + r];
    } else {
        float3 p_orig = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
        rot = rotations[idx];
        scale = scales[idx];
        R = sb_quat_to_rotmat(rot);
        SbMat3 S = sb_scale_to_mat(scale, 1.0f);
        SbMat3 L = sb_mul33(R, S);
        SbMat4x3 MT = sb_splat2world_T(L, p_orig);
        SbMat4 world2ndc = sb_world2ndc(s_pm);
        SbMat3x4 ndc2pix = sb_ndc2pix(Wq, Hq);
        Pm = sb_mul44_34(world2ndc, ndc2pix);  // backward associates (world2ndc*ndc2pix) first, :909-910
        T = sb_mul43_34(MT, Pm);
        float3 l2 = {L.c[2][0], L.c[2][1], L.c[2][2]};
        normal = sb_transform_vec4x3(l2, s_vm);
    }

    SbMat3 dL_dT;
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
        for (int r = 0; r < 3; r++) dL_dT.c[c][r] = a[3 * c + r];
    // the reference adds the low-pass mean2D gradient with two identical atomics (Q2)
    float2 g2 = {a[9] + a[9], a[10] + a[10]};
    if (g2.x != 0 || g2.y != 0) {
        const float distance = T.c[2][0] * T.c[2][0] + T.c[2][1] * T.c[2][1] - T.c[2][2] * T.c[2][2];
        const float f = 1 / (distance);
        const float dpx_dT00 = f * T.c[2][0];
        const float dpx_dT01 = f * T.c[2][1];
        const float dpx_dT02 = -f * T.c[2][2];
        const float dpy_dT10 = f * T.c[2][0];
        const float dpy_dT11 = f * T.c[2][1];
        const float dpy_dT12 = -f * T.c[2][2];
        const float dpx_dT30 = T.c[0][0] * (f - 2 * f * f * T.c[2][0] * T.c[2][0]);
        const float dpx_dT31 = T.c[0][1] * (f - 2 * f * f * T.c[2][1] * T.c[2][1]);
        const float dpx_dT32 = -T.c[0][2] * (f + 2 * f * f * T.c[2][2] * T.c[2][2]);
        const float dpy_dT30 = T.c[1][0] * (f - 2 * f * f * T.c[2][0] * T.c[2][0]);
        const float dpy_dT31 = T.c[1][1] * (f - 2 * f * f * T.c[2][1] * T.c[2][1]);
        const float dpy_dT32 = -T.c[1][2] * (f + 2 * f * f * T.c[2][2] * T.c[2][2]);
        dL_dT.c[0][0] += g2.x * dpx_dT00;
        dL_dT.c[0][1] += g2.x * dpx_dT01;
        dL_dT.c[0][2] += g2.x * dpx_dT02;
        dL_dT.c[1][0] += g2.y * dpy_dT10;
        dL_dT.c[1][1] += g2.y * dpy_dT11;
        dL_dT.c[1][2] += g2.y * dpy_dT12;
        dL_dT.c[2][0] += g2.x * dpx_dT30 + g2.y * dpy_dT30;
        dL_dT.c[2][1] += g2.x * dpx_dT31 + g2.y * dpy_dT31;
        dL_dT.c[2][2] += g2.x * dpx_dT32 + g2.y * dpy_dT32;
    }
    if (PRECOMP) {
        // the (possibly updated) dL/dT is the output; the heuristic below reads it back (:950-961,1043-1044)
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int r = 0; r < 3; r++) a[3 * c + r] = dL_dT.c[c][r];
    } else {
        // dL_dM = P * transpose(dL_dT)   (mat3x4 * mat3 -> mat3x4), :965
        float dL_dM[3][4];
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int r = 0; r < 4; r++)
                dL_dM[c][r] = Pm.c[0][r] * dL_dT.c[0][c] + Pm.c[1][r] * dL_dT.c[1][c] + Pm.c[2][r] * dL_dT.c[2][c];
        float3 dLn = {a[11], a[12], a[13]};
        float3 dL_dtn = sb_transform_vec4x3_transpose(dLn, s_vm);
        float multiplier = normal.z < 0 ? 1 : -1;
        dL_dtn.x = multiplier * dL_dtn.x;
        dL_dtn.y = multiplier * dL_dtn.y;
        dL_dtn.z = multiplier * dL_dtn.z;
        SbMat3 dL_dR;
#pragma unroll
        for (int r = 0; r < 3; r++) {
            dL_dR.c[0][r] = dL_dM[0][r] * scale.x;
            dL_dR.c[1][r] = dL_dM[1][r] * scale.y;
        }
        dL_dR.c[2][0] = dL_dtn.x; dL_dR.c[2][1] = dL_dtn.y; dL_dR.c[2][2] = dL_dtn.z;
        o_rot = sb_quat_to_rotmat_vjp(rot, dL_dR);
        o_scale.x = dL_dM[0][0] * R.c[0][0] + dL_dM[0][1] * R.c[0][1] + dL_dM[0][2] * R.c[0][2];
        o_scale.y = dL_dM[1][0] * R.c[1][0] + dL_dM[1][1] * R.c[1][1] + dL_dM[1][2] * R.c[1][2];
        o_mean3D = make_float3(dL_dM[2][0], dL_dM[2][1], dL_dM[2][2]);
    }
}
