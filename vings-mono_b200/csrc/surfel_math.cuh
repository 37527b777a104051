// Ray-splat transform arithmetic for one surfel.
//
// Radii, tile rects and depth keys are discontinuous in their inputs, and parity with the
// reference on them is bit-exact, so every expression below keeps the association (and the
// multiply-by-constant terms) of the reference's GLM-based code so that nvcc contracts FMAs the
// same way:  DSR/cuda_rasterizer/forward.cu:83-157, auxiliary.h:66-76,91-130,225-302 and the GLM
// operators it instantiates (type_mat3x3.inl:486-519, type_mat4x3.inl:505-563, type_mat4x4.inl:612-627).
// Matrices are column-major: m.c[col][row].
#pragma once
#include "common.cuh"

struct SbMat3 { float c[3][3]; };     // 3 columns of 3
struct SbMat3x4 { float c[3][4]; };   // 3 columns of 4
struct SbMat4x3 { float c[4][3]; };   // 4 columns of 3
struct SbMat4 { float c[4][4]; };     // 4 columns of 4

__device__ __forceinline__ SbMat3 sb_quat_to_rotmat(float4 quat) {
    // quaternion stored (w,x,y,z) in (.x,.y,.z,.w); auxiliary.h:225-245
    float s = rsqrtf(quat.w * quat.w + quat.x * quat.x + quat.y * quat.y + quat.z * quat.z);
    float w = quat.x * s;
    float x = quat.y * s;
    float y = quat.z * s;
    float z = quat.w * s;
    SbMat3 R;
    R.c[0][0] = 1.f - 2.f * (y * y + z * z);
    R.c[0][1] = 2.f * (x * y + w * z);
    R.c[0][2] = 2.f * (x * z - w * y);
    R.c[1][0] = 2.f * (x * y - w * z);
    R.c[1][1] = 1.f - 2.f * (x * x + z * z);
    R.c[1][2] = 2.f * (y * z + w * x);
    R.c[2][0] = 2.f * (x * z + w * y);
    R.c[2][1] = 2.f * (y * z - w * x);
    R.c[2][2] = 1.f - 2.f * (x * x + y * y);
    return R;
}

__device__ __forceinline__ SbMat3 sb_scale_to_mat(float2 scale, float glob_scale) {
    // auxiliary.h:295-302 -- identity with the two in-plane scales; scale_modifier is never applied (Q1)
    SbMat3 S;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) S.c[i][j] = (i == j) ? 1.f : 0.f;
    S.c[0][0] = glob_scale * scale.x;
    S.c[1][1] = glob_scale * scale.y;
    return S;
}

__device__ __forceinline__ SbMat3 sb_mul33(const SbMat3& a, const SbMat3& b) {
    SbMat3 r;
#pragma unroll
    for (int col = 0; col < 3; col++)
#pragma unroll
        for (int row = 0; row < 3; row++)
            r.c[col][row] = a.c[0][row] * b.c[col][0] + a.c[1][row] * b.c[col][1] + a.c[2][row] * b.c[col][2];
    return r;
}

// mat4x3 * mat4 -> mat4x3
__device__ __forceinline__ SbMat4x3 sb_mul43_44(const SbMat4x3& a, const SbMat4& b) {
    SbMat4x3 r;
#pragma unroll
    for (int col = 0; col < 4; col++)
#pragma unroll
        for (int row = 0; row < 3; row++)
            r.c[col][row] = a.c[0][row] * b.c[col][0] + a.c[1][row] * b.c[col][1] + a.c[2][row] * b.c[col][2] +
                            a.c[3][row] * b.c[col][3];
    return r;
}

// mat4x3 * mat3x4 -> mat3
__device__ __forceinline__ SbMat3 sb_mul43_34(const SbMat4x3& a, const SbMat3x4& b) {
    SbMat3 r;
#pragma unroll
    for (int col = 0; col < 3; col++)
#pragma unroll
        for (int row = 0; row < 3; row++)
            r.c[col][row] = a.c[0][row] * b.c[col][0] + a.c[1][row] * b.c[col][1] + a.c[2][row] * b.c[col][2] +
                            a.c[3][row] * b.c[col][3];
    return r;
}

// mat4 * mat3x4 -> mat3x4
__device__ __forceinline__ SbMat3x4 sb_mul44_34(const SbMat4& a, const SbMat3x4& b) {
    SbMat3x4 r;
#pragma unroll
    for (int col = 0; col < 3; col++)
#pragma unroll
        for (int row = 0; row < 4; row++)
            r.c[col][row] = a.c[0][row] * b.c[col][0] + a.c[1][row] * b.c[col][1] + a.c[2][row] * b.c[col][2] +
                            a.c[3][row] * b.c[col][3];
    return r;
}

__device__ __forceinline__ SbMat4 sb_world2ndc(const float* __restrict__ pm) {
    // forward.cu:106-111 -- the row-major tensor read so that columns are (pm[c], pm[4+c], pm[8+c], pm[12+c])
    SbMat4 m;
#pragma unroll
    for (int col = 0; col < 4; col++)
#pragma unroll
        for (int row = 0; row < 4; row++) m.c[col][row] = pm[4 * row + col];
    return m;
}

__device__ __forceinline__ SbMat3x4 sb_ndc2pix(int W, int H) {
    // forward.cu:113-117 (double divisions rounded to float by the vec4 constructor)
    SbMat3x4 m;
    m.c[0][0] = (float)(float(W) / 2.0); m.c[0][1] = 0.f; m.c[0][2] = 0.f; m.c[0][3] = (float)(float(W - 1) / 2.0);
    m.c[1][0] = 0.f; m.c[1][1] = (float)(float(H) / 2.0); m.c[1][2] = 0.f; m.c[1][3] = (float)(float(H - 1) / 2.0);
    m.c[2][0] = 0.f; m.c[2][1] = 0.f; m.c[2][2] = 0.f; m.c[2][3] = 1.f;
    return m;
}

__device__ __forceinline__ SbMat4x3 sb_splat2world_T(const SbMat3& L, float3 p) {
    // transpose of mat3x4( (L0,0), (L1,0), (p,1) ); forward.cu:100-104,119
    SbMat4x3 m;
    m.c[0][0] = L.c[0][0]; m.c[0][1] = L.c[1][0]; m.c[0][2] = p.x;
    m.c[1][0] = L.c[0][1]; m.c[1][1] = L.c[1][1]; m.c[1][2] = p.y;
    m.c[2][0] = L.c[0][2]; m.c[2][1] = L.c[1][2]; m.c[2][2] = p.z;
    m.c[3][0] = 0.f;       m.c[3][1] = 0.f;       m.c[3][2] = 1.f;
    return m;
}

__device__ __forceinline__ float3 sb_transform_vec4x3(float3 p, const float* __restrict__ m) {
    // auxiliary.h:112-120
    float3 t = {
        m[0] * p.x + m[4] * p.y + m[8] * p.z,
        m[1] * p.x + m[5] * p.y + m[9] * p.z,
        m[2] * p.x + m[6] * p.y + m[10] * p.z,
    };
    return t;
}

__device__ __forceinline__ float3 sb_transform_vec4x3_transpose(float3 p, const float* __restrict__ m) {
    // auxiliary.h:122-130
    float3 t = {
        m[0] * p.x + m[1] * p.y + m[2] * p.z,
        m[4] * p.x + m[5] * p.y + m[6] * p.z,
        m[8] * p.x + m[9] * p.y + m[10] * p.z,
    };
    return t;
}

// forward.cu:83-126. T columns = (Tu, Tv, Tw); normal in view space with the dual-visible flip.
__device__ __forceinline__ void sb_compute_transmat(float3 p_orig, float2 scale, float4 rot,
                                                    const float* __restrict__ projmatrix,
                                                    const float* __restrict__ viewmatrix, int W, int H,
                                                    SbMat3& T, float3& normal) {
    SbMat3 R = sb_quat_to_rotmat(rot);
    SbMat3 S = sb_scale_to_mat(scale, 1.0f);
    SbMat3 L = sb_mul33(R, S);
    SbMat4x3 s2wT = sb_splat2world_T(L, p_orig);
    SbMat4 world2ndc = sb_world2ndc(projmatrix);
    SbMat3x4 ndc2pix = sb_ndc2pix(W, H);
    SbMat4x3 A = sb_mul43_44(s2wT, world2ndc);
    T = sb_mul43_34(A, ndc2pix);
    float3 l2 = {L.c[2][0], L.c[2][1], L.c[2][2]};
    normal = sb_transform_vec4x3(l2, viewmatrix);
    float multiplier = normal.z < 0 ? 1 : -1;
    normal.x = multiplier * normal.x;
    normal.y = multiplier * normal.y;
    normal.z = multiplier * normal.z;
}

__device__ __forceinline__ float sb_sumf3(float3 a) { return a.x + a.y + a.z; }
__device__ __forceinline__ float3 sb_mul3(float3 a, float3 b) { return make_float3(a.x * b.x, a.y * b.y, a.z * b.z); }

// forward.cu:130-157
__device__ __forceinline__ bool sb_compute_aabb(const SbMat3& T, float2& point_image, float2& extent) {
    float3 T0 = {T.c[0][0], T.c[0][1], T.c[0][2]};
    float3 T1 = {T.c[1][0], T.c[1][1], T.c[1][2]};
    float3 T3 = {T.c[2][0], T.c[2][1], T.c[2][2]};
    float3 temp_point = {1.0f, 1.0f, -1.0f};
    float distance = sb_sumf3(sb_mul3(sb_mul3(T3, T3), temp_point));
    float inv = (1 / distance);
    float3 f = make_float3(inv * temp_point.x, inv * temp_point.y, inv * temp_point.z);
    if (distance == 0.0) return false;
    point_image = make_float2(sb_sumf3(sb_mul3(sb_mul3(f, T0), T3)), sb_sumf3(sb_mul3(sb_mul3(f, T1), T3)));
    float2 temp = make_float2(sb_sumf3(sb_mul3(sb_mul3(f, T0), T0)), sb_sumf3(sb_mul3(sb_mul3(f, T1), T1)));
    float2 half_extend = make_float2(point_image.x * point_image.x - temp.x, point_image.y * point_image.y - temp.y);
    extent = make_float2(sqrtf(fmaxf(1e-4, half_extend.x)), sqrtf(fmaxf(1e-4, half_extend.y)));
    return true;
}

// auxiliary.h:66-76
__device__ __forceinline__ void sb_get_rect(float2 p, int max_radius, uint2& rect_min, uint2& rect_max, uint2 grid) {
    rect_min.x = min(grid.x, max((int)0, (int)((p.x - max_radius) / SB_TILE_X)));
    rect_min.y = min(grid.y, max((int)0, (int)((p.y - max_radius) / SB_TILE_Y)));
    rect_max.x = min(grid.x, max((int)0, (int)((p.x + max_radius + SB_TILE_X - 1) / SB_TILE_X)));
    rect_max.y = min(grid.y, max((int)0, (int)((p.y + max_radius + SB_TILE_Y - 1) / SB_TILE_Y)));
}

// ---- conservative culling data (new; no counterpart in the reference) -----------------------------
// A (pixel, surfel) pair can only contribute if alpha = min(0.99, opa*exp(-rho/2)) >= 1/255 with
// rho = min(rho3d, rho2d) (forward.cu:412-435), i.e. rho <= 2 ln(255 opa).  The blend kernels skip a
// pair without running the reference arithmetic when BOTH rho3d and rho2d are provably above a padded
// cutoff tau; pairs that pass are evaluated with the reference's exact expression order, so skipping
// never changes a result.
//  * rho3d > tau  <=>  p.x^2 + p.y^2 > tau p.z^2 where p = (px Tw - Tu) x (py Tw - Tv) is AFFINE in the
//    pixel:  p = Tu x Tv + px (Tv x Tw) + py (Tw x Tu).  Re-centred on the surfel's screen centre xy and
//    with (dx,dy) = xy - pixel:  p = C + A dx + B dy.  Scaled by 1/|C.z| (the test is homogeneous).
//  * the screen-space box of {rho3d <= tau} follows compute_aabb (forward.cu:130-157) with the cutoff
//    weights (tau, tau, -1); it is united with the low-pass disk {rho2d <= tau} and padded by a pixel.
struct SbCull {
    float3 A, B, C;
    float tau;
    float4 bbox;  // xmin, xmax, ymin, ymax in pixel coordinates (pixel centres are integers)
};

__device__ __forceinline__ float3 sb_cross(float3 a, float3 b) {
    return make_float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}

__device__ __forceinline__ SbCull sb_cull_data(float3 Tu, float3 Tv, float3 Tw, float2 xy, float opa) {
    SbCull c;
    const float BIG = 1e30f;
    // padded cutoff: 1% + 0.01 covers the rounding of the affine form against the reference's k x l
    float t = 255.0f * opa;
    float tau = t > 1.0f ? 2.0f * logf(t) : 0.0f;
    tau = 1.01f * tau + 0.01f;
    c.tau = tau;
    float3 a = sb_cross(Tv, Tw), b = sb_cross(Tw, Tu), c0 = sb_cross(Tu, Tv);
    // p at the pixel (xy - d) = c0 + (xy.x - dx) a + (xy.y - dy) b
    float3 C = make_float3(c0.x + xy.x * a.x + xy.y * b.x, c0.y + xy.x * a.y + xy.y * b.y, c0.z + xy.x * a.z + xy.y * b.z);
    float n = 1.0f / fmaxf(fabsf(C.z), 1e-30f);
    c.A = make_float3(-a.x * n, -a.y * n, -a.z * n);
    c.B = make_float3(-b.x * n, -b.y * n, -b.z * n);
    c.C = make_float3(C.x * n, C.y * n, C.z * n);
    // box of the tau-level ellipse
    float xmin = -BIG, xmax = BIG, ymin = -BIG, ymax = BIG;
    float dist = tau * (Tw.x * Tw.x + Tw.y * Tw.y) - Tw.z * Tw.z;
    if (dist < 0.0f) {
        float inv = 1.0f / dist;
        float fx = tau * inv, fz = -inv;
        float cx = fx * (Tu.x * Tw.x + Tu.y * Tw.y) + fz * Tu.z * Tw.z;
        float cy = fx * (Tv.x * Tw.x + Tv.y * Tw.y) + fz * Tv.z * Tw.z;
        float hx = cx * cx - (fx * (Tu.x * Tu.x + Tu.y * Tu.y) + fz * Tu.z * Tu.z);
        float hy = cy * cy - (fx * (Tv.x * Tv.x + Tv.y * Tv.y) + fz * Tv.z * Tv.z);
        if (hx >= 0.0f && hy >= 0.0f) {  // false for NaN
            // pad: one pixel + a relative term for the cancellation in cx*cx - (...)
            float ex = sqrtf(hx) + 1.0f + 1e-3f * fabsf(cx), ey = sqrtf(hy) + 1.0f + 1e-3f * fabsf(cy);
            xmin = cx - ex; xmax = cx + ex; ymin = cy - ey; ymax = cy + ey;
        }
    }
    float rd = sqrtf(0.5f * tau) + 0.5f;  // low-pass disk rho2d = 2 |xy - pixel|^2 <= tau
    c.bbox = make_float4(fminf(xmin, xy.x - rd), fmaxf(xmax, xy.x + rd), fminf(ymin, xy.y - rd), fmaxf(ymax, xy.y + rd));
    if (!(c.bbox.x <= c.bbox.y) || !(c.bbox.z <= c.bbox.w)) c.bbox = make_float4(-BIG, BIG, -BIG, BIG);
    return c;
}

