// Per-(pixel, surfel) ray-splat intersection, bit-pinned.
//
// The gates of the blend -- alpha < 1/255, T(1-alpha) < 1e-4, T > 0.5, rho3d <= rho2d, depth < near,
// p.z == 0 (DSR/cuda_rasterizer/forward.cu:405-438, backward.cu:673-694) -- switch whole surfels on and
// off per pixel, and the reference's live backward re-blends pairs its forward skipped, so
// a one-ulp difference in the gated quantities can change a gradient by O(1) around edge-on surfels.
// The operation order below is therefore the one nvcc 12.9 emits at sm_100 for the reference (read off
// the SASS of oracle/_ref: renderCUDA<3> and PerGaussianRenderCUDA<3>), written with non-contractable
// intrinsics.  Forward and backward of the reference agree on k, l, p, s, rho3d and the depth dot
// product, but NOT on rho2d: the forward contracts it as fma(dy,dy,dx*dx), the backward as
// fma(dx,dx,dy*dy).  Both are provided; the backward kernel uses the backward form for its own gates
// and the forward form to decide what the forward did.
#pragma once
#include "common.cuh"

struct SbPair {
    float3 k, l, p;
    float2 s;      // p.xy / p.z (IEEE division)
    float rho3d;   // s.x^2 + s.y^2
    float sTw;     // s.x*Tw.x + s.y*Tw.y
};

__device__ __forceinline__ SbPair sb_pair(float2 pixf, float3 Tu, float3 Tv, float3 Tw) {
    SbPair g;
    g.k.x = __fmaf_rn(pixf.x, Tw.x, -Tu.x);
    g.k.y = __fmaf_rn(pixf.x, Tw.y, -Tu.y);
    g.k.z = __fmaf_rn(pixf.x, Tw.z, -Tu.z);
    g.l.x = __fmaf_rn(pixf.y, Tw.x, -Tv.x);
    g.l.y = __fmaf_rn(pixf.y, Tw.y, -Tv.y);
    g.l.z = __fmaf_rn(pixf.y, Tw.z, -Tv.z);
    // cross(k, l) (auxiliary.h:167): a.y*b.z - a.z*b.y, a.z*b.x - a.x*b.z, a.x*b.y - a.y*b.x
    g.p.x = __fmaf_rn(g.k.y, g.l.z, -__fmul_rn(g.k.z, g.l.y));
    g.p.y = __fmaf_rn(g.k.z, g.l.x, -__fmul_rn(g.k.x, g.l.z));
    g.p.z = __fmaf_rn(g.k.x, g.l.y, -__fmul_rn(g.k.y, g.l.x));
    g.s.x = __fdiv_rn(g.p.x, g.p.z);
    g.s.y = __fdiv_rn(g.p.y, g.p.z);
    g.rho3d = __fmaf_rn(g.s.x, g.s.x, __fmul_rn(g.s.y, g.s.y));
    g.sTw = __fmaf_rn(Tw.x, g.s.x, __fmul_rn(Tw.y, g.s.y));
    return g;
}

// rho2d = FilterInvSquare * |xy - pixel|^2 with d = xy - pixf (auxiliary.h:39, forward.cu:413-414)
__device__ __forceinline__ float sb_rho2d_forward(float2 d) {
    const float r = __fmaf_rn(d.y, d.y, __fmul_rn(d.x, d.x));
    return __fadd_rn(r, r);
}
__device__ __forceinline__ float sb_rho2d_backward(float2 d) {
    const float r = __fmaf_rn(d.x, d.x, __fmul_rn(d.y, d.y));
    return __fadd_rn(r, r);
}

// depth of the intersection, or of the surfel centre when the low-pass filter wins (forward.cu:416)
__device__ __forceinline__ float sb_pair_depth(const SbPair& g, float rho2d, float Tw_z) {
    return (g.rho3d <= rho2d) ? __fadd_rn(Tw_z, g.sTw) : Tw_z;
}

// m = far/(far-near) * (1 - near/depth) as the reference evaluates it: (-near/depth + 1) * 1.002004
__device__ __forceinline__ float sb_ndc_depth(float depth) {
    return __fmul_rn(__fadd_rn(__fdiv_rn(-SB_NEAR, depth), 1.0f), SB_FAR / (SB_FAR - SB_NEAR));
}

// ---- warp-level cull, interval form: can a whole 8x4 pixel block skip a surfel? ----
// One lane answers for ONE surfel and the warp's whole block (centre (bcx, bcy), pixel centres within +-3.5 / +-1.5 of
// it).  q = C + A d.x + B d.y is affine in the pixel, so over the block every component lies in an interval around its
// value at the block centre:  |q_i - q_i(centre)| <= r_i = 3.5 |A_i| + 1.5 |B_i|.  Then
//      E = q.x^2 + q.y^2 - tau q.z^2  >=  max(|q.x| - r.x, 0)^2 + max(|q.y| - r.y, 0)^2 - tau (|q.z| + r.z)^2
// and the block is skipped when that bound is positive and the rectangle's nearest point to the surfel's centre is
// outside the low-pass disc -- i.e. when the lane-level test would reject all 32 pixels.  Unlike sb_block_culled below
// this holds for ANY form (ellipse, hyperbola of a surfel that crosses the camera plane -- their boxes are unbounded and
// they are the bulk of the pairs of the 1080p-class configs), needs no division, and is centred on the block, so its
// slack does not grow with the distance to a far-away surfel centre.
// Rounding: the lane-level test sees q with an error of a few ulps of G_i = |C_i| + |A_i d.x| + |B_i d.y| (the terms
// cancel for far-away centres); every interval is widened by 2e-6 G_i (~33 ulps), and the final comparison carries a
// relative 1e-4 for the rounding of the three squares.  NaNs compare false: the block is visited.
// CPU model + conservativeness test against the lane-level test: tests/test_block_cull_cpu.py.
__device__ __forceinline__ bool sb_block_far(float4 c0, float4 c1, float4 c2, float bcx, float bcy) {
    const float tau = c2.w;
    const float dxc = c0.w - bcx, dyc = c1.w - bcy;   // d = xy - pixel at the block centre
    const float adx = fabsf(dxc), ady = fabsf(dyc);
    const float nx = fmaxf(adx - 3.5f, 0.0f), ny = fmaxf(ady - 1.5f, 0.0f);
    const float r2 = nx * nx + ny * ny;
    if (!((r2 + r2) > tau * 1.0001f + 1e-4f)) return false;  // the low-pass disc reaches the block
    const float Av[3] = {c0.x, c0.y, c0.z}, Bv[3] = {c1.x, c1.y, c1.z}, Cv[3] = {c2.x, c2.y, c2.z};
    float b[3];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const float q = Cv[i] + Av[i] * dxc + Bv[i] * dyc;
        float r = 3.5f * fabsf(Av[i]) + 1.5f * fabsf(Bv[i]);
        const float g = fabsf(Cv[i]) + fabsf(Av[i]) * adx + fabsf(Bv[i]) * ady + r;
        r += 2e-6f * g;
        b[i] = (i < 2) ? fmaxf(fabsf(q) - r, 0.0f) : fabsf(q) + r;
    }
    return (b[0] * b[0] + b[1] * b[1]) > 1.0001f * tau * (b[2] * b[2]);
}

// ---- warp-level cull, exact form (opt-in, -DSB_BLOCK_CULL; ellipses only): can a whole pixel block skip a surfel? ----
// The lane-level test of the blend kernels rejects pixel (px,py) iff  E(d) > 0  &&  2|d|^2 > tau  with
// d = xy - pixel and E(d) = q.x^2 + q.y^2 - tau q.z^2, q = C + A d.x + B d.y (record quads 0-2).  One lane
// evaluates this for ONE surfel against the warp's whole block [bx0,bx1] x [by0,by1] (pixel centres, inclusive):
// the block is skipped when the minimum of E over the rectangle is positive (E is a quadratic form; when it is
// an ellipse -- positive definite -- its minimum over a rectangle is at the stationary point if that is inside,
// else on one of the four edges, each a clamped 1-D parabola) and the rectangle's nearest point to the centre
// is outside the low-pass disc.  Hyperbolic forms (surfels crossing the camera plane) are never skipped here.
// The margin covers fp32 rounding of both evaluations (expanded coefficients here, factored form per pixel):
// every error term is bounded by eps * S with S the sum of the magnitudes of the terms of E.
__device__ __forceinline__ bool sb_block_culled(float4 c0, float4 c1, float4 c2, float bx0, float bx1, float by0,
                                                float by1) {
    const float tau = c2.w;
    const float xl = c0.w - bx1, xh = c0.w - bx0, yl = c1.w - by1, yh = c1.w - by0;  // range of d over the block
    const float nx = fmaxf(xl, fminf(0.0f, xh)), ny = fmaxf(yl, fminf(0.0f, yh));
    const float r2 = nx * nx + ny * ny;
    if (!((r2 + r2) > tau * 1.0001f + 1e-4f)) return false;  // the low-pass disc reaches the block
    const float a = c0.x * c0.x + c0.y * c0.y - tau * c0.z * c0.z;
    const float c = c1.x * c1.x + c1.y * c1.y - tau * c1.z * c1.z;
    const float b = 2.0f * (c0.x * c1.x + c0.y * c1.y - tau * c0.z * c1.z);
    const float d = 2.0f * (c2.x * c0.x + c2.y * c0.y - tau * c2.z * c0.z);
    const float e = 2.0f * (c2.x * c1.x + c2.y * c1.y - tau * c2.z * c1.z);
    const float f = c2.x * c2.x + c2.y * c2.y - tau * c2.z * c2.z;
    const float det = 4.0f * a * c - b * b;
    if (!(a > 0.0f && c > 0.0f && det > 0.0f)) return false;  // not an ellipse (or NaN): visit
    const float X = fmaxf(fabsf(xl), fabsf(xh)), Y = fmaxf(fabsf(yl), fabsf(yh));
    const float sx_ = fabsf(c2.x) + fabsf(c0.x) * X + fabsf(c1.x) * Y;
    const float sy_ = fabsf(c2.y) + fabsf(c0.y) * X + fabsf(c1.y) * Y;
    const float sz_ = fabsf(c2.z) + fabsf(c0.z) * X + fabsf(c1.z) * Y;
    const float margin = 1e-4f * (sx_ * sx_ + sy_ * sy_ + tau * sz_ * sz_);
    // vertical edges dx = X0: g(dy) = c dy^2 + (b X0 + e) dy + (a X0^2 + d X0 + f)
    const float ic = -0.5f / c, ia = -0.5f / a;
    float m;
    {
        const float l1 = b * xl + e, k1 = (a * xl + d) * xl + f;
        const float t1 = fminf(fmaxf(l1 * ic, yl), yh);
        m = (c * t1 + l1) * t1 + k1;
        const float l2 = b * xh + e, k2 = (a * xh + d) * xh + f;
        const float t2 = fminf(fmaxf(l2 * ic, yl), yh);
        m = fminf(m, (c * t2 + l2) * t2 + k2);
    }
    // horizontal edges dy = Y0: h(dx) = a dx^2 + (b Y0 + d) dx + (c Y0^2 + e Y0 + f)
    {
        const float l1 = b * yl + d, k1 = (c * yl + e) * yl + f;
        const float t1 = fminf(fmaxf(l1 * ia, xl), xh);
        m = fminf(m, (a * t1 + l1) * t1 + k1);
        const float l2 = b * yh + d, k2 = (c * yh + e) * yh + f;
        const float t2 = fminf(fmaxf(l2 * ia, xl), xh);
        m = fminf(m, (a * t2 + l2) * t2 + k2);
    }
    // stationary point
    const float sx = (b * e - 2.0f * c * d) / det, sy = (b * d - 2.0f * a * e) / det;
    if (sx >= xl && sx <= xh && sy >= yl && sy <= yh) return false;  // the ellipse's centre is inside the block
    return m > margin;
}
