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
