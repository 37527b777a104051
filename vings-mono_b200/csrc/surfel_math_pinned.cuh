// Bit-pinned forward ray-splat transform.
//
// Radii, tile rects and depth keys must match the reference build BIT FOR BIT (BASELINE.json), and they
// are discontinuous functions of T.  Which of a*b + c*d becomes the FMUL and which the FFMA is decided
// by nvcc's instruction selection from the surrounding code, so restating the reference's expressions
// (surfel_math.cuh) reproduces its bits only for ~97 % of the surfels.  This file therefore spells out,
// with non-contractable intrinsics, the exact operation sequence nvcc 12.9 emits at sm_100 for
// compute_transmat + compute_aabb of the reference (DSR/cuda_rasterizer/forward.cu:83-157, with GLM's
// operators, auxiliary.h:225-245,295-302), read off the SASS of the oracle build (oracle/_ref) with
// cuobjdump and a small dataflow lister (tools/sass_dataflow.py).  The multiply-by-zero terms GLM's
// matrix products leave behind are kept: they are exact no-ops for finite inputs and free.
//   q0..q3 = rotation (w,x,y,z as stored), s0,s1 = scales, p0..p2 = mean, pm/vm = the 16 floats of the
//   projection / view tensors.  T = (Tu, Tv, Tw) row-major 9 floats, n = view-space normal (flipped).
#pragma once
#include "common.cuh"

__device__ __forceinline__ void sb_transmat_pinned(float q0, float q1, float q2, float q3, float s0, float s1,
                                                   float p0, float p1, float p2, const float* __restrict__ pm,
                                                   const float* __restrict__ vm, int W, int H, float* T, float3& n) {
    const float t1 = __fmul_rn(q3, q3);
    const float t2 = __fmaf_rn(q0, q0, t1);
    const float t3 = __fmaf_rn(q1, q1, t2);
    const float t4 = __fmaf_rn(q2, q2, t3);
    const float t7 = rsqrtf(t4);
    const float t8 = __fmul_rn(q0, t7);
    const float t9 = __fmul_rn(q3, t7);
    const float t10 = __fmul_rn(q2, t7);
    const float t11 = __fmul_rn(q1, t7);
    const float t12 = __fmul_rn(t8, t9);
    const float t13 = __fmul_rn(t8, t10);
    const float t14 = __fmul_rn(t8, t11);
    const float t15 = __fmul_rn(t10, t10);
    const float t16 = __fmul_rn(t9, t9);
    const float t17 = __fmaf_rn(t10, t9, t14);
    const float t18 = __fmaf_rn(t10, t9, -t14);
    const float t19 = __fadd_rn(t15, t16);
    const float t20 = __fmaf_rn(t11, t10, t12);
    const float t21 = __fmaf_rn(t11, t10, -t12);
    const float t22 = __fmaf_rn(t11, t9, -t13);
    const float t23 = __fadd_rn(t19, t19);
    const float t24 = __fadd_rn(t17, t17);
    const float t25 = __fmaf_rn(t11, t9, t13);
    const float t26 = __fmaf_rn(t11, t11, t16);
    const float t27 = __fmaf_rn(t11, t11, t15);
    const float t28 = __fadd_rn(-t23, 1.0f);
    const float t29 = __fadd_rn(t22, t22);
    const float t30 = __fmul_rn(0.0f, t24);
    const float t31 = __fmul_rn(s1, t24);
    const float t32 = __fadd_rn(t21, t21);
    const float t33 = __fmul_rn(s0, t28);
    const float t34 = __fmul_rn(0.0f, t28);
    const float t35 = __fmaf_rn(0.0f, t29, t30);
    const float t36 = __fadd_rn(t26, t26);
    const float t37 = __fmaf_rn(s0, t29, t30);
    const float t38 = __fmaf_rn(0.0f, t29, t31);
    const float t39 = __fmaf_rn(0.0f, t32, t33);
    const float t40 = __fmaf_rn(0.0f, t32, t34);
    const float t41 = __fmaf_rn(s1, t32, t34);
    const float t42 = __fadd_rn(-t36, 1.0f);
    const float t43 = __fadd_rn(t20, t20);
    const float t44 = __fmul_rn(0.0f, t42);
    const float t45 = __fmul_rn(s1, t42);
    const float t46 = __fadd_rn(t27, t27);
    const float t47 = __fadd_rn(t18, t18);
    const float t48 = __fmaf_rn(s0, t43, t44);
    const float t49 = __fmaf_rn(0.0f, t43, t45);
    const float t50 = __fadd_rn(-t46, 1.0f);
    const float t51 = __fadd_rn(t25, t25);
    const float t52 = __fmaf_rn(0.0f, t47, t48);
    const float t53 = __fmaf_rn(0.0f, t47, t49);
    const float t54 = __fmaf_rn(0.0f, t43, t44);
    const float t55 = __fmaf_rn(0.0f, t50, t37);
    const float t56 = __fadd_rn(t51, t40);
    const float t57 = __fmaf_rn(0.0f, t51, t39);
    const float t58 = __fmaf_rn(0.0f, t51, t41);
    const float t59 = __fmul_rn(pm[4], p1);
    const float t60 = __fmul_rn(t52, pm[4]);
    const float t61 = __fmul_rn(t53, pm[4]);
    const float t62 = __fmaf_rn(0.0f, t50, t38);
    const float t63 = __fmaf_rn(pm[0], p0, t59);
    const float t64 = __fmaf_rn(t57, pm[0], t60);
    const float t65 = __fmaf_rn(t58, pm[0], t61);
    const float t66 = __fadd_rn(t50, t35);
    const float t67 = __fadd_rn(t47, t54);
    const float t68 = __fmaf_rn(pm[8], p2, t63);
    const float t69 = __fmaf_rn(t55, pm[8], t64);
    const float t70 = __fmaf_rn(t62, pm[8], t65);
    const float t71 = __fmul_rn(pm[5], p1);
    const float t72 = __fmul_rn(t52, pm[5]);
    const float t73 = __fmul_rn(t53, pm[5]);
    const float t74 = __fmaf_rn(pm[1], p0, t71);
    const float t75 = __fmaf_rn(t58, pm[1], t73);
    const float t76 = __fmaf_rn(t57, pm[1], t72);
    const float t77 = __fmaf_rn(pm[9], p2, t74);
    const float t78 = __fmaf_rn(t55, pm[9], t76);
    const float t79 = __fmaf_rn(t62, pm[9], t75);
    const float t80 = __fadd_rn(pm[13], t77);
    const float t81 = __fadd_rn(pm[12], t68);
    const float t82 = __fmaf_rn(0.0f, pm[12], t69);
    const float t83 = __fmaf_rn(0.0f, pm[12], t70);
    const float t84 = __fmaf_rn(0.0f, pm[13], t78);
    const float t85 = __fmaf_rn(0.0f, pm[13], t79);
    const float t86 = __fmul_rn(t52, pm[6]);
    const float t87 = __fmaf_rn(t57, pm[2], t86);
    const float t88 = __fmul_rn(t52, pm[7]);
    const float t89 = __fmaf_rn(t57, pm[3], t88);
    const float t90 = __fmul_rn(pm[6], p1);
    const float t91 = __fmul_rn(pm[7], p1);
    const float t92 = __fmul_rn(0.0f, t80);
    const float t93 = __fmaf_rn(t55, pm[11], t89);
    const float t94 = __fmaf_rn(t55, pm[10], t87);
    const float t95 = __fmul_rn(t53, pm[6]);
    const float t96 = __fmaf_rn(pm[2], p0, t90);
    const float t97 = __fmul_rn(t53, pm[7]);
    const float t98 = __fmaf_rn(t58, pm[2], t95);
    const float t99 = (float)H;
    const float t100 = __fmaf_rn(t58, pm[3], t97);
    const float t101 = __fmul_rn(t99, 0.5f);
    const float t102 = (float)W;
    const float t103 = __fmul_rn(t101, t85);
    const float t104 = __fmaf_rn(pm[3], p0, t91);
    const float t105 = __fmul_rn(t102, 0.5f);
    const float t106 = __fmul_rn(0.0f, t85);
    const float t107 = __fmaf_rn(0.0f, t83, t103);
    const float t108 = __fmul_rn(t101, t80);
    const float t109 = __fmaf_rn(t105, t81, t92);
    const float t110 = __fmaf_rn(t105, t83, t106);
    const float t111 = __fmaf_rn(0.0f, t83, t106);
    const float t112 = __fmaf_rn(0.0f, t81, t108);
    const float t113 = __fmaf_rn(0.0f, t81, t92);
    const float t114 = __fmaf_rn(t62, pm[11], t100);
    const float t115 = __fmaf_rn(pm[10], p2, t96);
    const float t116 = __fmaf_rn(pm[11], p2, t104);
    const float t117 = __fmul_rn(t101, t84);
    const float t118 = __fmul_rn(0.0f, t84);
    const float t119 = __fadd_rn(pm[14], t115);
    const float t120 = __fmaf_rn(0.0f, t82, t117);
    const float t121 = __fmaf_rn(t62, pm[10], t98);
    const float t122 = __fmaf_rn(t105, t82, t118);
    const float t123 = __fmaf_rn(0.0f, t82, t118);
    const float t124 = (float)(H - 1);
    const float t125 = __fmaf_rn(0.0f, t119, t112);
    const float t126 = (float)(W - 1);
    const float t127 = __fmaf_rn(0.0f, t119, t109);
    const float t128 = __fmaf_rn(0.0f, t119, t113);
    const float t129 = __fadd_rn(pm[15], t116);
    const float t130 = __fmaf_rn(0.0f, pm[15], t93);
    const float t131 = __fmul_rn(t124, 0.5f);
    const float t132 = __fmaf_rn(0.0f, pm[15], t114);
    const float t133 = __fmul_rn(t126, 0.5f);
    const float t134 = __fmaf_rn(t131, t129, t125);
    const float t135 = __fmaf_rn(t133, t129, t127);
    const float t136 = __fmaf_rn(0.0f, pm[14], t94);
    const float t137 = __fmaf_rn(0.0f, pm[14], t121);
    const float t138 = __fmaf_rn(0.0f, t136, t120);
    const float t139 = __fmaf_rn(0.0f, t137, t107);
    const float t140 = __fmaf_rn(0.0f, t136, t122);
    const float t141 = __fmaf_rn(0.0f, t136, t123);
    const float t142 = __fadd_rn(t129, t128);
    const float t143 = __fmaf_rn(t131, t130, t138);
    const float t144 = __fmaf_rn(t131, t132, t139);
    const float t145 = __fmaf_rn(0.0f, t137, t110);
    const float t146 = __fmaf_rn(0.0f, t137, t111);
    const float t147 = __fmaf_rn(t133, t132, t145);
    const float t148 = __fmaf_rn(t133, t130, t140);
    const float t149 = __fadd_rn(t132, t146);
    T[2] = t135;
    T[5] = t134;
    T[8] = t142;
    T[0] = t148;
    T[1] = t147;
    T[3] = t143;
    T[4] = t144;
    T[7] = t149;
    const float t150 = __fmul_rn(t67, vm[6]);
    const float t151 = __fadd_rn(t130, t141);
    const float t152 = __fmaf_rn(t56, vm[2], t150);
    T[6] = t151;
    const float t153 = __fmaf_rn(t66, vm[10], t152);
    const float t154 = __fmul_rn(t67, vm[4]);
    const float t155 = __fmul_rn(t67, vm[5]);
    const float t156 = __fmaf_rn(t56, vm[1], t155);
    const float t157 = (t153 < 0.0f) ? 1.0f : -1.0f;
    const float t158 = __fmaf_rn(t56, vm[0], t154);
    const float t159 = __fmaf_rn(t66, vm[8], t158);
    const float t160 = __fmaf_rn(t66, vm[9], t156);
    const float t161 = __fmul_rn(t159, t157);
    const float t162 = __fmul_rn(t153, t157);
    const float t163 = __fmul_rn(t160, t157);
    n = make_float3(t161, t163, t162);
}

// compute_aabb (forward.cu:130-157) + radius (forward.cu:228) + getRect (auxiliary.h:66-76), in the
// operation order of the oracle build's SASS.  f = (1/d)*(1,1,-1) is folded into the operands
// (inv*T, inv*-T.z); 1/d is rcp.rn; the divisions by the tile size are exact multiplications by 1/16.
// Returns false when the reference returns early (d == 0).
__device__ __forceinline__ bool sb_aabb_pinned(const float* T, float2& point_image, float& radius) {
    const float d0 = __fmul_rn(T[7], T[7]);
    const float d1 = __fmaf_rn(T[6], T[6], d0);
    const float dist = __fmaf_rn(-T[8], T[8], d1);
    if (!(dist != 0.0f)) return false;
    const float inv = __frcp_rn(dist);
    const float f0x = __fmul_rn(inv, T[0]);
    const float f0y = __fmul_rn(inv, T[1]);
    const float f0z = __fmul_rn(inv, -T[2]);
    float tx = __fmul_rn(f0x, T[0]);
    float cx = __fmul_rn(f0x, T[6]);
    tx = __fmaf_rn(f0y, T[1], tx);
    cx = __fmaf_rn(f0y, T[7], cx);
    tx = __fmaf_rn(f0z, T[2], tx);
    cx = __fmaf_rn(f0z, T[8], cx);
    const float f1x = __fmul_rn(inv, T[3]);
    const float f1y = __fmul_rn(inv, T[4]);
    const float f1z = __fmul_rn(inv, -T[5]);
    float ty = __fmul_rn(f1x, T[3]);
    float cy = __fmul_rn(f1x, T[6]);
    ty = __fmaf_rn(f1y, T[4], ty);
    cy = __fmaf_rn(f1y, T[7], cy);
    ty = __fmaf_rn(f1z, T[5], ty);
    cy = __fmaf_rn(f1z, T[8], cy);
    const float hx = fmaxf(__fmaf_rn(cx, cx, -tx), 1e-4f);
    const float hy = fmaxf(__fmaf_rn(cy, cy, -ty), 1e-4f);
    const float ext = fmaxf(__fsqrt_rn(hy), __fsqrt_rn(hx));
    radius = __fmul_rn(ceilf(ext), 3.0f);
    point_image = make_float2(cx, cy);
    return true;
}

__device__ __forceinline__ void sb_rect_pinned(float2 p, int radius_i, uint2 grid, uint2& rmin, uint2& rmax) {
    const float r = (float)radius_i;
    const int x0 = (int)__fmul_rn(__fadd_rn(p.x, -r), 0.0625f);
    const int y0 = (int)__fmul_rn(__fadd_rn(p.y, -r), 0.0625f);
    const int x1 = (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(p.x, r), 16.0f), -1.0f), 0.0625f);
    const int y1 = (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(p.y, r), 16.0f), -1.0f), 0.0625f);
    rmin.x = min(grid.x, (unsigned)max(0, x0));
    rmin.y = min(grid.y, (unsigned)max(0, y0));
    rmax.x = min(grid.x, (unsigned)max(0, x1));
    rmax.y = min(grid.y, (unsigned)max(0, y1));
}
