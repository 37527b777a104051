"""CPU: the warp-level block cull of the blend kernels (csrc/pair_math.cuh::sb_block_culled) is conservative.

A float32 numpy port of sb_cull_data / the lane-level test / sb_block_culled is run on the surfels of the
synthetic scenes (T matrices from the oracle's preprocess) against every 8x4 pixel block around each surfel:
a block may be skipped only if the lane-level test rejects all 32 of its pixels.  The CUDA code itself is
covered by the bit-exact parity tests (images, contributor counts) on the GPU.
"""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vings-mono_b200"), os.path.join(ROOT, "oracle")]
import surfel_emulation as emu  # noqa: E402
import synthetic  # noqa: E402

f32 = np.float32


def cross(a, b):
    return np.stack([a[:, 1] * b[:, 2] - a[:, 2] * b[:, 1], a[:, 2] * b[:, 0] - a[:, 0] * b[:, 2],
                     a[:, 0] * b[:, 1] - a[:, 1] * b[:, 0]], 1).astype(f32)


def cull_data(Tu, Tv, Tw, xy, opa):
    """surfel_math.cuh::sb_cull_data (quads 0-2 of the record)"""
    t = f32(255.0) * opa
    tau = np.where(t > 1, f32(2.0) * np.log(np.maximum(t, f32(1e-30))), f32(0)).astype(f32)
    tau = (f32(1.01) * tau + f32(0.01)).astype(f32)
    a, b, c0 = cross(Tv, Tw), cross(Tw, Tu), cross(Tu, Tv)
    C = (c0 + xy[:, 0:1] * a + xy[:, 1:2] * b).astype(f32)
    n = (f32(1.0) / np.maximum(np.abs(C[:, 2]), f32(1e-30)))[:, None].astype(f32)
    return (-a * n).astype(f32), (-b * n).astype(f32), (C * n).astype(f32), tau


def lane_culled(A, B, C, tau, xy, px, py):
    """the per-pixel test of blend_forward.cu / blend_backward.cu; px, py broadcast against the surfel axis"""
    dx, dy = (xy[:, 0][:, None] - px).astype(f32), (xy[:, 1][:, None] - py).astype(f32)
    r2 = (dx * dx + dy * dy).astype(f32)
    q = [(C[:, i][:, None] + A[:, i][:, None] * dx + B[:, i][:, None] * dy).astype(f32) for i in range(3)]
    E = ((q[0] * q[0] + q[1] * q[1]) - tau[:, None] * (q[2] * q[2])).astype(f32)
    return (E > 0) & ((r2 + r2) > tau[:, None])


def block_culled(A, B, C, tau, xy, bx0, bx1, by0, by1):
    """pair_math.cuh::sb_block_culled, same operation order, float32"""
    with np.errstate(all="ignore"):
        xl, xh = (xy[:, 0] - bx1).astype(f32), (xy[:, 0] - bx0).astype(f32)
        yl, yh = (xy[:, 1] - by1).astype(f32), (xy[:, 1] - by0).astype(f32)
        nx, ny = np.maximum(xl, np.minimum(f32(0), xh)), np.maximum(yl, np.minimum(f32(0), yh))
        r2 = (nx * nx + ny * ny).astype(f32)
        disc_out = (r2 + r2) > tau * f32(1.0001) + f32(1e-4)
        a = (A[:, 0] * A[:, 0] + A[:, 1] * A[:, 1] - tau * A[:, 2] * A[:, 2]).astype(f32)
        c = (B[:, 0] * B[:, 0] + B[:, 1] * B[:, 1] - tau * B[:, 2] * B[:, 2]).astype(f32)
        b = (f32(2) * (A[:, 0] * B[:, 0] + A[:, 1] * B[:, 1] - tau * A[:, 2] * B[:, 2])).astype(f32)
        d = (f32(2) * (C[:, 0] * A[:, 0] + C[:, 1] * A[:, 1] - tau * C[:, 2] * A[:, 2])).astype(f32)
        e = (f32(2) * (C[:, 0] * B[:, 0] + C[:, 1] * B[:, 1] - tau * C[:, 2] * B[:, 2])).astype(f32)
        f = (C[:, 0] * C[:, 0] + C[:, 1] * C[:, 1] - tau * C[:, 2] * C[:, 2]).astype(f32)
        det = (f32(4) * a * c - b * b).astype(f32)
        ellipse = (a > 0) & (c > 0) & (det > 0)
        X, Y = np.maximum(np.abs(xl), np.abs(xh)), np.maximum(np.abs(yl), np.abs(yh))
        s = [np.abs(C[:, i]) + np.abs(A[:, i]) * X + np.abs(B[:, i]) * Y for i in range(3)]
        margin = (f32(1e-4) * (s[0] * s[0] + s[1] * s[1] + tau * s[2] * s[2])).astype(f32)
        ic, ia = (f32(-0.5) / c).astype(f32), (f32(-0.5) / a).astype(f32)
        m = np.full_like(a, np.inf)
        for x0 in (xl, xh):
            l, k = (b * x0 + e).astype(f32), ((a * x0 + d) * x0 + f).astype(f32)
            t = np.minimum(np.maximum(l * ic, yl), yh).astype(f32)
            m = np.minimum(m, ((c * t + l) * t + k).astype(f32))
        for y0 in (yl, yh):
            l, k = (b * y0 + d).astype(f32), ((c * y0 + e) * y0 + f).astype(f32)
            t = np.minimum(np.maximum(l * ia, xl), xh).astype(f32)
            m = np.minimum(m, ((a * t + l) * t + k).astype(f32))
        sx, sy = ((b * e - f32(2) * c * d) / det).astype(f32), ((b * d - f32(2) * a * e) / det).astype(f32)
        centre_in = (sx >= xl) & (sx <= xh) & (sy >= yl) & (sy <= yh)
        return disc_out & ellipse & ~centre_in & (m > margin)


@pytest.mark.parametrize("cfg,P", [("C1", 4000), ("C3", 4000), ("C2", 4000)])
def test_block_cull_never_drops_a_passing_pixel(cfg, P):
    scene, views = synthetic.make_scene(cfg, P=P, W=320, H=240)
    inp = synthetic.view_inputs(scene, views[-1])
    pre = emu.preprocess(inp["means3D"], inp["scales"], inp["rotations"], inp["opacities"], inp["viewmatrix"],
                         inp["projmatrix"], inp["W"], inp["H"])
    vis = pre["visible"].numpy()
    T = pre["T"].numpy().astype(f32)[vis]
    xy = pre["means2D"].numpy().astype(f32)[vis]
    opa = inp["opacities"].numpy().astype(f32).reshape(-1)[vis]
    A, B, C, tau = cull_data(T[:, 0], T[:, 1], T[:, 2], xy, opa)
    n = len(xy)
    assert n > 500
    lx, ly = np.arange(8, dtype=f32), np.arange(4, dtype=f32)
    culled_blocks = kept_blocks = 0
    # every block of the 8x4 lattice within +-5 blocks in x and +-10 blocks in y of the surfel's centre
    cbx, cby = np.floor(xy[:, 0] / 8), np.floor(xy[:, 1] / 4)
    for ox in range(-5, 6):
        for oy in range(-10, 11):
            bx0, by0 = ((cbx + ox) * 8).astype(f32), ((cby + oy) * 4).astype(f32)
            bc = block_culled(A, B, C, tau, xy, bx0, bx0 + f32(7), by0, by0 + f32(3))
            px = (bx0[:, None] + np.tile(lx, 4)[None, :]).astype(f32)
            py = (by0[:, None] + np.repeat(ly, 8)[None, :]).astype(f32)
            lc = lane_culled(A, B, C, tau, xy, px, py)
            bad = bc & ~lc.all(1)
            assert not bad.any(), (cfg, ox, oy, int(bad.sum()))
            culled_blocks += int(bc.sum())
            kept_blocks += int((~bc & lc.all(1)).sum())  # visited although every lane rejects
    total = n * 11 * 21
    # the test is not vacuous: it skips most far blocks, and few all-rejecting blocks are still visited
    assert culled_blocks > 0.5 * total, (culled_blocks, total)
    assert kept_blocks < 0.2 * total, (kept_blocks, total)
