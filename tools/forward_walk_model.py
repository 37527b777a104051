"""CPU model of the forward blend's walk on sampled tiles of a synthetic config: where its warp instructions go and what
the warp-level interval cull (csrc/pair_math.cuh::sb_block_far) removes.  No GPU needed.

    python tools/forward_walk_model.py C5 [tiles=16] [view=-1]

For every sampled tile the depth-ordered list is rebuilt from the oracle's preprocess, all 256 pixels are blended in
float64 to find where each pixel terminates (the walk's length per warp), and for every (warp, 32-surfel batch) the walk
visits it counts: surfels passing the record's box test, surfels the interval cull removes, surfels some live lane
passes (phase-2 visits).  Instruction weights are rough per-visit counts read off the SASS (they reproduce the measured
per-pair totals to ~1.7x; the SHARES are what the model is for).  Output of the runs behind DESIGN.md section 5:
profiles/r02_forward_walk_model.txt.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vings-mono_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import surfel_emulation as emu  # noqa: E402
import synthetic  # noqa: E402
from test_block_cull_cpu import block_far, cull_data  # noqa: E402

f32 = np.float32
P1_TRIP = 33.0    # lane-level cull, two surfels per trip: 6 LDS.128 + ~22 FP + bit bookkeeping
P2_VISIT = 75.0   # a phase-2 visit some lane blends (reference arithmetic, two IEEE divisions, expf)
FAR_TEST = 56.0   # sb_block_far per (warp, batch): 3 LDS.128 + ~50 FP (SASS delta of the kernel)
FIXED = 45.0      # per (warp, batch): ring wait, box test, votes, hand-back


def record_bbox(Tu, Tv, Tw, xy, tau):
    """surfel_math.cuh::sb_cull_data, the box of quad 7 (float32)"""
    BIG = f32(1e30)
    with np.errstate(all="ignore"):
        dist = tau * (Tw[:, 0] ** 2 + Tw[:, 1] ** 2) - Tw[:, 2] ** 2
        inv = 1 / dist
        fx, fz = tau * inv, -inv
        cx = fx * (Tu[:, 0] * Tw[:, 0] + Tu[:, 1] * Tw[:, 1]) + fz * Tu[:, 2] * Tw[:, 2]
        cy = fx * (Tv[:, 0] * Tw[:, 0] + Tv[:, 1] * Tw[:, 1]) + fz * Tv[:, 2] * Tw[:, 2]
        hx = cx * cx - (fx * (Tu[:, 0] ** 2 + Tu[:, 1] ** 2) + fz * Tu[:, 2] ** 2)
        hy = cy * cy - (fx * (Tv[:, 0] ** 2 + Tv[:, 1] ** 2) + fz * Tv[:, 2] ** 2)
        ok = (dist < 0) & (hx >= 0) & (hy >= 0)
        ex = np.sqrt(np.maximum(hx, 0)) + 1 + 1e-3 * np.abs(cx)
        ey = np.sqrt(np.maximum(hy, 0)) + 1 + 1e-3 * np.abs(cy)
    xmin, xmax = np.where(ok, cx - ex, -BIG), np.where(ok, cx + ex, BIG)
    ymin, ymax = np.where(ok, cy - ey, -BIG), np.where(ok, cy + ey, BIG)
    rd = np.sqrt(0.5 * tau) + 0.5
    return (np.minimum(xmin, xy[:, 0] - rd), np.maximum(xmax, xy[:, 0] + rd), np.minimum(ymin, xy[:, 1] - rd),
            np.maximum(ymax, xy[:, 1] + rd))


def main():
    cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
    n_tiles = int(sys.argv[2]) if len(sys.argv) > 2 else 16
    vi = int(sys.argv[3]) if len(sys.argv) > 3 else -1
    scene, views = synthetic.make_scene(cfg)
    inp = synthetic.view_inputs(scene, views[vi])
    W, H = inp["W"], inp["H"]
    pre = emu.preprocess(inp["means3D"], inp["scales"], inp["rotations"], inp["opacities"], inp["viewmatrix"],
                         inp["projmatrix"], W, H)
    vis = pre["visible"].numpy()
    T = pre["T"].numpy().astype(f32)[vis]
    xy = pre["means2D"].numpy().astype(f32)[vis]
    opa = inp["opacities"].numpy().astype(f32).reshape(-1)[vis]
    depth, rect = pre["depth"].numpy()[vis], pre["rect"].numpy()[vis]
    tt = pre["tiles_touched"].numpy()[vis]
    gx, gy = pre["grid"]
    Tu, Tv, Tw = T[:, 0], T[:, 1], T[:, 2]
    A, B, C, tau = cull_data(Tu, Tv, Tw, xy, opa)
    bx0, bx1, by0, by1 = record_bbox(Tu, Tv, Tw, xy, tau)
    o = np.argsort(-tt)
    half = int(np.searchsorted(np.cumsum(tt[o]) / tt.sum(), 0.5)) + 1
    print("%s view %d: V %d visible surfels, R %d pairs, %d tiles; %d surfels (%.2f %%) own half of the pairs, %.0f %% of "
          "those have an unbounded box" % (cfg, vi, len(xy), tt.sum(), gx * gy, half, 100.0 * half / len(xy),
                                           100.0 * (bx1[o[:half]] > 1e20).mean()))
    order = np.argsort(depth.view(np.uint32), kind="stable")
    tiles = np.random.default_rng(1).choice(gx * gy, n_tiles, replace=False)
    lxs, lys = np.tile(np.arange(8), 4), np.repeat(np.arange(4), 8)
    acc = dict(list=0, walked=0, wb=0, box=0, far=0, live=0, p1=0.0, p1_far=0.0, p2=0, bad=0)
    for tile in tiles:
        tx, ty = tile % gx, tile // gx
        m = (rect[:, 0] <= tx) & (rect[:, 2] > tx) & (rect[:, 1] <= ty) & (rect[:, 3] > ty)
        lst = order[m[order]]
        n = len(lst)
        acc["list"] += n
        if n == 0:
            continue
        px = np.tile(tx * 16 + np.arange(16), 16).astype(f32)
        py = np.repeat(ty * 16 + np.arange(16), 16).astype(f32)
        k = px[:, None, None] * Tw[lst][None] - Tu[lst][None]
        l = py[:, None, None] * Tw[lst][None] - Tv[lst][None]
        p = np.cross(k, l)
        with np.errstate(all="ignore"):
            sx, sy = p[..., 0] / p[..., 2], p[..., 1] / p[..., 2]
            rho3d = sx * sx + sy * sy
            dx, dy = (xy[lst, 0][None] - px[:, None]).astype(f32), (xy[lst, 1][None] - py[:, None]).astype(f32)
            rho2d = 2 * (dx * dx + dy * dy)
            rho = np.where(np.isnan(rho3d), rho2d, np.minimum(rho3d, rho2d))
            dep = np.where(rho3d <= rho2d, sx * Tw[lst][None, :, 0] + sy * Tw[lst][None, :, 1] + Tw[lst][None, :, 2],
                           Tw[lst][None, :, 2])
            alpha = np.minimum(0.99, opa[lst][None] * np.exp(-0.5 * rho))
            ok = (p[..., 2] != 0) & ~(dep < 0.2) & ~(alpha < 1 / 255)
            q = [(C[lst, i][None] + A[lst, i][None] * dx + B[lst, i][None] * dy).astype(f32) for i in range(3)]
            E = ((q[0] * q[0] + q[1] * q[1]) - tau[lst][None] * (q[2] * q[2])).astype(f32)
            lanepass = ~((E > 0) & (rho2d > tau[lst][None]))
        donepos = np.full(256, n, dtype=np.int64)   # list position at which the pixel terminates
        stop = (np.cumprod(1 - np.where(ok, alpha, 0.0), axis=1) < 1e-4) & ok
        has = stop.any(1)
        donepos[has] = stop.argmax(1)[has]
        donepos[~((px < W) & (py < H))] = -1
        tile_end = min((n + 31) // 32, int(donepos.max()) // 32 + 1)
        acc["walked"] += min(n, tile_end * 32)
        for w in range(8):
            wx0, wy0 = tx * 16 + (w & 1) * 8, ty * 16 + (w >> 1) * 4
            pix = ((w >> 1) * 4 + lys) * 16 + (w & 1) * 8 + lxs
            dpos = donepos[pix]
            wend = min(tile_end, int(dpos.max()) // 32 + 1) if dpos.max() >= 0 else 0
            if wend == 0:
                continue
            box = (bx0[lst] <= wx0 + 7) & (bx1[lst] >= wx0) & (by0[lst] <= wy0 + 3) & (by1[lst] >= wy0)
            far = block_far(A[lst], B[lst], C[lst], tau[lst], xy[lst], f32(wx0), f32(wy0))
            lp = lanepass[pix]
            acc["bad"] += int((box & far & lp.any(0)).sum())
            pad = lambda v: np.pad(v[:wend * 32], (0, max(0, wend * 32 - len(v)))).reshape(wend, 32)
            bb, ff = pad(box), pad(box & far)
            live = dpos[:, None] >= (np.arange(wend) * 32)[None, :]            # [lane, batch]: live at the batch start
            lpp = np.pad(lp[:, :wend * 32], ((0, 0), (0, max(0, wend * 32 - lp.shape[1])))).reshape(32, wend, 32)
            am = (lpp & live[:, :, None]).any(0) & bb
            acc["wb"] += wend
            acc["box"] += int(bb.sum())
            acc["far"] += int(ff.sum())
            acc["live"] += int((bb & pad(lp.any(0))).sum())
            acc["p1"] += float((np.ceil(bb.sum(1) / 2) * P1_TRIP).sum())
            acc["p1_far"] += float((np.ceil((bb.sum(1) - ff.sum(1)) / 2) * P1_TRIP).sum())
            acc["p2"] += int(am.sum())
    n = acc["list"]
    fixed, p2 = acc["wb"] * FIXED, acc["p2"] * P2_VISIT
    far_cost = acc["wb"] * FAR_TEST
    print("  %d tiles, %d list entries, %.0f %% walked; per (warp, batch): %.1f surfels pass the box test, a lane passes %.1f %% "
          "of them, the interval cull removes %.1f %%; non-conservative culls: %d" %
          (n_tiles, n, 100.0 * acc["walked"] / n, acc["box"] / acc["wb"], 100.0 * acc["live"] / acc["box"],
           100.0 * acc["far"] / acc["box"], acc["bad"]))
    print("  model warp instructions per pair: box test only %.0f (lane cull %.0f, blends %.0f, fixed %.0f)  ->  with the "
          "interval cull %.0f (lane cull %.0f, interval test %.0f)" %
          ((acc["p1"] + p2 + fixed) / n, acc["p1"] / n, p2 / n, fixed / n, (acc["p1_far"] + far_cost + p2 + fixed) / n,
           acc["p1_far"] / n, far_cost / n))


if __name__ == "__main__":
    main()
