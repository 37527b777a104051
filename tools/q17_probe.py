"""Diagnostic for reference quirk Q17: on which surfels, and in which direction, does the repo's
backward -- uncapped, and capped at the reference's stored max_contrib -- differ from the reference's?
Usage: python tools/q17_probe.py C2 [P]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "vings-mono_b200"), os.path.join(ROOT, "tests")]
import numpy as np
import torch

import synthetic
import util


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "C2"
    P = int(sys.argv[2]) if len(sys.argv) > 2 else None
    mine, ref = util.load_mine(), util.load_ref()
    scene, views = synthetic.make_scene(name, P=P)
    inp = util.cuda_inputs(synthetic.view_inputs(scene, views[-1]))
    W, H = inp["W"], inp["H"]
    gc_, ga = synthetic.upstream_grads(W, H, synthetic.CONFIGS[name]["seed"])
    mask = None
    orf = util.run_impl(ref, inp, gc_, ga, mask, ref_state=True)
    rs = orf.pop("state")
    orf2 = util.run_impl(ref, inp, gc_, ga, mask, ref_state=True)
    rs2 = orf2.pop("state")
    print("max_contrib equal across two reference forwards:", np.array_equal(rs["max_contrib"], rs2["max_contrib"]),
          " reference run-to-run d_scores rel:", util.rel_err(orf2["d_scores"], orf["d_scores"]))
    om = util.run_impl(mine, inp, gc_, ga, mask)
    ids, nbad = util.tainted_surfels(rs, W, H)
    cap = torch.from_numpy(rs["max_contrib"].astype(np.int64)).to(torch.int32).cuda()
    mine.debug_walk_cap(cap)
    oc = util.run_impl(mine, inp, gc_, ga, mask)
    mine.debug_walk_cap(None)
    t = torch.from_numpy(ids)
    keep = torch.ones(om["d_scores"].shape[0], dtype=torch.bool)
    keep[t] = False
    print("%s: tainted %d surfels in %d bad tiles" % (name, len(ids), nbad))
    for k in ("d_scores", "d_colors", "d_opacities", "d_means3D"):
        a, b, c = orf[k], om[k], oc[k]
        print("%-12s mass(tainted)/mass(all) %.3e | keep: uncapped %.2e capped %.2e | tainted: uncapped %.2e capped %.2e"
              % (k, float(a[t].abs().sum() / a.abs().sum()), util.rel_err(b[keep], a[keep]), util.rel_err(c[keep], a[keep]),
                 util.rel_err(b[t], a[t]), util.rel_err(c[t], a[t])))
    # score 0 = sum over pixels of alpha * T: a plain positive sum -- who counts more?
    s_ref, s_cap, s_unc = orf["d_scores"][t, 0].double(), oc["d_scores"][t, 0].double(), om["d_scores"][t, 0].double()
    d = s_cap - s_ref
    print("score0 on tainted: sum ref %.6e capped %.6e uncapped %.6e; capped-ref: %d larger, %d smaller, %d equal(1e-6 rel)"
          % (float(s_ref.sum()), float(s_cap.sum()), float(s_unc.sum()), int((d > 1e-6 * s_ref.abs()).sum()),
             int((d < -1e-6 * s_ref.abs()).sum()), int((d.abs() <= 1e-6 * s_ref.abs()).sum())))
    order = torch.argsort(d.abs(), descending=True)[:12]
    # for the worst ones: in which tiles do they sit, at which list position, and what are that tile's numbers?
    ranges, pl, mc = rs["ranges"].astype(np.int64), rs["point_list"], rs["max_contrib"].astype(np.int64)
    gx = (W + 15) // 16
    last = np.zeros((((H + 15) // 16) * 16, gx * 16), np.int64)
    last[:H, :W] = rs["n_contrib"][0].reshape(H, W)
    true_max = last.reshape((H + 15) // 16, 16, gx, 16).max(axis=(1, 3)).reshape(-1)
    tile_of = np.zeros(len(pl), np.int64)
    for tl in np.nonzero(ranges[:, 1] > ranges[:, 0])[0]:
        tile_of[ranges[tl, 0]:ranges[tl, 1]] = tl
    for o in order.tolist():
        sid = int(ids[o])
        pos = np.nonzero(pl == sid)[0]
        desc = []
        for p_ in pos[:8]:
            tl = tile_of[p_]
            desc.append("tile %d pos %d/%d mc %d true %d" % (tl, p_ - ranges[tl, 0], ranges[tl, 1] - ranges[tl, 0], mc[tl], true_max[tl]))
        print("surfel %d: score0 ref %.4e capped %.4e uncapped %.4e | %s" % (sid, float(s_ref[o]), float(s_cap[o]), float(s_unc[o]), "; ".join(desc)))


if __name__ == "__main__":
    main()
