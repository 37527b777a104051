"""Diagnostic (GPU box): isolates single pixels with pixel_mask and compares per-(pixel, surfel) scores
(alpha*T as the backward sees it) between this package and the reference."""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "vings-mono_b200"), os.path.join(ROOT, "tests")]
import numpy as np
import torch
import synthetic
import util

name = sys.argv[1]
tx0, ty0, tx1, ty1 = [int(v) for v in sys.argv[2:6]]
scene, views = synthetic.make_scene(name)
inp = util.cuda_inputs(synthetic.view_inputs(scene, views[-1]))
W, H = inp["W"], inp["H"]
gc, ga = synthetic.upstream_grads(W, H, synthetic.CONFIGS[name]["seed"])
mine, ref = util.load_mine(), util.load_ref()
mi = util.mine_intermediates(mine, inp)
shown = 0
for py in range(ty0 * 16, ty1 * 16):
    for px in range(tx0 * 16, tx1 * 16):
        mask = torch.zeros(H * W, dtype=torch.bool, device="cuda")
        mask[py * W + px] = True
        a = util.run_impl(mine, inp, gc, ga, mask)["d_scores"][:, 0]
        b = util.run_impl(ref, inp, gc, ga, mask)["d_scores"][:, 0]
        d = (a - b).abs()
        if float(d.max()) > 1e-4 and shown < 4:
            shown += 1
            tile = (py // 16) * ((W + 15) // 16) + px // 16
            r0, r1 = mi["ranges"][tile]
            plist = mi["point_list"][r0:r1]
            last = int(mi["n_contrib"][0][py * W + px])
            print("pixel", px, py, "tile", tile, "range", r0, r1, "last", last, "final T", float(mi["final_T"][0][py * W + px]))
            for pos, sid in enumerate(plist.tolist()):
                if a[sid] != 0 or b[sid] != 0:
                    print("   pos %4d surfel %6d mine %.6e ref %.6e %s" % (pos, sid, float(a[sid]), float(b[sid]),
                                                                           "<<<" if abs(float(a[sid] - b[sid])) > 1e-5 else ""))
print("done, differing pixels shown:", shown)
