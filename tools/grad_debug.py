"""Diagnostic (GPU box): per-surfel gradient differences vs the reference for one config; prints the worst surfels."""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "vings-mono_b200"), os.path.join(ROOT, "tests")]
import numpy as np
import torch
import synthetic
import util

name = sys.argv[1] if len(sys.argv) > 1 else "C1"
scene, views = synthetic.make_scene(name)
inp = util.cuda_inputs(synthetic.view_inputs(scene, views[-1]))
gc, ga = synthetic.upstream_grads(inp["W"], inp["H"], synthetic.CONFIGS[name]["seed"])
mine, ref = util.load_mine(), util.load_ref()
om = util.run_impl(mine, inp, gc, ga)
orf = util.run_impl(ref, inp, gc, ga)
mi = util.mine_intermediates(mine, inp)
d = (om["d_scores"] - orf["d_scores"]).abs()
print("lib", os.environ.get("SB200_LIB", "default"), "score0 max diff", float(d[:, 0].max()), "n>1e-3", int((d[:, 0] > 1e-3).sum()),
      "rel means3D", util.rel_err(om["d_means3D"], orf["d_means3D"]))
worst = torch.argsort(d[:, 0], descending=True)[:6]
for i in worst.tolist():
    print(i, "score mine/ref", om["d_scores"][i].tolist(), orf["d_scores"][i].tolist(), "radius", int(mi["radii"][i]),
          "rect", mi["rect"][i].tolist(), "xy", mi["means2D"][i].tolist(), "tau", float(mi["cull_tau"][i]),
          "bbox", mi["bbox"][i].tolist(), "opa", float(inp["opacities"][i]))
if "--emu" in sys.argv:
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import surfel_emulation as emu
    inp_cpu = synthetic.view_inputs(scene, views[-1])
    pre, bins, fwd, grads = emu.forward_backward(inp_cpu, gc, ga)
    for i in worst.tolist():
        print(i, "emu score", grads["scores"][i].tolist())
    print("emu vs ref rel", util.rel_err(grads["scores"], orf["d_scores"]), "emu vs mine", util.rel_err(grads["scores"], om["d_scores"]))
ri = util.ref_intermediates(ref, inp)
ids, nbad = util.tainted_surfels(ri, inp["W"], inp["H"])
print("reference max_contrib bug: affected tiles", nbad, "tainted surfels", len(ids))
keep = torch.ones(om["d_scores"].shape[0], dtype=torch.bool)
keep[torch.from_numpy(ids)] = False
for k in ("d_means3D", "d_means2D", "d_colors", "d_opacities", "d_scales", "d_rotations", "d_scores"):
    print("  untainted %-12s rel %.3e  maxabs %.3e" % (k, util.rel_err(om[k][keep], orf[k][keep]),
                                                  float((om[k][keep] - orf[k][keep]).abs().max())))
