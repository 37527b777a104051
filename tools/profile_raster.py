"""Profiling driver (GPU box): a few rasterizer fwd+bwd iterations on one window view of a config, for ncu.
Usage: python tools/profile_raster.py C2 [iters] [view]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "vings-mono_b200"), os.path.join(ROOT, "tests")]
import torch

import synthetic
import util


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "C2"
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    vi = int(sys.argv[3]) if len(sys.argv) > 3 else -1
    scene, views = synthetic.make_scene(name)
    inp = util.cuda_inputs(synthetic.view_inputs(scene, views[vi]))
    gc, ga = synthetic.upstream_grads(inp["W"], inp["H"], synthetic.CONFIGS[name]["seed"])
    mine = util.load_mine()
    for _ in range(iters):
        out = util.run_impl(mine, inp, gc, ga)
    print("R", mine.last_counters["R"], "V", int((out["radii"] > 0).sum()))


if __name__ == "__main__":
    main()
