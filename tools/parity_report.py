"""Diagnostic (GPU box): this package vs the reference build on seeded synthetic configs.
Usage: python tools/parity_report.py C1 [P] [view]   -- prints bit-exactness of the integer stages and
error statistics of images / gradients, plus the reference's own run-to-run jitter."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "vings-mono_b200"), os.path.join(ROOT, "tests")]
import numpy as np
import torch

import synthetic
import util


def stats(name, a, b):
    a, b = torch.as_tensor(a).double(), torch.as_tensor(b).double()
    d = (a - b).abs()
    denom = b.abs().sum().clamp(min=1e-30)
    print("  %-14s max|d| %.3e  sum|d|/sum|ref| %.3e  max|ref| %.3e  nbad(>1e-4*max) %d" % (
        name, float(d.max()) if d.numel() else 0, float(d.sum() / denom), float(b.abs().max()) if b.numel() else 0,
        int((d > 1e-4 * b.abs().max()).sum()) if d.numel() else 0))


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "C1"
    P = int(sys.argv[2]) if len(sys.argv) > 2 and int(sys.argv[2]) > 0 else None
    vi = int(sys.argv[3]) if len(sys.argv) > 3 else -1
    scene, views = synthetic.make_scene(name, P=P)
    inp = util.cuda_inputs(synthetic.view_inputs(scene, views[vi]))
    seed = synthetic.CONFIGS[name]["seed"]
    gc, ga = synthetic.upstream_grads(inp["W"], inp["H"], seed)
    mine, ref = util.load_mine(), util.load_ref()
    print("config", name, "P", scene["means3D"].shape[0], "W,H", inp["W"], inp["H"], "ref:", ref is not None)
    mi = util.mine_intermediates(mine, inp)
    print("mine: R", mi["R"], "V", int(mi["visible"].sum()), "max n_contrib", int(mi["n_contrib"][0].max()),
          "mean n_contrib", float(mi["n_contrib"][0].mean()))
    if ref is not None:
        ri = util.ref_intermediates(ref, inp)
        print("ref:  R", ri["R"], "B", ri["B"])
        vis = ri["radii"] > 0
        for k in ("radii", "tiles_touched"):
            print("  exact %-14s %s  (mismatch %d)" % (k, bool((mi[k] == ri[k]).all()), int((mi[k] != ri[k]).sum())))
        for k in ("depths", "means2D", "transMat", "normal_opacity"):
            a, b = mi[k][vis].view(np.uint32), ri[k][vis].view(np.uint32)
            print("  bits  %-14s %s  (mismatch %d of %d)" % (k, bool((a == b).all()), int((a != b).sum()), a.size))
        if mi["R"] == ri["R"]:
            for k in ("keys", "point_list", "ranges"):
                print("  exact %-14s %s  (mismatch %d)" % (k, bool((mi[k] == ri[k]).all()), int((mi[k] != ri[k]).sum())))
        print("  exact n_contrib     ", bool((mi["n_contrib"][0] == ri["n_contrib"][0]).all()),
              int((mi["n_contrib"][0] != ri["n_contrib"][0]).sum()), " median:",
              int((mi["n_contrib"][1] != ri["n_contrib"][1]).sum()))
        stats("color", mi["color"], ri["color"])
        for c, nm in enumerate(("depth", "alpha", "nx", "ny", "nz", "median", "dist")):
            stats("allmap." + nm, mi["allmap"][c], ri["allmap"][c])
        stats("final_T", mi["final_T"], ri["final_T"])
        print("  color bit-identical pixels: %.6f" % float((mi["color"] == ri["color"]).float().mean()))
    t0 = time.time()
    om = util.run_impl(mine, inp, gc, ga)
    print("mine fwd+bwd wall %.3fs" % (time.time() - t0))
    if ref is not None:
        o1 = util.run_impl(ref, inp, gc, ga)
        o2 = util.run_impl(ref, inp, gc, ga)
        print("reference run-to-run jitter (atomic order):")
        for k in o1:
            if k.startswith("d_"):
                stats(k, o2[k], o1[k])
        print("mine vs reference:")
        for k in o1:
            stats(k, om[k], o1[k])
        om2 = util.run_impl(mine, inp, gc, ga)
        print("mine run-to-run jitter:")
        for k in o1:
            if k.startswith("d_"):
                stats(k, om2[k], om[k])


if __name__ == "__main__":
    main()
