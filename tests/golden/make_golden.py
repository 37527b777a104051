"""Generates the golden fixtures of tests/golden/*.npz by running the UNMODIFIED reference rasterizer
(oracle/_ref, sm_100 build of /root/reference/submodules/diff-surfel-rasterization) on a B200:

    gpurun -- python tests/golden/make_golden.py        # writes gpurun_out/golden/*.npz; copy to tests/golden/

Each fixture stores the seeded inputs and everything the reference produced for them: radii, per-surfel
intermediates, unsorted/sorted keys, sorted ids, tile ranges, contributor counts, images, its own
max_contrib array, all gradients and scores.  The reference has no tests or vectors of its own
(SURVEY.md section 4), so these runs are what pins the oracle (oracle/surfel_emulation.py) and the kernels."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "vings-mono_b200"), os.path.join(ROOT, "tests")]
import numpy as np
import torch

import synthetic
import util

CASES = {
    # name: (config, P, W, H, bg, mask fraction kept)
    "outdoor_small": ("C1", 1500, 96, 64, (0.0, 0.0, 0.0), 1.0),
    "indoor_ragged": ("C3", 2500, 120, 72, (0.2, 0.4, 0.6), 0.7),
}


def case_inputs(name):
    cfg, P, W, H, bg, keep = CASES[name]
    scene, views = synthetic.make_scene(cfg, P=P, W=W, H=H)
    inp = synthetic.view_inputs(scene, views[-1])
    inp["bg"] = torch.tensor(bg, dtype=torch.float32)
    gen = torch.Generator(device="cpu").manual_seed(99)
    mask = torch.rand(W * H, generator=gen) < keep
    gc, ga = synthetic.upstream_grads(W, H, 7)
    return inp, mask, gc, ga


def main():
    out_dir = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out_dir, exist_ok=True)
    ref = util.load_ref()
    assert ref is not None, "oracle/_ref is not built"
    for name in CASES:
        inp_cpu, mask, gc, ga = case_inputs(name)
        inp = util.cuda_inputs(inp_cpu)
        ri = util.ref_intermediates(ref, inp)
        out = util.run_impl(ref, inp, gc, ga, mask.cuda())
        ids, nbad = util.tainted_surfels(ri, inp["W"], inp["H"])
        full = name != "outdoor_c1"   # the big case keeps only what the CPU suite checks cheaply
        d = dict(W=inp["W"], H=inp["H"], tanfovx=inp["tanfovx"], tanfovy=inp["tanfovy"], mask=mask.numpy(),
                 radii=ri["radii"], tiles_touched=ri["tiles_touched"], R=ri["R"], B=ri["B"],
                 ranges=ri["ranges"], max_contrib=ri["max_contrib"], tainted=ids,
                 keys_crc=np.array([int(np.bitwise_xor.reduce(ri["keys"] * np.uint64(0x9E3779B97F4A7C15)))], dtype=np.uint64),
                 point_list_crc=np.array([int(np.bitwise_xor.reduce(
                     (ri["point_list"].astype(np.uint64) + np.arange(ri["R"], dtype=np.uint64) * np.uint64(1000003))
                     * np.uint64(0x9E3779B97F4A7C15)))], dtype=np.uint64))
        for k in ("means3D", "scales", "rotations", "opacities", "colors", "bg", "viewmatrix", "projmatrix", "campos"):
            d["in_" + k] = inp_cpu[k].numpy()
        d["g_color"], d["g_allmap"] = gc.numpy(), ga.numpy()
        if full:
            for k in ("depths", "means2D", "transMat", "normal_opacity", "keys", "keys_unsorted", "vals_unsorted",
                      "point_list", "n_contrib", "final_T"):
                d[k] = ri[k]
            d["color"], d["allmap"] = out["color"].numpy(), out["allmap"].numpy()
        else:
            d["color"] = out["color"].numpy().astype(np.float16)
            d["n_contrib"] = ri["n_contrib"][0].astype(np.uint16)
        for k in util.GRAD_NAMES:
            d["d_" + k] = out["d_" + k].numpy()
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), **d)
        print(name, "R", ri["R"], "B", ri["B"], "tainted", len(ids), "tiles", nbad,
              "size", os.path.getsize(os.path.join(out_dir, name + ".npz")))


if __name__ == "__main__":
    main()
