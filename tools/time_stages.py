"""Per-stage device times of the rasterizer (fwd+bwd) over the window views of a config, for A/B runs of kernel
variants:  SB200_LIB=.../libsurfel_b200_<variant>.so python tools/time_stages.py C2 [rounds]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "vings-mono_b200"), os.path.join(ROOT, "tests")]
import torch

import synthetic
import util


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "C2"
    rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    scene, views = synthetic.make_scene(name)
    mine = util.load_mine()
    gc, ga = synthetic.upstream_grads(scene["W"], scene["H"], synthetic.CONFIGS[name]["seed"])
    inps = [util.cuda_inputs(synthetic.view_inputs(scene, v)) for v in views]
    for inp in inps:  # warm-up
        util.run_impl(mine, inp, gc, ga)
    torch.cuda.synchronize()
    mine.profile_enable(True)
    mine.profile_read()
    for _ in range(rounds):
        for inp in inps:
            util.run_impl(mine, inp, gc, ga)
    st = mine.profile_read()
    tot = 0.0
    out = []
    for k, (ms, n) in st.items():
        if n:
            out.append("%s %.4f" % (k, ms / n))
            tot += ms / n
    print(os.environ.get("SB200_LIB", "default").split("/")[-1], "|", " ".join(out), "| sum %.4f" % tot)


if __name__ == "__main__":
    main()
