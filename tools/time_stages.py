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
    rounds = int(sys.argv[2]) if len(sys.argv) > 2 and sys.argv[2].isdigit() else 3
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
    if "--per-view" in sys.argv:   # how much the window's views differ (rank skew of the multi-GPU step)
        for vi, inp in enumerate(inps):
            mine.profile_read()
            for _ in range(rounds):
                util.run_impl(mine, inp, gc, ga)
            st = mine.profile_read()
            print("view %d: R %d  sum %.4f ms  (fwd %.4f bwd %.4f)" % (vi, mine.last_counters["R"], sum(ms / n for ms, n in st.values() if n),
                                                                      st["blend_fwd"][0] / st["blend_fwd"][1], st["blend_bwd"][0] / st["blend_bwd"][1]))


if __name__ == "__main__":
    main()
