"""Why is the backward blend slower inside the mapping step than in the rasterizer-only loop?  (VERDICT r1, item 3 iv)

Times the blend kernels (library stage events) in three settings on config C2:
  A  rasterizer-only loop on the INITIAL map            (what bench.py's `raster` leg and tools/time_stages.py measure)
  B  the mapping step itself                             (what bench.py's `stages` report)
  C  rasterizer-only loop on the map as the mapping steps LEFT it (same surfels, trained parameters)
and prints the pair count R and the number of (warp, surfel) visits the forward's hit masks hold for each.
If C ~ B the gap is the workload (the map trains: opacities and scales move, more pairs contribute); if C ~ A it
is the kernel's surroundings (cache state, co-running work).
Usage: python tools/insitu_probe.py [C2] [steps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "vings-mono_b200"), os.path.join(ROOT, "tests")]
import torch

import mapper_step
import synthetic
import util


def raster_loop(mine, mapper, views, W, H, seed, rounds=2):
    gc, ga = synthetic.upstream_grads(W, H, seed)
    with torch.no_grad():
        inp0 = dict(means3D=mapper._xyz.detach().clone(), colors=mapper._rgb.detach().clone(),
                    opacities=torch.sigmoid(mapper._opacity.detach()), scales=torch.exp(mapper._scaling.detach()),
                    rotations=torch.nn.functional.normalize(mapper._rotation.detach()), bg=mapper.bg, W=W, H=H)
    inps = []
    for v in views:
        d = dict(inp0)
        d.update(viewmatrix=v.viewmatrix.cuda(), projmatrix=v.projmatrix.cuda(), campos=v.campos.cuda(),
                 tanfovx=v.tanfovx, tanfovy=v.tanfovy)
        inps.append(d)
    for d in inps:
        util.run_impl(mine, d, gc, ga)
    mine.profile_enable(True)
    mine.profile_read()
    Rs = []
    for _ in range(rounds):
        for d in inps:
            util.run_impl(mine, d, gc, ga)
            Rs.append(mine.last_counters["R"])
    st = mine.profile_read()
    mine.profile_enable(False)
    return {k: ms / n for k, (ms, n) in st.items() if n}, sum(Rs) / len(Rs)


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "C2"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    mine = util.load_mine()
    scene, views = synthetic.make_scene(name)
    cfg = synthetic.CONFIGS[name]
    dev = torch.device("cuda")
    W, H = scene["W"], scene["H"]
    m = mapper_step.SurfelMapper(mine, scene, dev, supervision=cfg["sup"], fused=True)
    gts = [{k: t.cuda() for k, t in synthetic.make_gt(scene, v, i).items()} for i, v in enumerate(views)]
    vds = [mapper_step.view_to_device(v, dev) for v in views]
    a, Ra = raster_loop(mine, m, views, W, H, cfg["seed"])
    print("A raster-only, initial map : fwd %.3f bwd %.3f ms  R %.0f" % (a["blend_fwd"], a["blend_bwd"], Ra))
    for s in range(3):   # warm-up steps (also what bench.py does before timing)
        for i in range(len(views)):
            m.train_view(vds[i], gts[i], kf_id=i)
    mine.profile_enable(True)
    mine.profile_read()
    Rs = []
    for s in range(steps):
        for i in range(len(views)):
            m.train_view(vds[i], gts[i], kf_id=i)
            Rs.append(m._ws.R_seen)
    b = {k: ms / n for k, (ms, n) in mine.profile_read().items() if n}
    mine.profile_enable(False)
    print("B mapping step (%d steps)   : fwd %.3f bwd %.3f ms  R %.0f" % (steps, b["blend_fwd"], b["blend_bwd"], sum(Rs) / len(Rs)))
    c, Rc = raster_loop(mine, m, views, W, H, cfg["seed"])
    print("C raster-only, trained map : fwd %.3f bwd %.3f ms  R %.0f" % (c["blend_fwd"], c["blend_bwd"], Rc))


if __name__ == "__main__":
    main()
