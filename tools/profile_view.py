"""Profiling driver (GPU box): ONE fused mapping view of a config between cudaProfilerStart/Stop, after warm-up views,
so that `ncu --profile-from-start off` captures exactly the kernels of one view in order (activate, preprocess,
surfel sort, pair emission, tile scan, pair sort, forward blend, map_loss, backward blend, backward preprocess, update).

    ncu --set full --import-source on --clock-control none --profile-from-start off -o gpurun_out/view_C2 \
        python tools/profile_view.py C2
    python tools/traffic_from_ncu.py gpurun_out/view_C2.ncu-rep C2 "$(cat gpurun_out/view_C2.csrc_hash)" > profiles/traffic.json
Writes gpurun_out/view_<cfg>.csrc_hash (the content hash of csrc/ the capture was made from)."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "vings-mono_b200")]
import torch

import mapper_step
import synthetic


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "C2"
    warm = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    import diff_surfel_rasterization as mod
    mod._load()
    dev = torch.device("cuda")
    scene, views = synthetic.make_scene(name)
    cfg = synthetic.CONFIGS[name]
    m = mapper_step.SurfelMapper(mod, scene, dev, supervision=cfg["sup"], fused=True)
    gts = [{k: t.to(dev) for k, t in synthetic.make_gt(scene, v, i).items()} for i, v in enumerate(views)]
    vds = [mapper_step.view_to_device(v, dev) for v in views]
    for i in range(warm):
        m.train_view(vds[i % len(views)], gts[i % len(views)], kf_id=i % len(views))
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    v = warm % len(views)
    m.train_view(vds[v], gts[v], kf_id=v)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
    spec = importlib.util.spec_from_file_location("sb200_build", os.path.join(ROOT, "vings-mono_b200", "build.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "view_%s.csrc_hash" % name), "w") as f:
        f.write(b.csrc_hash())
    print("R", m._ws.R_seen, "V", int((m._ws.radii > 0).sum()), "P", m.P)


if __name__ == "__main__":
    main()
