#!/usr/bin/env python
"""Benchmark of the surfel-rasterizer hot path inside VINGS-Mono's mapping step (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config C2]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

metric   mapping keyframe-views/s: one "view" = render -> get_loss -> backward -> score bookkeeping ->
         masked Adam step, exactly the iteration of scripts/gaussian/gaussian_base.py:353-404.
step     one pass over the rank's 8 window views (config C2: 1.2 M surfels, 8-keyframe window at 1241x376,
         RGB+depth loss).  With N ranks every optimizer step takes N distinct views of the window (one per
         rank), all-reduces the fused [P,16] gradient/score/visibility buffer over NCCL and applies the same
         Adam step on every rank, so per-rank work is fixed ("weak") and value = 8*N views / step time.
value    device-resident: window ground truth and poses already in HBM.
e2e      the same step through the public API with HOST inputs: every view's ground truth (rgb, depth,
         depth covariance) and camera matrices are copied from pinned host memory inside the timed
         region and the view's loss is read back.
extras   raster: rasterizer-only fwd+bwd iterations/s (BASELINE.json's first metric), per-stage device times,
         roofline of the dominant kernel, cpu_baseline (the CPU emulation, oracle/, on a bounded sample).

--impl reference runs the *unmodified* reference rasterizer (oracle/_ref, built for sm_100 from
/root/reference by oracle/build_ref.py) through the same mapping step on the same GPU: the reference has
no CPU implementation of this path (rasterize_points.cu:28-29 rejects non-CUDA tensors), and its own
sm_100 build is the number BASELINE.json says to beat.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.join(ROOT, "vings-mono_b200"), os.path.join(ROOT, "tests"), ROOT]

import torch  # noqa: E402

VIEWS_PER_STEP = 8


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="C2")
    ap.add_argument("--surfels", type=int, default=0, help="override P (debug only; makes the line invalid)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--unfused", action="store_true",
                    help="b200 arm: torch autograd + torch loss around the drop-in rasterizer instead of the fused "
                         "mapping blocks (debug / ablation)")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "25"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def mark(self):
        return time.time()

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()

    def summary(self, t0, t1):
        sm, mx, reasons = [], 0, set()
        for t, line in self.rows:
            if t < t0 or t > t1 + 0.2:
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = max(mx, float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


SHAPE_NAMES = {"C1": "small synthetic", "C2": "KITTI-shaped", "C3": "RTG Hotel-shaped indoor",
               "C4": "Hierarchical SmallCity-shaped", "C5": "Waymo-shaped 5-camera"}


def load_module(impl):
    if impl == "reference":
        import util
        mod = util.load_ref()
        if mod is None:
            raise RuntimeError("oracle/_ref is not built (python oracle/build_ref.py needs /root/reference)")
        return mod
    import diff_surfel_rasterization as mod
    mod._load()  # fail loudly if the CUDA library is missing -- there is no fallback
    return mod


def stage_bytes(stage, c):
    """Algorithmic bytes per launch (SURVEY.md section 8d / DESIGN.md section 5)."""
    P, V, R, HW, tiles, n_p = c["P"], c["V"], c["R"], c["HW"], c["tiles"], c["sort_passes"]
    return {
        # inputs 52 + radii, tile count, sort key out per surfel; record 128 + depth 4 + rect 8 per visible surfel
        "preprocess": 64 * P + 140 * V,
        # histogram 4P; pass 0 reads P keys, writes V (key, id); passes 1-2 move 16 B; the last reads 8, writes 4
        "surfel_sort": 8 * P + 52 * V,
        # sorted id, tile count, rect per visible surfel; (tile, id) out and one tile-count RED per pair
        "duplicate": 16 * V + 12 * R,
        # pair sort by tile id: (tile, id) in and out per pass, the last pass writes ids only
        "sort": (16 * (n_p - 1) + 12) * R,
        "ranges": 16 * tiles,
        "blend_fwd": 4 * R + 76 * V + 60 * HW,
        "blend_bwd": 4 * R + 76 * V + 101 * HW + 80 * V,
        "preprocess_bwd": 156 * V + 72 * P,
        "adam": 28 * 13 * V + P,
        # fused mapping blocks: loss = rendered maps 40 + ground truth 20 in, gradients 40 out;
        # activate = 28 in + 28 out; update = bookkeeping 49/surfel + (13 grads + 39 param/moment rw) per visible
        "map_loss": 100 * HW,
        "activate": 56 * P,
        "update": 49 * P + 364 * V,
        "churn": 0,
    }[stage]


def iteration_bytes(c):
    # the reference algorithm's compulsory traffic (64-bit key sort, ref_sort_passes passes)
    return 136 * c["P"] + 464 * c["V"] + (36 + 24 * c["ref_sort_passes"]) * c["R"] + 161 * c["HW"] + 8 * c["tiles"]


def higher_msb(n):
    msb, step = 16, 16
    while step > 1:
        step //= 2
        msb = msb + step if n >> msb else msb - step
    if n >> msb:
        msb += 1
    return msb


def cpu_baseline():
    """The CPU emulation (oracle/surfel_emulation.py, a port: the reference has no CPU path) on a bounded
    sample: config C1 (10 k surfels, one 320x240 view), rasterizer fwd+bwd only."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import surfel_emulation as emu
    import synthetic
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    scene, views = synthetic.make_scene("C1")
    inp = synthetic.view_inputs(scene, views[0])
    gc, ga = synthetic.upstream_grads(inp["W"], inp["H"], 0)
    t = time.time()
    emu.forward_backward(inp, gc, ga)
    dt = time.time() - t
    return {"value": 1.0 / dt, "unit": "views/s", "cores": cores, "kind": "port",
            "sample": "rasterizer fwd+bwd only, config C1 (10k surfels, one 320x240 view, R~46k pairs), 1 iteration "
                      "of the pure-torch CPU emulation; the mapping-step config itself is out of reach on CPU"}


def run(emit):
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference" and rank != 0:
        return 0  # the reference has no multi-GPU path: rank 0 alone measures it
    if args.impl == "reference":
        world = 1
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    pg = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
        pg = dist.group.WORLD

    import synthetic
    import mapper_step
    try:
        mod = load_module(args.impl)
    except RuntimeError as e:
        if args.impl == "reference":
            emit(json.dumps({"impl": "reference", "unavailable": str(e)}))
            return 0
        raise
    cfg = synthetic.CONFIGS[args.config]
    scene, views = synthetic.make_scene(args.config, P=args.surfels or None)
    K = len(views)
    W, H, P = scene["W"], scene["H"], scene["means3D"].shape[0]
    mapper = mapper_step.SurfelMapper(mod, scene, dev, supervision=cfg["sup"], world_size=world, process_group=pg,
                                      fused=False if (args.unfused or args.impl == "reference") else None)

    # window ground truth: pinned host copies (e2e leg) and device copies (value leg)
    gt_host = [{k: v.pin_memory() for k, v in synthetic.make_gt(scene, views[i], i).items()} for i in range(K)]
    gt_dev = [{k: v.to(dev) for k, v in g.items()} for g in gt_host]
    view_dev = [mapper_step.view_to_device(v, dev) for v in views]
    view_host = [dict(viewmatrix=v.viewmatrix.pin_memory(), projmatrix=v.projmatrix.pin_memory(),
                      campos=v.campos.pin_memory(), w2c=torch.linalg.inv(v.c2w).float().pin_memory(),
                      tanfovx=v.tanfovx, tanfovy=v.tanfovy) for v in views]
    loss_host = torch.zeros(VIEWS_PER_STEP, dtype=torch.float32).pin_memory()
    h2d_view = sum(t.numel() * t.element_size() for t in gt_host[0].values()) + (16 + 16 + 3 + 16) * 4

    def step_resident(s):
        for i in range(VIEWS_PER_STEP):
            v = (s * VIEWS_PER_STEP + i + rank) % K
            mapper.train_view(view_dev[v], gt_dev[v], kf_id=v)

    copy_stream = torch.cuda.Stream(device=dev)

    def stage(v):
        """Queue the host->device copies of one view's inputs on the copy stream (pinned source, so they run while
        the previous view computes); returns the device tensors and the event the compute stream waits for."""
        with torch.cuda.stream(copy_stream):
            gt = {k: t.to(dev, non_blocking=True) for k, t in gt_host[v].items()}
            vw = {k: (t.to(dev, non_blocking=True) if isinstance(t, torch.Tensor) else t) for k, t in view_host[v].items()}
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        return gt, vw, ev

    def step_e2e(s):
        cur = torch.cuda.current_stream()
        ids = [(s * VIEWS_PER_STEP + i + rank) % K for i in range(VIEWS_PER_STEP)]
        copy_stream.wait_stream(cur)       # every copy of the step starts inside the step
        nxt = stage(ids[0])
        for i in range(VIEWS_PER_STEP):
            gt, vw, ev = nxt
            if i + 1 < VIEWS_PER_STEP:
                nxt = stage(ids[i + 1])    # prefetch the next view of this step behind the current one
            cur.wait_event(ev)
            for t in list(gt.values()) + [t for t in vw.values() if isinstance(t, torch.Tensor)]:
                t.record_stream(cur)
            loss = mapper.train_view(vw, gt, kf_id=ids[i])
            loss_host[i:i + 1].copy_(loss.reshape(1), non_blocking=True)
        cur.synchronize()  # the step's result is on the host

    def barrier():
        if world > 1:
            torch.distributed.barrier()

    # every timed pass starts from the same map: parameters, Adam moments and score bookkeeping are put back first, so
    # `value`, the instrumented repeat, `e2e` and the rasterizer-only loop all measure the same work
    book_names = ("_local_scores", "_global_scores", "_stable_mask", "_globalkf_id", "_globalkf_max_scores")
    snap = {k: getattr(mapper, k).detach().clone() for k in mapper_step.GROUPS + book_names}

    @torch.no_grad()
    def restore():
        for k, t in snap.items():
            cur = getattr(mapper, k)
            (cur.data if isinstance(cur, torch.nn.Parameter) else cur).copy_(t)
        for k in mapper_step.GROUPS:
            st = mapper.optimizer.state.get(getattr(mapper, k))
            if st:
                st["exp_avg"].zero_()
                st["exp_avg_sq"].zero_()

    def timed(fn, steps, warmup):
        restore()
        for s in range(warmup):
            fn(s)
        torch.cuda.synchronize()
        barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.time()
        e0.record()
        for s in range(steps):
            fn(warmup + s)
        e1.record()
        torch.cuda.synchronize()
        t1 = time.time()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
        return float(ms[0]), t0, t1

    clocks = ClockSampler(local_rank) if rank == 0 else None
    mine = args.impl == "b200"
    l0 = mod.launch_count() if mine else 0
    ms_val, t0, t1 = timed(step_resident, args.steps, args.warmup)
    launches = (mod.launch_count() - l0) if mine else 0
    # warm-up launches are in the counter too; scale launches to the timed steps only
    launches = launches * args.steps // (args.steps + args.warmup)
    # per-stage device times (roofline): the same K steps once more with the library's stage events switched on
    # (two event records per stage cost ~1.5 % of a view, so they stay out of the pass `value` is taken from)
    R_seen = []
    stage_ms = {}
    if mine:
        orig_train = mapper.train_view

        def train_and_count(*a, **k):
            out = orig_train(*a, **k)
            R_seen.append(mod.last_counters.get("R", 0))
            return out
        mapper.train_view = train_and_count
        mod.profile_enable(True)
        mod.profile_read()
        timed(step_resident, args.steps, args.warmup)
        stage_ms = mod.profile_read()
        mod.profile_enable(False)
        mapper.train_view = orig_train
    R_timed = R_seen
    ms_e2e, _, t2 = timed(step_e2e, args.steps, args.warmup)
    clk = clocks.summary(t0, t2) if clocks else None
    if clocks:
        clocks.stop()

    views_per_step = VIEWS_PER_STEP * world
    value = views_per_step * args.steps / (ms_val / 1e3)
    e2e = views_per_step * args.steps / (ms_e2e / 1e3)
    out = {
        "metric": "mapping keyframe-views/s (render + loss + backward + score bookkeeping + masked Adam per view)",
        "value": value, "unit": "views/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_val / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "%s: %s mapping step, %d surfels, %d-keyframe window at %dx%d, %s loss; "
                               "one step = %d views per rank" % (args.config, SHAPE_NAMES.get(args.config, "synthetic"),
                                                                 P, K, W, H, cfg["sup"], VIEWS_PER_STEP),
                   "mapping_step": "fused (activate, rasterizer, map_loss, surfel_update kernels)" if mapper.fused
                                   else "torch autograd + torch loss around the rasterizer module",
                   "surfels": P, "window": K, "width": W, "height": H, "views_per_step_per_rank": VIEWS_PER_STEP,
                   "parallelism": "window views partitioned over %d rank(s), surfels replicated, fused [P,16] "
                                  "gradient all-reduce (NCCL)" % world if world > 1 else "single GPU",
                   "l2": "no explicit flush: one view touches %.0f MB of surfel parameters, Adam state, workspace "
                         "and gradients (> 126 MB L2) and consecutive views differ" % ((40 + 104 + 100 + 80 + 120) * P / 1e6)},
        "e2e": {"value": e2e, "unit": "views/s", "h2d_bytes_per_step": h2d_view * VIEWS_PER_STEP,
                "d2h_bytes_per_step": (4 + 8) * VIEWS_PER_STEP,  # loss + the rasterizer's own pair-count read-back
                "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": launches, "clocks": clk,
    }
    out["config"]["e2e_inputs"] = ("pinned host -> device copies of every view's ground truth and camera on a copy stream, "
                                   "the next view of the step prefetched behind the current one; loss read back per view")
    out["config"]["stage_timing"] = "per-stage CUDA events from an instrumented repeat of the same steps, not from the pass `value` times"
    if args.impl == "reference":
        out["impl"] = "reference"
        out["cpu_baseline"] = {"value": value, "unit": "views/s", "cores": os.cpu_count(), "kind": "reference",
                               "sample": "the unmodified reference rasterizer (oracle/_ref, sm_100 build of "
                                         "submodules/diff-surfel-rasterization) through the same mapping step on the "
                                         "same GPU; the reference has no CPU implementation of this path"}

    # ---- rasterizer-only fwd+bwd iterations/s (BASELINE.json's first metric), same window views, the initial map ----
    restore()
    import util
    g_color, g_allmap = synthetic.upstream_grads(W, H, cfg["seed"])
    g_color, g_allmap = g_color.to(dev), g_allmap.to(dev)
    with torch.no_grad():
        act = dict(means3D=mapper._xyz.detach(), colors=mapper._rgb.detach(), opacities=torch.sigmoid(mapper._opacity),
                   scales=torch.exp(mapper._scaling), rotations=torch.nn.functional.normalize(mapper._rotation))
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in act.items()}
    mask = torch.ones(H * W, dtype=torch.bool, device=dev)

    def raster_iter(v):
        vw = view_dev[v % K]
        st = mod.GaussianRasterizationSettings(
            image_height=H, image_width=W, tanfovx=vw["tanfovx"], tanfovy=vw["tanfovy"], bg=mapper.bg,
            scale_modifier=1.0, viewmatrix=vw["viewmatrix"], projmatrix=vw["projmatrix"], sh_degree=0,
            campos=vw["campos"], prefiltered=False, debug=False, pixel_mask=mask)
        m2d = torch.zeros(P, 3, device=dev, requires_grad=True)
        sc = torch.zeros(P, 2, device=dev, requires_grad=True)
        color, radii, allmap = mod.GaussianRasterizer(raster_settings=st)(
            means3D=leaves["means3D"], means2D=m2d, opacities=leaves["opacities"], shs=None,
            colors_precomp=leaves["colors"], scales=leaves["scales"], rotations=leaves["rotations"],
            cov3D_precomp=None, scores=sc)
        torch.autograd.backward([color, allmap], [g_color, g_allmap])
        for t in leaves.values():
            t.grad = None
        return radii

    V_seen = []
    for v in range(K):  # warm-up pass; also collects the visible-surfel count per view (untimed)
        V_seen.append(int((raster_iter(v) > 0).sum()))
    torch.cuda.synchronize()
    n_it = 3 * K
    if mine:
        mod.profile_enable(True)
        mod.profile_read()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for v in range(n_it):
        raster_iter(v)
    e1.record()
    torch.cuda.synchronize()
    raster_ms = e0.elapsed_time(e1) / n_it
    raster_stage = mod.profile_read() if mine else {}
    if mine:
        mod.profile_enable(False)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    counters = {"P": P, "V": sum(V_seen) / len(V_seen), "HW": W * H, "tiles": tiles,
                "R": (sum(R_timed) / len(R_timed)) if R_timed else None,
                "sort_passes": (higher_msb(tiles) + 7) // 8, "ref_sort_passes": (32 + higher_msb(tiles) + 7) // 8}
    out["raster"] = {"metric": "rasterizer fwd+bwd iterations/s (one view, fixed upstream gradients)",
                     "value": 1e3 / raster_ms, "unit": "iters/s", "ms_per_iter": raster_ms}
    out["counters"] = counters

    if mine and counters["R"]:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "MEASURED_PEAKS.json hbm_gbs (of measured)"
        else:
            peak, peak_src = 6650.0, "B200_PROFILING.md fallback (of fallback)"
        stages = {}
        for name, (ms, cnt) in stage_ms.items():
            if cnt:
                # from the instrumented repeat of the timed steps (see above)
                per = ms / cnt
                stages[name] = {"ms_per_launch": per, "launches": cnt,
                                "GBps": stage_bytes(name, counters) / (per * 1e-3) / 1e9}
        dom = max((n for n in stages if n != "adam"), key=lambda n: stages[n]["ms_per_launch"])
        traffic, winst = None, None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            tj = json.load(open(tpath))
            traffic = tj.get(args.config, {}).get(dom)
            winst = tj.get(args.config + "_warp_inst", {}).get(dom)
        out["roofline"] = {"bound": "hbm", "kernel": dom, "achieved": stages[dom]["GBps"], "peak": peak,
                           "unit": "GB/s", "frac": stages[dom]["GBps"] / peak, "traffic": traffic,
                           "peak_source": peak_src,
                           "algorithmic_bytes_per_launch": stage_bytes(dom, counters)}
        if winst:
            # the blend kernels are bound by instruction issue, not HBM: executed warp
            # instructions per launch (ncu capture, profiles/traffic.json) / live launch time vs 148 SMs x 4 schedulers
            sm_mhz = (clk or {}).get("sm_mhz") or 1965.0
            peak_issue = 148 * 4 * sm_mhz * 1e6
            out["roofline"]["issue"] = {"warp_inst_per_launch": winst, "achieved_ginst_s": winst / (stages[dom]["ms_per_launch"] * 1e-3) / 1e9,
                                        "peak_ginst_s": peak_issue / 1e9,
                                        "frac": winst / (stages[dom]["ms_per_launch"] * 1e-3) / peak_issue}
        out["stages"] = stages
        a_iter = iteration_bytes(counters)
        out["raster"]["algorithmic_bytes_per_iter"] = a_iter
        out["raster"]["hbm_frac"] = a_iter / (raster_ms * 1e-3) / 1e9 / peak
        out["raster"]["stages_ms"] = {n: ms / cnt for n, (ms, cnt) in raster_stage.items() if cnt}
    if mine and rank == 0 and world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline()
    if rank == 0:
        emit(json.dumps(out))
    if world > 1:
        torch.distributed.destroy_process_group()
    return 0


def main():
    """stdout carries the ONE JSON line and nothing else: while the bench runs, file descriptor 1 points at stderr, so
    what libraries print there (NCCL's version banner on communicator creation) cannot precede the line."""
    lines = []
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        rc = run(lines.append)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    for ln in lines:
        print(ln, flush=True)
    return rc


if __name__ == "__main__":
    sys.exit(main())
