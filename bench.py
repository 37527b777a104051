#!/usr/bin/env python
"""Benchmark of the surfel-rasterizer hot path inside VINGS-Mono's mapping step (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config C2]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

metric   mapping keyframe-views/s: one "view" = render -> get_loss -> backward -> score bookkeeping ->
         masked Adam step, exactly the iteration of scripts/gaussian/gaussian_base.py:353-404.
step     one pass over the rank's 8 window views (config C2: 1.2 M surfels, 8-keyframe window at 1241x376,
         RGB+depth loss).  With N ranks every optimizer step takes N distinct views of the window (one per
         rank), exchanges the per-surfel gradient / score rows over NVLink and applies the same Adam step on every
         rank, so per-rank work is fixed ("weak") and value = 8*N views / step time.
value    device-resident: window ground truth and poses already in HBM.
e2e      the same step through the public API with HOST inputs: every view's ground truth (rgb, depth,
         depth covariance) and camera matrices are copied from pinned host memory inside the timed
         region and the view's loss is read back.
extras   raster   rasterizer-only fwd+bwd iterations/s through the drop-in autograd interface (BASELINE.json's first metric)
         dropin   the mapping step through the reference's OWN boundary: GaussianRasterizer under torch autograd +
                  the caller's torch loss + SparseGaussianAdam (what a VINGS-Mono checkout gets by swapping the module)
         strong   BASELINE.json configs[3]: ONE 16-keyframe window of the 5 M-surfel 1080p scene split over the ranks
                  (16/N optimizer steps per window pass) -- strong scaling of the mapping step
         stages / roofline / counters / cpu_baseline (the CPU emulation, oracle/, on a bounded sample)
--config C5 puts the densify / prune churn inside the timed step: every step prunes a random 1 % of the surfels
and appends 40 000 new ones (scripts/gaussian/gaussian_model.py:92-161), so P varies from step to step.

--impl reference runs the *unmodified* reference rasterizer (oracle/_ref, built for sm_100 from
/root/reference by oracle/build_ref.py) through the same mapping step on the same GPU: the reference has
no CPU implementation of this path (rasterize_points.cu:28-29 rejects non-CUDA tensors), and its own
sm_100 build is the number BASELINE.json says to beat.
"""
import argparse
import importlib.util
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.join(ROOT, "vings-mono_b200"), ROOT]

import torch  # noqa: E402

VIEWS_PER_STEP = 8
CHURN_PRUNE_FRAC, CHURN_ADD = 0.01, 40_000


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="C2")
    ap.add_argument("--surfels", type=int, default=0, help="override P (debug only; makes the line invalid)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the dropin / strong / cpu_baseline legs")
    ap.add_argument("--strong-config", default="C4")
    ap.add_argument("--unfused", action="store_true",
                    help="b200 arm: torch autograd + torch loss around the drop-in rasterizer instead of the fused "
                         "mapping blocks for the headline value too (debug / ablation)")
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
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import ref_loader   # oracle/: only the reference arm goes there
        mod = ref_loader.load_ref()
        if mod is None:
            raise RuntimeError("oracle/_ref is not built (python oracle/build_ref.py needs /root/reference)")
        return mod
    import diff_surfel_rasterization as mod
    mod._load()  # fail loudly if the CUDA library is missing -- there is no fallback
    return mod


def csrc_hash():
    spec = importlib.util.spec_from_file_location("sb200_build", os.path.join(ROOT, "vings-mono_b200", "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.csrc_hash()


def stage_bytes(stage, c):
    """Algorithmic bytes per launch (SURVEY.md section 8d / DESIGN.md section 4)."""
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
        # (+ R bytes of hit masks out / in: one 32-bit word per 32-surfel batch and warp)
        "blend_fwd": 5 * R + 76 * V + 60 * HW,
        "blend_bwd": 5 * R + 76 * V + 101 * HW + 80 * V,
        "preprocess_bwd": 156 * V + 72 * P,
        "adam": 28 * 13 * V + P,
        # fused mapping blocks: loss = rendered maps 40 + ground truth 20 in, gradients 40 out;
        # activate = 28 in + 28 out; update = bookkeeping 49/surfel + (13 grads + 39 param/moment rw) per visible
        "map_loss": 100 * HW,
        "activate": 56 * P,
        "update": 49 * P + 364 * V,
        "churn": 0,
        # multi-GPU exchange (csrc/dp_exchange.cu): per rank, stage rows of the slice's visible surfels in from every
        # rank + result rows out; result rows of every touched surfel in + the Adam traffic of sb_surfel_update
        "dp_reduce": 64 * V + P + (80 * min(P, 1.5 * V) + 9 * P) / max(c.get("world", 1), 1),
        "dp_post": P + (80 + 312) * min(P, 1.5 * V),
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
    """The CPU emulation (oracle/surfel_emulation.py, a port: the reference has no CPU path) on a bounded sample of the
    hot path: rasterizer fwd+bwd on config C1 (the config BASELINE.json names for it), one warm-up + 3 timed iterations."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import surfel_emulation as emu
    import synthetic
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    scene, views = synthetic.make_scene("C1")
    inp = synthetic.view_inputs(scene, views[0])
    gc, ga = synthetic.upstream_grads(inp["W"], inp["H"], 0)
    emu.forward_backward(inp, gc, ga)   # warm-up
    n = 3
    t = time.time()
    for _ in range(n):
        emu.forward_backward(inp, gc, ga)
    dt = (time.time() - t) / n
    return {"value": 1.0 / dt, "unit": "iters/s", "cores": cores, "kind": "port",
            "sample": "rasterizer fwd+bwd iterations (NOT mapping views: compare with raster.value at C1, "
                      "profiles/) on config C1 = 10k surfels, one 320x240 view, R~46k pairs; pure-torch CPU emulation "
                      "(oracle/surfel_emulation.py), 1 warm-up + %d timed iterations on %d host threads; the mapping-step "
                      "configs are out of reach on CPU" % (n, cores)}


class Workload:
    """A config's scene, window views, ground truth (pinned host + device) and its mapper."""

    def __init__(self, name, mod, dev, rank, world, pg, fused, P_override=None):
        import mapper_step
        import synthetic
        self.name, self.dev, self.rank, self.world = name, dev, rank, world
        self.cfg = synthetic.CONFIGS[name]
        self.scene, self.views = synthetic.make_scene(name, P=P_override)
        self.K = len(self.views)
        self.W, self.H, self.P0 = self.scene["W"], self.scene["H"], self.scene["means3D"].shape[0]
        self.mapper = mapper_step.SurfelMapper(mod, self.scene, dev, supervision=self.cfg["sup"], world_size=world,
                                               process_group=pg, fused=fused, rank=rank)
        self.gt_host = [{k: v.pin_memory() for k, v in synthetic.make_gt(self.scene, self.views[i], i).items()}
                        for i in range(self.K)]
        self.gt_dev = [{k: v.to(dev) for k, v in g.items()} for g in self.gt_host]
        self.view_dev = [mapper_step.view_to_device(v, dev) for v in self.views]
        self.view_host = [dict(viewmatrix=v.viewmatrix.pin_memory(), projmatrix=v.projmatrix.pin_memory(),
                               campos=v.campos.pin_memory(), w2c=torch.linalg.inv(v.c2w).float().pin_memory(),
                               tanfovx=v.tanfovx, tanfovy=v.tanfovy) for v in self.views]
        self.h2d_view = sum(t.numel() * t.element_size() for t in self.gt_host[0].values()) + (16 + 16 + 3 + 16) * 4
        book = ("_local_scores", "_global_scores", "_stable_mask", "_globalkf_id", "_globalkf_max_scores")
        self.snap = {k: getattr(self.mapper, k).detach().clone() for k in mapper_step.GROUPS + book}
        self.churn_pool = None

    def ids(self, s, i):
        """Window view of every rank at view slot i of step s (rank order) -- also the keyframe ids of the step."""
        if os.environ.get("SB200_BENCH_SAME_VIEW"):   # diagnostic only (invalid as a bench line): no rank skew
            return [(s * VIEWS_PER_STEP + i) % self.K] * self.world
        return [(s * VIEWS_PER_STEP + i + q) % self.K for q in range(self.world)]

    @torch.no_grad()
    def restore(self):
        """Every timed pass starts from the same map: parameters, Adam moments and score bookkeeping are put back."""
        import mapper_step
        m = self.mapper
        if m.P != self.P0:   # the churn legs changed the surfel count: rebuild storage from the snapshot
            raw = {k: self.snap[k] for k in mapper_step.GROUPS}
            m.set_raw(raw, book={k: v for k, v in self.snap.items() if k not in mapper_step.GROUPS})
            return
        for k, t in self.snap.items():
            cur = getattr(m, k)
            (cur.data if isinstance(cur, torch.nn.Parameter) else cur).copy_(t)
        for k in mapper_step.GROUPS:
            st = m.optimizer.state.get(getattr(m, k))
            if st:
                st["exp_avg"].zero_()
                st["exp_avg_sq"].zero_()

    # ---- churn (C5): prune a random 1 %, append 40 000 new surfels; identical on every rank (seeded) ----
    def prepare_churn(self, n_steps):
        """New surfels for n_steps densifications, drawn the way the cloud itself was (one more source keyframe each)."""
        import synthetic
        extra, _ = synthetic.make_scene(self.name, P=self.P0 + CHURN_ADD * min(n_steps, 8))
        g = torch.Generator(device="cpu").manual_seed(99)
        self.churn_pool = dict(xyz=extra["means3D"][self.P0:].to(self.dev), rgb=extra["colors"][self.P0:].to(self.dev),
                               rot=torch.randn(extra["means3D"].shape[0] - self.P0, 4, generator=g).to(self.dev))
        self.churn_gen = torch.Generator(device=self.dev)

    def churn(self, s):
        m = self.mapper
        self.churn_gen.manual_seed(1234 + s)
        mask = torch.rand(m.P, device=self.dev, generator=self.churn_gen) < CHURN_PRUNE_FRAC
        m.prune(mask)
        n_slots = self.churn_pool["xyz"].shape[0] // CHURN_ADD
        lo = (s % n_slots) * CHURN_ADD
        m.add_surfels(self.churn_pool["xyz"][lo:lo + CHURN_ADD], self.churn_pool["rgb"][lo:lo + CHURN_ADD],
                      self.churn_pool["rot"][lo:lo + CHURN_ADD])


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

    import mapper_step
    import synthetic
    try:
        mod = load_module(args.impl)
    except RuntimeError as e:
        if args.impl == "reference":
            emit(json.dumps({"impl": "reference", "unavailable": str(e)}))
            return 0
        raise
    mine = args.impl == "b200"
    fused = False if (args.unfused or not mine) else None
    wl = Workload(args.config, mod, dev, rank, world, pg, fused, P_override=args.surfels or None)
    mapper, K, W, H, P = wl.mapper, wl.K, wl.W, wl.H, wl.P0
    cfg = wl.cfg
    churn = args.config == "C5"
    if churn:
        wl.prepare_churn(args.steps + args.warmup)
    loss_host = torch.zeros(VIEWS_PER_STEP, dtype=torch.float32).pin_memory()
    P_seen = []

    def step_resident(s):
        if churn:
            wl.churn(s)
            P_seen.append(mapper.P)
        for i in range(VIEWS_PER_STEP):
            ids = wl.ids(s, i)
            mapper.train_view(wl.view_dev[ids[rank]], wl.gt_dev[ids[rank]], kf_id=ids[rank], kf_ids=ids)

    copy_stream = torch.cuda.Stream(device=dev)

    def stage(v):
        """Queue the host->device copies of one view's inputs on the copy stream (pinned source, so they run while
        the previous view computes); returns the device tensors and the event the compute stream waits for."""
        with torch.cuda.stream(copy_stream):
            gt = {k: t.to(dev, non_blocking=True) for k, t in wl.gt_host[v].items()}
            vw = {k: (t.to(dev, non_blocking=True) if isinstance(t, torch.Tensor) else t) for k, t in wl.view_host[v].items()}
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        return gt, vw, ev

    def step_e2e(s):
        cur = torch.cuda.current_stream()
        if churn:
            wl.churn(s)
        all_ids = [wl.ids(s, i) for i in range(VIEWS_PER_STEP)]
        copy_stream.wait_stream(cur)       # every copy of the step starts inside the step
        nxt = stage(all_ids[0][rank])
        for i in range(VIEWS_PER_STEP):
            gt, vw, ev = nxt
            if i + 1 < VIEWS_PER_STEP:
                nxt = stage(all_ids[i + 1][rank])    # prefetch the next view of this step behind the current one
            cur.wait_event(ev)
            for t in list(gt.values()) + [t for t in vw.values() if isinstance(t, torch.Tensor)]:
                t.record_stream(cur)
            loss = mapper.train_view(vw, gt, kf_id=all_ids[i][rank], kf_ids=all_ids[i])
            loss_host[i:i + 1].copy_(loss.reshape(1), non_blocking=True)
        cur.synchronize()  # the step's result is on the host

    def barrier():
        if world > 1:
            torch.distributed.barrier()

    def timed(fn, steps, warmup, restore=wl.restore):
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

    def replicas_identical(m):
        """Bit-level checksum of every parameter, moment and bookkeeping array, compared across the ranks."""
        if world <= 1:
            return None
        sums = []
        for k in mapper_step.GROUPS:
            p = getattr(m, k)
            st = m.optimizer.state[p]
            sums += [p.data, st["exp_avg"], st["exp_avg_sq"]]
        sums += [m._local_scores, m._global_scores, m._globalkf_max_scores, m._globalkf_id]
        cs = torch.stack([t.contiguous().view(torch.uint8).to(torch.int64).sum() * 31 +
                          t.contiguous().view(torch.uint8)[::7].to(torch.int64).sum() for t in sums])
        lo, hi = cs.clone(), cs.clone()
        torch.distributed.all_reduce(lo, op=torch.distributed.ReduceOp.MIN)
        torch.distributed.all_reduce(hi, op=torch.distributed.ReduceOp.MAX)
        return bool(torch.equal(lo, hi))

    clocks = ClockSampler(local_rank) if rank == 0 else None
    l0 = mod.launch_count() if mine else 0
    ms_val, t0, t1 = timed(step_resident, args.steps, args.warmup)
    launches = (mod.launch_count() - l0) if mine else 0
    # warm-up launches are in the counter too; scale launches to the timed steps only
    launches = launches * args.steps // (args.steps + args.warmup)
    same_bits = replicas_identical(mapper)
    if getattr(mapper, "_dpx", None) is not None:
        mapper._dpx.check()   # no kernel of the exchange ever gave up waiting for a peer
    P_timed = P_seen[args.warmup:] if churn else [P]
    # per-stage device times (roofline): the same K steps once more with the library's stage events switched on
    # (two event records per stage cost ~1.5 % of a view, so they stay out of the pass `value` is taken from)
    R_seen = []
    stage_ms = {}
    dp_wait = None
    if mine:
        orig_train = mapper.train_view

        def train_and_count(*a, **k):
            out = orig_train(*a, **k)
            R_seen.append(mod.last_counters.get("R", 0))
            return out
        mapper.train_view = train_and_count
        mod.profile_enable(True)
        mod.profile_read()
        dpx = getattr(mapper, "_dpx", None)
        if dpx is not None:
            dpx.wait_ms()
        timed(step_resident, args.steps, args.warmup)
        stage_ms = mod.profile_read()
        mod.profile_enable(False)
        if dpx is not None:
            w = dpx.wait_ms()
            n_views = VIEWS_PER_STEP * (args.steps + args.warmup)
            dp_wait = {"before_reduce": w[0] / n_views, "before_post": w[1] / n_views}
            # every rank's own numbers: its compute per view (all stages but the exchange's reduce / post), and how long
            # its reduce / post kernels waited for the slowest peer -- tells view / GPU skew from exchange work
            own = sum(ms / cnt for n, (ms, cnt) in stage_ms.items() if cnt and n not in ("dp_reduce", "dp_post"))
            mine_row = torch.tensor([own, dp_wait["before_reduce"], dp_wait["before_post"],
                                     stage_ms["blend_fwd"][0] / max(stage_ms["blend_fwd"][1], 1),
                                     stage_ms["blend_bwd"][0] / max(stage_ms["blend_bwd"][1], 1)], device=dev, dtype=torch.float64)
            rows = [torch.empty_like(mine_row) for _ in range(world)]
            torch.distributed.all_gather(rows, mine_row)
            dp_wait["by_rank"] = {"compute_ms_per_view": [round(float(r[0]), 4) for r in rows],
                                  "wait_before_reduce": [round(float(r[1]), 4) for r in rows],
                                  "wait_before_post": [round(float(r[2]), 4) for r in rows],
                                  "blend_fwd_ms": [round(float(r[3]), 4) for r in rows],
                                  "blend_bwd_ms": [round(float(r[4]), 4) for r in rows]}
        mapper.train_view = orig_train
    R_timed = R_seen
    ms_e2e, _, t2 = timed(step_e2e, args.steps, args.warmup)
    clk = clocks.summary(t0, t2) if clocks else None
    if clocks:
        clocks.stop()

    views_per_step = VIEWS_PER_STEP * world
    value = views_per_step * args.steps / (ms_val / 1e3)
    e2e = views_per_step * args.steps / (ms_e2e / 1e3)
    workload = "%s: %s mapping step, %d surfels, %d-keyframe window at %dx%d, %s loss; one step = %d views per rank" % (
        args.config, SHAPE_NAMES.get(args.config, "synthetic"), P, K, W, H, cfg["sup"], VIEWS_PER_STEP)
    if churn:
        workload += "; every step first prunes a random %.0f %% of the surfels and appends %d new ones (kNN scales)" % (
            100 * CHURN_PRUNE_FRAC, CHURN_ADD)
    out = {
        "metric": "mapping keyframe-views/s (render + loss + backward + score bookkeeping + masked Adam per view)",
        "value": value, "unit": "views/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_val / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload,
                   "mapping_step": "fused (activate, rasterizer, map_loss, surfel_update kernels)" if mapper.fused
                                   else "torch autograd + torch loss around the rasterizer module",
                   "surfels": P, "window": K, "width": W, "height": H, "views_per_step_per_rank": VIEWS_PER_STEP,
                   "parallelism": mapper.parallelism(),
                   "l2": "no explicit flush: one view touches %.0f MB of surfel parameters, Adam state, workspace "
                         "and gradients (> 126 MB L2) and consecutive views differ" % ((40 + 104 + 100 + 80 + 120) * P / 1e6)},
        "e2e": {"value": e2e, "unit": "views/s", "h2d_bytes_per_step": wl.h2d_view * VIEWS_PER_STEP,
                "d2h_bytes_per_step": (4 + 8) * VIEWS_PER_STEP,  # loss + the rasterizer's own pair-count read-back
                "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": launches, "clocks": clk,
    }
    out["config"]["e2e_inputs"] = ("pinned host -> device copies of every view's ground truth and camera on a copy stream, "
                                   "the next view of the step prefetched behind the current one; loss read back per view")
    out["config"]["stage_timing"] = "per-stage CUDA events from an instrumented repeat of the same steps, not from the pass `value` times"
    if churn:
        out["churn"] = {"surfels_per_timed_step": P_timed, "prune_frac": CHURN_PRUNE_FRAC, "added_per_step": CHURN_ADD}
    if same_bits is not None:
        out["replicas_bit_identical"] = same_bits   # parameters, moments, score bookkeeping after the timed steps
    if dp_wait:
        # ms per view rank 0's reduce / post kernels spent WAITING for the slowest peer (rank skew), included in the
        # dp_reduce / dp_post stage times
        out["dp_wait_ms_per_view"] = dp_wait
    if args.impl == "reference":
        out["impl"] = "reference"
        out["cpu_baseline"] = {"value": value, "unit": "views/s", "cores": os.cpu_count(), "kind": "reference",
                               "sample": "the unmodified reference rasterizer (oracle/_ref, sm_100 build of "
                                         "submodules/diff-surfel-rasterization) through the same mapping step on the "
                                         "same GPU; the reference has no CPU implementation of this path"}
        out["dropin"] = {"value": value, "unit": "views/s",
                         "what": "the reference arm IS the drop-in boundary: its module under torch autograd + torch loss"}

    # ---- rasterizer-only fwd+bwd iterations/s (BASELINE.json's first metric), same window views, the initial map ----
    wl.restore()
    g_color, g_allmap = synthetic.upstream_grads(W, H, cfg["seed"])
    g_color, g_allmap = g_color.to(dev), g_allmap.to(dev)
    with torch.no_grad():
        act = dict(means3D=mapper._xyz.detach(), colors=mapper._rgb.detach(), opacities=torch.sigmoid(mapper._opacity),
                   scales=torch.exp(mapper._scaling), rotations=torch.nn.functional.normalize(mapper._rotation))
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in act.items()}
    mask = torch.ones(H * W, dtype=torch.bool, device=dev)

    def raster_iter(v):
        vw = wl.view_dev[v % K]
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
    n_it = max(3 * K, 24)
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
    counters = {"P": P, "V": sum(V_seen) / len(V_seen), "HW": W * H, "tiles": tiles, "world": world,
                "R": (sum(R_timed) / len(R_timed)) if R_timed else None,
                "sort_passes": (higher_msb(tiles) + 7) // 8, "ref_sort_passes": (32 + higher_msb(tiles) + 7) // 8}
    out["raster"] = {"metric": "rasterizer fwd+bwd iterations/s (one view, fixed upstream gradients) through the "
                               "GaussianRasterizer autograd interface", "value": 1e3 / raster_ms, "unit": "iters/s",
                     "ms_per_iter": raster_ms}
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
        dom = max((n for n in stages if n not in ("adam", "churn")), key=lambda n: stages[n]["ms_per_launch"])
        traffic, winst, stale = None, None, None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            tj = json.load(open(tpath))
            traffic = tj.get(args.config, {}).get(dom)
            winst = tj.get(args.config + "_warp_inst", {}).get(dom)
            # the ncu capture describes the kernels only if it was made from the same sources
            stale = tj.get("_csrc_hash") != csrc_hash()
        out["roofline"] = {"bound": "hbm", "kernel": dom, "achieved": stages[dom]["GBps"], "peak": peak,
                           "unit": "GB/s", "frac": stages[dom]["GBps"] / peak, "traffic": traffic,
                           "traffic_stale": stale, "peak_source": peak_src,
                           "algorithmic_bytes_per_launch": stage_bytes(dom, counters)}
        if winst:
            # the blend kernels are bound by instruction issue, not HBM: executed warp
            # instructions per launch (ncu capture, profiles/traffic.json) / live launch time vs 148 SMs x 4 schedulers
            sm_mhz = (clk or {}).get("sm_mhz") or 1965.0
            peak_issue = 148 * 4 * sm_mhz * 1e6
            out["roofline"]["issue"] = {"warp_inst_per_launch": winst, "achieved_ginst_s": winst / (stages[dom]["ms_per_launch"] * 1e-3) / 1e9,
                                        "peak_ginst_s": peak_issue / 1e9, "stale": stale,
                                        "frac": winst / (stages[dom]["ms_per_launch"] * 1e-3) / peak_issue}
        out["stages"] = stages
        a_iter = iteration_bytes(counters)
        out["raster"]["algorithmic_bytes_per_iter"] = a_iter
        out["raster"]["hbm_frac"] = a_iter / (raster_ms * 1e-3) / 1e9 / peak
        out["raster"]["stages_ms"] = {n: ms / cnt for n, (ms, cnt) in raster_stage.items() if cnt}

    extras = not args.no_extras
    # ---- dropin: the mapping step through the reference's own boundary (module under torch autograd + torch loss) ----
    if mine and extras and world == 1 and not args.unfused and not churn:
        dm = mapper_step.SurfelMapper(mod, wl.scene, dev, supervision=cfg["sup"], fused=False)

        def step_dropin(s):
            for i in range(VIEWS_PER_STEP):
                v = wl.ids(s, i)[0]
                dm.train_view(wl.view_dev[v], wl.gt_dev[v], kf_id=v)
        n_d = max(2, args.steps // 2)
        ms_d, _, _ = timed(step_dropin, n_d, 1, restore=lambda: None)
        out["dropin"] = {"value": VIEWS_PER_STEP * n_d / (ms_d / 1e3), "unit": "views/s", "steps": n_d, "warmup": 1,
                         "what": "same workload and metric through the reference's own plugin boundary: this repo's "
                                 "GaussianRasterizer + SparseGaussianAdam under torch autograd with the caller's torch "
                                 "loss (mapper_step.SurfelMapper(fused=False)) -- a VINGS-Mono checkout that only swaps "
                                 "the module; compare with the reference arm's value"}
    # ---- strong: BASELINE.json configs[3], one 16-keyframe window split over the ranks ----
    if extras and args.strong_config != "none" and args.config == "C2" and not args.unfused:
        torch.cuda.empty_cache()
        sw = Workload(args.strong_config, mod, dev, rank, world, pg, fused)
        sm_ = sw.mapper
        Ks = sw.K
        if Ks % world == 0:
            def window_pass(s):
                for j in range(Ks // world):   # one optimizer step = one view per rank
                    ids = [(j * world + q) % Ks for q in range(world)]
                    sm_.train_view(sw.view_dev[ids[rank]], sw.gt_dev[ids[rank]], kf_id=ids[rank], kf_ids=ids)
            n_w = 2
            ms_w, _, _ = timed(window_pass, n_w, 1, restore=sw.restore)
            out["strong"] = {"value": Ks * n_w / (ms_w / 1e3), "unit": "views/s", "scaling": "strong",
                             "ms_per_window": ms_w / n_w, "window_passes": n_w, "warmup": 1, "n_gpus": world,
                             "views_per_rank_per_window": Ks // world, "optimizer_steps_per_window": Ks // world,
                             "replicas_bit_identical": replicas_identical(sm_),
                             "workload": "%s: %s, %d surfels, ONE %d-keyframe window at %dx%d partitioned over %d rank(s); "
                                         "total work per window fixed" % (args.strong_config, SHAPE_NAMES[args.strong_config],
                                                                          sw.P0, Ks, sw.W, sw.H, world),
                             "parallelism": sm_.parallelism()}
    if mine and rank == 0 and world == 1 and extras and not args.no_cpu_baseline:
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
