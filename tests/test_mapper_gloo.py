"""CPU (gloo, world_size 2): the multi-rank mapping step -- window views partitioned over ranks, surfels
replicated, ONE all-reduce of the fused [P,16] gradient/score/visibility buffer, identical Adam step on every
rank.  Checked against a single process that performs the same batched step."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vings-mono_b200"), os.path.join(ROOT, "tests")]


BOOK = ("_local_scores", "_globalkf_max_scores", "_globalkf_id")


def _setup():
    import emu_module
    import mapper_step
    import synthetic
    torch.set_num_threads(2)
    scene, views = synthetic.make_scene("C1", P=300, K=2, W=64, H=48)
    gts = [synthetic.make_gt(scene, v, i) for i, v in enumerate(views)]
    vds = [mapper_step.view_to_device(v, torch.device("cpu")) for v in views]
    return emu_module, mapper_step, scene, vds, gts


def _worker(rank, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=2)
    mod, ms, scene, vds, gts = _setup()
    mapper = ms.SurfelMapper(mod, scene, torch.device("cpu"), "rgb+depth", world_size=2, process_group=dist.group.WORLD,
                             rank=rank)
    for it in range(2):
        ids = [(it + q) % 2 + 10 * it for q in range(2)]       # keyframe ids of both ranks' views, rank order
        v = (it + rank) % 2
        # step 0 passes the schedule, step 1 lets the mapper gather it (all_gather_object)
        loss = mapper.train_view(vds[v], gts[v], kf_id=ids[rank], kf_ids=ids if it == 0 else None)
    res = {k: getattr(mapper, k).detach().clone() for k in ms.GROUPS}
    res["scores"] = mapper._global_scores.clone()
    for k in BOOK:
        res[k] = getattr(mapper, k).clone()
    res["fused"] = mapper._fused.clone()
    res["loss"] = float(loss)
    out[rank] = res
    dist.destroy_process_group()


def test_two_rank_step_equals_single_process_batched_step():
    port = 29500 + os.getpid() % 2000
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(port, out), nprocs=2, join=True)
    a, b = out[0], out[1]
    mod, ms, scene, vds, gts = _setup()
    # replicas stay bit-identical (the all-reduce returns the same bits everywhere, Adam is deterministic)
    # -- the whole score bookkeeping included: add_records is applied for every view in rank order on every rank
    for k in list(ms.GROUPS) + ["scores"] + list(BOOK):
        assert torch.equal(a[k], b[k]), k
    # single-process restatement of the batched step: per-view weighting, summed grads/scores, OR-ed visibility
    ref = ms.SurfelMapper(mod, scene, torch.device("cpu"), "rgb+depth")
    for it in range(2):
        grads = {k: torch.zeros_like(getattr(ref, k)) for k in ms.GROUPS}
        score_sum = torch.zeros(ref.P, 2)
        vis = torch.zeros(ref.P, dtype=torch.bool)
        g0 = ref._global_scores[:, 0].clone()   # every view of a step is weighted with the scores from BEFORE the step
        for r in range(2):
            v = (it + r) % 2
            pred = ref.render(vds[v])
            ms.get_loss(ref.weights, pred, gts[v], ref.window).backward()
            cur = ref._zeros.grad.detach()
            gw = (cur[:, 0] / (torch.clamp(g0 + cur[:, 0], 0, 1e4) + 1e-6 + cur[:, 0])).unsqueeze(1)
            for k in ms.GROUPS:
                grads[k] += getattr(ref, k).grad * gw
                getattr(ref, k).grad = None
            score_sum += cur
            vis |= pred["radii"] > 0
            ref._add_records(cur, (it + r) % 2 + 10 * it)    # view by view, in rank order
        last_grads = torch.cat([grads[k].reshape(-1) for k in ms.GROUPS])
        for k in ms.GROUPS:
            getattr(ref, k).grad = grads[k]
        ref.optimizer.step(vis, ref.P)
        ref.optimizer.zero_grad()
    # the all-reduced buffer of the last step: 13 gradient columns, then scores, then the visibility count
    n13 = 13 * ref.P
    assert torch.allclose(a["fused"][:n13], last_grads, rtol=1e-4, atol=1e-9)
    assert torch.allclose(a["fused"][n13:n13 + 2 * ref.P].view(-1, 2), score_sum, rtol=1e-5, atol=1e-7)
    assert torch.equal(a["fused"][n13 + 2 * ref.P:] > 0, vis)
    # Adam without bias correction moves a coordinate by ~3.2*lr*sign(g) on its first steps, so a gradient
    # that rounds to the other side of zero flips a whole step: compare per element, allow 2 % such flips
    for k in ms.GROUPS:
        close = torch.isclose(a[k], getattr(ref, k).detach(), rtol=1e-5, atol=1e-7)
        assert float(close.float().mean()) > 0.98, (k, float(close.float().mean()))
        assert float((a[k] - getattr(ref, k).detach()).abs().max()) <= 2 * 2 * 3.2 * ms.LR[k] + 1e-6, k
    assert torch.allclose(a["scores"], ref._global_scores, rtol=1e-5, atol=1e-6)
    assert torch.allclose(a["_local_scores"], ref._local_scores, rtol=1e-5, atol=1e-6)
    assert torch.allclose(a["_globalkf_max_scores"], ref._globalkf_max_scores, rtol=1e-5, atol=1e-7)
    assert float((a["_globalkf_id"] != ref._globalkf_id).float().mean()) < 0.02   # ties broken by ulp-level score noise
    assert int(a["_globalkf_id"].max()) >= 10                                     # ids of the second step were recorded
    # and something was actually learned/moved
    assert float((a["_opacity"] - torch.logit(scene["opacities"].clamp(1e-6, 1 - 1e-6))).abs().max()) > 0


def test_single_rank_step_runs_and_updates_only_visible_surfels():
    mod, ms, scene, vds, gts = _setup()
    mapper = ms.SurfelMapper(mod, scene, torch.device("cpu"), "rgb+depth+normal+dist")
    before = mapper._xyz.detach().clone()
    loss = mapper.train_view(vds[0], gts[0], kf_id=3)
    assert torch.isfinite(loss)
    pred = mapper.render(vds[0])
    moved = (mapper._xyz.detach() != before).any(dim=1)
    assert moved.any() and not moved[pred["radii"] == 0].any()
    assert int(mapper._globalkf_id.max()) == 3
