"""Rows N1/N2 of SURVEY.md section 8f: the fused mapping loss and the fused per-surfel update.

Oracles:
* tests/golden/loss_*.npz -- produced by the reference's own Python (scripts/gaussian/loss_utils.get_loss,
  normal_utils.depth_propagate_normal, tf.TFer) by tests/golden/make_loss_golden.py;
* mapper_step.SurfelMapper(fused=False) -- the torch restatement of the caller (itself checked against the
  golden vectors here, on CPU) driving the same rasterizer.
Tolerances are fp32 rounding: the fused kernels sum in a different association (separable filter, block
reductions) than torch's conv2d / masked means.
"""
import ctypes
import os

import numpy as np
import pytest
import torch

import mapper_step
import synthetic
import util

GOLDEN_LOSS = ("loss_rgb_depth", "loss_normal_dist")
WKEYS = ("rgb_loss", "normal_loss", "alpha_loss", "depth_loss", "dist_loss")


def load_loss_case(name):
    z = np.load(os.path.join(util.GOLDEN, name + ".npz"))
    c = {k: torch.from_numpy(z[k]) for k in ("color", "allmap", "gt_rgb", "gt_depth", "gt_cov", "w2c", "g_color",
                                              "g_allmap")}
    c["loss"] = float(z["loss"])
    fu, fv, cu, cv = [float(x) for x in z["intr"]]
    c.update(fx=fv, fy=fu, cx=cv, cy=cu)  # the reference's u is the row (configs/rtg/hotel.yaml:31-34)
    c["weights"] = {k: float(w) for k, w in zip(WKEYS, z["weights"])}
    return c


def torch_loss(c, dev="cpu"):
    """mapper_step's restatement of render_raw post-processing + get_loss, with autograd."""
    color = c["color"].to(dev).clone().requires_grad_(True)
    allmap = c["allmap"].to(dev).clone().requires_grad_(True)
    w2c = c["w2c"].to(dev)
    alpha = allmap[1:2]
    depth = torch.nan_to_num(allmap[0:1] / alpha, 0, 0)
    pred = dict(accum=alpha, rgb=color, depth=depth, dist=allmap[6:7])
    pred["normal"] = (allmap[2:5].permute(1, 2, 0) @ w2c[:3, :3]).permute(2, 0, 1)
    sn = mapper_step.depth_propagate_normal(depth.squeeze(0), c["fx"], c["fy"], c["cx"], c["cy"]).permute(2, 0, 1)
    pred["surf_normal"] = (sn.permute(1, 2, 0) @ w2c[:3, :3]).permute(2, 0, 1)
    gt = dict(rgb=c["gt_rgb"].to(dev), depth=c["gt_depth"].to(dev), depth_cov=c["gt_cov"].to(dev))
    loss = mapper_step.get_loss(c["weights"], pred, gt, mapper_step._ssim_window(3, dev))
    loss.backward()
    return float(loss.detach()), color.grad, allmap.grad


@pytest.mark.parametrize("name", GOLDEN_LOSS)
def test_torch_restatement_matches_reference_loss(name):
    """The caller-side torch code of mapper_step (used by the reference arm of bench.py and by fused=False)
    reproduces the reference's get_loss and its gradients."""
    c = load_loss_case(name)
    loss, g_color, g_allmap = torch_loss(c)
    assert abs(loss - c["loss"]) <= 1e-6 * max(1.0, abs(c["loss"]))
    assert torch.equal(torch.isnan(g_allmap), torch.isnan(c["g_allmap"]))
    assert util.rel_err(g_color, c["g_color"]) < 1e-6
    assert util.rel_err(torch.nan_to_num(g_allmap), torch.nan_to_num(c["g_allmap"])) < 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("name", GOLDEN_LOSS)
def test_map_loss_matches_reference_golden(name):
    mine = util.load_mine()
    c = load_loss_case(name)
    cu = lambda k: c[k].cuda()
    loss, g_color, g_allmap = mine.mapping.map_loss(cu("color"), cu("allmap"), cu("gt_rgb"), cu("gt_depth"),
                                                    cu("gt_cov"), cu("w2c"), c["cx"], c["cy"], c["fx"], c["fy"],
                                                    c["weights"])
    assert abs(float(loss[0]) - c["loss"]) <= 2e-6 * max(1.0, abs(c["loss"]))
    assert torch.isfinite(g_color).all() and torch.isfinite(g_allmap).all()
    assert util.rel_err(g_color.cpu(), c["g_color"]) < 2e-5
    ref = c["g_allmap"]
    nan = torch.isnan(ref)
    # torch yields NaN (0/0) where no surfel covers the pixel; the kernel writes 0 there (surfel_b200.h)
    # (channel 0 gets 0; channel 1 keeps its finite alpha-loss term)
    assert not nan[2:].any() and torch.equal(nan[0], nan[1])
    assert float(g_allmap[0].cpu()[nan[0]].abs().max()) == 0.0
    sky = c["gt_rgb"].sum(0) == 0
    expect = torch.where(sky, torch.full_like(ref[1], c["weights"]["alpha_loss"] / float(sky.sum())), torch.zeros_like(ref[1]))
    assert torch.allclose(g_allmap[1].cpu()[nan[1]], expect[nan[1]], rtol=1e-6, atol=0)
    for ch in range(7):
        r, m = ref[ch][~nan[ch]], g_allmap[ch].cpu()[~nan[ch]]
        assert util.rel_err(m, r) < 2e-5, (ch, util.rel_err(m, r))
        assert float((m - r).abs().max()) <= 1e-5 * max(float(r.abs().max()), 1e-12) + 1e-9, ch


@pytest.mark.gpu
def test_map_loss_edge_cases():
    """No valid pixel / no sky pixel: means over empty sets are NaN like torch's, gradients stay finite zeros."""
    mine = util.load_mine()
    c = load_loss_case("loss_rgb_depth")
    gt_rgb = c["gt_rgb"].clone()
    gt_rgb[:, :4, :] = 0.5  # no sky
    out = mine.mapping.map_loss(c["color"].cuda(), c["allmap"].cuda(), gt_rgb.cuda(), c["gt_depth"].cuda(),
                                c["gt_cov"].cuda(), c["w2c"].cuda(), c["cx"], c["cy"], c["fx"], c["fy"], c["weights"])
    assert torch.isnan(out[0][0]) and torch.isnan(out[0][4]) and torch.isfinite(out[0][1])
    assert torch.isfinite(out[1]).all() and torch.isfinite(out[2]).all()
    out = mine.mapping.map_loss(c["color"].cuda(), c["allmap"].cuda(), torch.zeros_like(c["gt_rgb"]).cuda(),
                                c["gt_depth"].cuda(), c["gt_cov"].cuda(), c["w2c"].cuda(), c["cx"], c["cy"], c["fx"],
                                c["fy"], c["weights"])
    assert torch.isnan(out[0][0])
    assert float(out[1].abs().max()) == 0.0  # no valid pixel: no colour gradient
    assert torch.isfinite(out[2]).all()


@pytest.mark.gpu
@pytest.mark.parametrize("name", GOLDEN_LOSS)
def test_map_loss_full_size_matches_torch(name):
    """At C2's image size against the torch restatement on the GPU (random maps)."""
    mine = util.load_mine()
    c = load_loss_case(name)
    W, H = 1241, 376
    g = torch.Generator(device="cpu").manual_seed(5)
    alpha = torch.rand(H, W, generator=g).clamp(0.05, 0.999)
    alpha[torch.rand(H, W, generator=g) < 0.02] = 0.0
    yy, xx = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    depth = 3.0 + 0.01 * xx + 0.02 * yy + torch.sin(xx / 30.0) + 0.02 * torch.rand(H, W, generator=g)
    allmap = torch.zeros(7, H, W)
    allmap[0], allmap[1] = depth * alpha, alpha
    allmap[2:5] = torch.nn.functional.normalize(torch.randn(3, H, W, generator=g), dim=0) * alpha
    allmap[6] = 0.01 * torch.rand(H, W, generator=g)
    color = torch.rand(3, H, W, generator=g)
    gt_rgb = (0.5 * color + 0.5 * torch.rand(3, H, W, generator=g))
    gt_rgb[:, :50] = 0.0
    gt_depth = (depth + 0.1 * torch.randn(H, W, generator=g)).unsqueeze(0)
    gt_depth[0][torch.rand(H, W, generator=g) < 0.1] = 0.0
    big = dict(c, color=color, allmap=allmap, gt_rgb=gt_rgb, gt_depth=gt_depth, gt_cov=0.5 + torch.rand(1, H, W, generator=g),
               fx=0.8 * W, fy=0.8 * W, cx=W / 2.0, cy=H / 2.0)
    loss_t, gc_t, ga_t = torch_loss(big, dev="cuda")
    loss, g_color, g_allmap = mine.mapping.map_loss(color.cuda(), allmap.cuda(), gt_rgb.cuda(), gt_depth.cuda(),
                                                    big["gt_cov"].cuda(), c["w2c"].cuda(), big["cx"], big["cy"],
                                                    big["fx"], big["fy"], c["weights"])
    assert abs(float(loss[0]) - loss_t) <= 1e-5 * max(1.0, abs(loss_t))
    assert util.rel_err(g_color, gc_t) < 1e-4
    nan = torch.isnan(ga_t)
    assert util.rel_err(torch.where(nan, torch.zeros_like(ga_t), g_allmap), torch.nan_to_num(ga_t)) < 1e-4


@pytest.mark.gpu
def test_activate_matches_torch():
    mine = util.load_mine()
    g = torch.Generator(device="cpu").manual_seed(3)
    o = (3 * torch.randn(5000, 1, generator=g)).cuda()
    s = (torch.randn(5000, 2, generator=g) - 3).cuda()
    r = torch.randn(5000, 4, generator=g).cuda()
    r[7] = 0.0
    ao, as_, ar = mine.mapping.activate(o, s, r)
    assert torch.allclose(ao, torch.sigmoid(o), rtol=2e-7, atol=1e-9)
    assert torch.allclose(as_, torch.exp(s), rtol=2e-7, atol=0)
    assert torch.allclose(ar, torch.nn.functional.normalize(r), rtol=1e-6, atol=1e-7)


@pytest.mark.gpu
@pytest.mark.parametrize("sup", ["rgb+depth", "rgb+depth+normal+dist"])
def test_fused_view_matches_torch_view(sup):
    """Three training views: torch autograd + torch loss + 5 Adam launches vs the fused path, ONE step at a time
    from bit-identical state (util.sync_mapper_state; see there why trajectories are not compared)."""
    mine = util.load_mine()
    P = 20000
    scene, views = synthetic.make_scene("C3", P=P, W=160, H=96)
    dev = torch.device("cuda")
    a = mapper_step.SurfelMapper(mine, scene, dev, supervision=sup, fused=False)
    b = mapper_step.SurfelMapper(mine, scene, dev, supervision=sup, fused=True)
    g = torch.Generator(device="cpu").manual_seed(1)
    stable = (torch.rand(P, generator=g) < 0.1).cuda()
    a._stable_mask = stable.clone()
    b._stable_mask.copy_(stable)
    init = {k: getattr(a, k).data.clone() for k in mapper_step.GROUPS}
    for i in range(3):
        v = views[-1 - (i % len(views))]
        gt = {k: t.cuda() for k, t in synthetic.make_gt(scene, v, i).items()}
        vd = mapper_step.view_to_device(v, dev)
        before = util.mapper_snapshot(b)
        la, lb = float(a.train_view(vd, gt, kf_id=i + 1)), float(b.train_view(vd, gt, kf_id=i + 1))
        assert abs(la - lb) <= 2e-5 * max(1.0, abs(la)), (i, la, lb)
        util.assert_one_step_agrees(a, b, before, ctx="%s view %d" % (sup, i))
        util.sync_mapper_state(a, b)
    for k in mapper_step.GROUPS:
        assert float((getattr(b, k).data - init[k]).abs().sum()) > 0, k
    # surfels frozen by _stable_mask did not move
    assert torch.equal(b._xyz.data[stable], init["_xyz"][stable])
    assert not torch.equal(b._xyz.data[~stable], init["_xyz"][~stable])


@pytest.mark.gpu
def test_update_dp_modes_equal_local_on_one_rank():
    """SB200_UPDATE_DP_PRE + (an all-reduce over one rank = identity) + SB200_UPDATE_DP_POST is the LOCAL pass on
    the same backward outputs: same weighting, same bookkeeping, same Adam -- bit for bit.  (Two separate
    backward passes would differ by atomic-order jitter, so the rasterizer runs once.)"""
    mine = util.load_mine()
    mp = mine.mapping
    scene, views = synthetic.make_scene("C3", P=20000, W=160, H=96)
    dev = torch.device("cuda")
    m = mapper_step.SurfelMapper(mine, scene, dev, supervision="rgb+depth+normal+dist", fused=True)
    for i in range(2):  # leaves non-trivial moments, scores and one view's gradients in m._ws
        v = views[-1 - i]
        m.train_view(mapper_step.view_to_device(v, dev), {k: t.cuda() for k, t in synthetic.make_gt(scene, v, i).items()},
                     kf_id=i + 1)
    g = torch.Generator(device="cpu").manual_seed(4)
    stable = (torch.rand(m.P, generator=g) < 0.1).cuda()

    def state():
        prm = [getattr(m, k).data.clone() for k in mapper_step.GROUPS]
        st = [m.optimizer.state[getattr(m, k)] for k in mapper_step.GROUPS]
        return (prm, [x["exp_avg"].clone() for x in st], [x["exp_avg_sq"].clone() for x in st],
                [m._local_scores.clone(), m._global_scores.clone(), m._globalkf_max_scores.clone(), m._globalkf_id.clone()])

    lr = [mapper_step.LR[k] for k in mapper_step.GROUPS]
    pa, ma, va, ba = state()
    mp.surfel_update(mp.UPDATE_LOCAL, pa, ma, va, lr, m._ws, stable, *ba, 7)
    pb, mb, vb, bb = state()
    fused = torch.zeros(16 * m.P, device=dev)
    mp.surfel_update(mp.UPDATE_DP_PRE, pb, mb, vb, lr, m._ws, stable, *bb, 7, fused=fused)
    assert all(torch.equal(x, y) for x, y in zip(pb + bb, state()[0] + state()[3]))  # PRE only fills the buffer
    assert torch.equal(fused[15 * m.P:] > 0, m._ws.radii > 0)
    mp.surfel_update(mp.UPDATE_DP_POST, pb, mb, vb, lr, m._ws, stable, *bb, 7, fused=fused)
    for x, y in zip(pa + ma + va + ba, pb + mb + vb + bb):
        assert torch.equal(x, y)
    assert not torch.equal(pa[0], m._xyz.data)


@pytest.mark.gpu
def test_forward_async_is_the_forward_and_survives_a_too_small_scratch():
    """sb200_forward_async (pair count kept on the device, no host wait inside the forward) gives the bits of the
    two-stage forward when the kept scratch is large enough; with a scratch that is too small it stays inside its
    buffers, confirm() reports it and grows the scratch."""
    mine = util.load_mine()
    mp = mine.mapping
    scene, views = synthetic.make_scene("C3", P=60000, W=320, H=192)
    inp = util.cuda_inputs(synthetic.view_inputs(scene, views[-1]))
    dev = inp["means3D"].device
    args = (inp["means3D"], inp["scales"], inp["rotations"], inp["opacities"], inp["colors"], inp["bg"],
            inp["viewmatrix"], inp["projmatrix"])
    a = mp.RasterWorkspace(inp["means3D"].shape[0], inp["W"], inp["H"], dev)
    a.forward(*args)
    R = a.R
    io = (ctypes.c_size_t * 3)()
    mine._load().sb200_image_layout(inp["W"], inp["H"], io)
    kept = io[2] + 8 * inp["W"] * inp["H"]   # ranges, final T/M1/M2, contributor counts (the launch order is not unique)
    ref = (a.color.clone(), a.allmap.clone(), a.radii.clone(), a.sorted_ids(R).clone(), a.image[:kept].clone())
    assert R > 5000
    # same workspace again, asynchronously (scratch holds 1.25 R pairs)
    a.color.zero_(); a.allmap.zero_(); a.point_list.zero_()
    a.forward_async(*args)
    assert a.confirm() and mp.last_counters["R"] == R
    torch.cuda.synchronize()
    got = (a.color, a.allmap, a.radii, a.sorted_ids(R), a.image[:kept])
    for x, y in zip(got, ref):
        assert torch.equal(x, y)
    # a scratch of R/3 pairs surrounded by canaries: nothing outside is written, confirm() says no and grows it
    b = mp.RasterWorkspace(inp["means3D"].shape[0], inp["W"], inp["H"], dev)
    small = R // 3
    nbytes = mine._load().sb200_binning_bytes(small)
    guard = 4096
    bin_buf = torch.full((nbytes + 2 * guard,), 0x5A, dtype=torch.uint8, device=dev)
    pl_bytes = mine._load().sb200_pair_list_bytes(small, inp["W"], inp["H"])
    pl_buf = torch.full((pl_bytes + 2 * guard,), 0x5A, dtype=torch.uint8, device=dev)
    b.binning, b.point_list, b.pair_capacity = bin_buf[guard:guard + nbytes], pl_buf[guard:guard + pl_bytes], small
    b.forward_async(*args)
    torch.cuda.synchronize()
    assert (bin_buf[:guard] == 0x5A).all() and (bin_buf[guard + nbytes:] == 0x5A).all()
    assert (pl_buf[:guard] == 0x5A).all() and (pl_buf[guard + pl_bytes:] == 0x5A).all()
    assert not b.confirm() and b.pair_capacity >= R
    b.forward(*args)
    torch.cuda.synchronize()
    for x, y in zip((b.color, b.allmap, b.radii, b.sorted_ids(R)), ref):
        assert torch.equal(x, y)


@pytest.mark.gpu
def test_peer_exchange_kernels_equal_the_batched_step_on_one_device():
    """The three kernels of the NVLink exchange (sb200_dp_stage / _reduce / _post, csrc/dp_exchange.cu) with the
    "ranks" emulated on ONE device (every rank's buffers are ordinary tensors here, the peer pointer arrays point at
    them): the result must be, bit for bit, (a) what the NCCL formulation gives when its two all-reduces are replaced
    by torch sums / maxima over the same per-view rows, and (b) for the score bookkeeping, what add_records
    (gaussian_model.py:163-168) gives when applied view by view in rank order."""
    mine = util.load_mine()
    mp, lib = mine.mapping, mine._load()
    N = 3
    scene, views = synthetic.make_scene("C3", P=30000, W=160, H=96)
    dev = torch.device("cuda")
    sup = "rgb+depth+normal+dist"
    m = mapper_step.SurfelMapper(mine, scene, dev, supervision=sup, fused=True)
    gts = [{k: t.cuda() for k, t in synthetic.make_gt(scene, v, i).items()} for i, v in enumerate(views[:N + 1])]
    vds = [mapper_step.view_to_device(v, dev) for v in views[:N + 1]]
    m.train_view(vds[N], gts[N], kf_id=9)            # non-trivial moments and scores
    P = m.P
    g = torch.Generator(device="cpu").manual_seed(4)
    stable = (torch.rand(P, generator=g) < 0.1).cuda()
    prm = [getattr(m, k).data for k in mapper_step.GROUPS]
    lr = [mapper_step.LR[k] for k in mapper_step.GROUPS]
    kf_ids = [5, 3, 8]
    # one backward per "rank" from the same parameters, each into its own workspace
    wss = []
    opac, scales, rot = mp.activate(prm[2], prm[3], prm[4])
    for q in range(N):
        ws = mp.RasterWorkspace(P, m.W, m.H, dev)
        ws.forward(prm[0], scales, rot, opac, prm[1], m.bg, vds[q]["viewmatrix"], vds[q]["projmatrix"])
        loss, d_color, d_allmap = mp.map_loss(ws.color, ws.allmap, gts[q]["rgb"], gts[q]["depth"], gts[q]["depth_cov"],
                                              vds[q]["w2c"], m.cx, m.cy, m.fx, m.fy, m.weights, ws=ws.loss_ws, out=ws.loss_out)
        ws.backward(d_color, d_allmap, prm[0], scales, rot, m.bg, vds[q]["viewmatrix"], vds[q]["projmatrix"],
                    vds[q]["tanfovx"], vds[q]["tanfovy"])
        wss.append(ws)
    assert sum(int((ws.radii > 0).sum()) for ws in wss) > P // 2

    def state():
        st = [m.optimizer.state[getattr(m, k)] for k in mapper_step.GROUPS]
        return ([p.clone() for p in prm], [x["exp_avg"].clone() for x in st], [x["exp_avg_sq"].clone() for x in st],
                [m._local_scores.clone(), m._global_scores.clone(), m._globalkf_max_scores.clone(), m._globalkf_id.clone()])

    # ---- (A) the peer-memory kernels, ranks emulated: N exchange buffers on this device, one handle per "rank".  The
    # kernels' flag hand-offs run for real (all stage kernels are queued before any reduce, all reduces before any
    # post, so every wait finds its flags already set) ----
    pa, ma, va, ba = state()
    vp = ctypes.c_void_p
    st_ = vp(torch.cuda.current_stream().cuda_stream)
    nbytes = lib.sb200_dp_buffer_bytes(P, N)
    bufs = [torch.zeros(nbytes, dtype=torch.uint8, device=dev) for _ in range(N)]
    bases = (vp * N)(*[t.data_ptr() for t in bufs])
    handles = [lib.sb200_dp_create(r, N, P, bases) for r in range(N)]
    assert all(handles), lib.sb200_last_error()
    S = lib.sb200_dp_slice(vp(handles[0]), P)
    assert S % 256 == 0 and S * N >= P
    for q in range(N):
        ws = wss[q]
        assert lib.sb200_dp_stage(vp(handles[q]), P, vp(ws.d_means3D.data_ptr()), vp(ws.d_colors.data_ptr()),
                                  vp(ws.d_opacity.data_ptr()), vp(ws.d_scales.data_ptr()), vp(ws.d_rot.data_ptr()),
                                  vp(ws.scores.data_ptr()), vp(ws.radii.data_ptr()), vp(prm[2].data_ptr()),
                                  vp(prm[3].data_ptr()), vp(prm[4].data_ptr()), st_) == 0, lib.sb200_last_error()
    kf_arr = (ctypes.c_int64 * N)(*kf_ids)
    for r in range(N):
        assert lib.sb200_dp_reduce(vp(handles[r]), P, vp(ba[1].data_ptr()), st_) == 0, lib.sb200_last_error()
    arr5 = lambda ts: (vp * 5)(*[t.data_ptr() for t in ts])
    # rank 0's replica is updated; the other "ranks" would do the same from the same result rows
    assert lib.sb200_dp_post(vp(handles[0]), P, kf_arr, arr5(pa), arr5(ma), arr5(va), vp(stable.data_ptr()),
                             vp(ba[0].data_ptr()), vp(ba[1].data_ptr()), vp(ba[2].data_ptr()), vp(ba[3].data_ptr()),
                             (ctypes.c_float * 5)(*lr), 0.9, 0.999, 1e-15, st_) == 0, lib.sb200_last_error()
    torch.cuda.synchronize()
    for r in range(N):
        assert lib.sb200_dp_error(vp(handles[r]), st_) == 0, lib.sb200_last_error()
        lib.sb200_dp_destroy(vp(handles[r]))

    # ---- (B) the NCCL formulation with the collectives replaced by torch reductions over the same rows ----
    pb, mb, vb, bb = state()
    fused = torch.zeros(16 * P, device=dev)
    s1_max = torch.zeros(P, device=dev)
    s0_max = torch.full((P,), -1.0, device=dev)
    kf_arg = torch.zeros(P, dtype=torch.int64, device=dev)
    for q in range(N):
        f = torch.zeros(16 * P, device=dev)
        mp.surfel_update(mp.UPDATE_DP_PRE, pb, mb, vb, lr, wss[q], stable, *bb, kf_ids[q], fused=f)
        fused += f
        cur = wss[q].scores
        s1_max = torch.maximum(s1_max, cur[:, 1])
        first = (cur[:, 0] > s0_max) & (wss[q].radii > 0)
        s0_max = torch.where(first, cur[:, 0], s0_max)
        kf_arg = torch.where(first, torch.full_like(kf_arg, kf_ids[q]), kf_arg)
    fused[13 * P:15 * P].view(P, 2)[:, 1] = s1_max
    kfm0, kid0 = bb[2].clone(), bb[3].clone()
    mp.surfel_update(mp.UPDATE_DP_POST, pb, mb, vb, lr, wss[0], stable, *bb, kf_ids[0], fused=fused)
    replace = kfm0 < s0_max
    bb[2].copy_(torch.where(replace, s0_max, kfm0))
    bb[3].copy_(torch.where(replace, kf_arg, kid0))
    names = ["p%d" % i for i in range(5)] + ["m%d" % i for i in range(5)] + ["v%d" % i for i in range(5)] + \
            ["local", "global", "kf_max", "kf_id"]
    for n_, x, y in zip(names, pa + ma + va + ba, pb + mb + vb + bb):
        assert torch.equal(x, y), (n_, float((x.double() - y.double()).abs().max()))
    assert not torch.equal(pa[0], prm[0]) and torch.equal(pa[0][stable], prm[0][stable])

    # ---- (C) add_records view by view, in rank order ----
    loc, glo, kfm, kid = m._local_scores.clone(), m._global_scores.clone(), m._globalkf_max_scores.clone(), m._globalkf_id.clone()
    for q in range(N):
        cur = wss[q].scores
        loc[:, 0] += cur[:, 0]
        loc[:, 1] = torch.where(cur[:, 1] > loc[:, 1], cur[:, 1], loc[:, 1])
        glo[:, 0] += cur[:, 0]
        glo = torch.clamp(glo, 0, 1e4)
        rep = kfm < cur[:, 0]
        kfm = torch.where(rep, cur[:, 0], kfm)
        kid = torch.where(rep, torch.full_like(kid, kf_ids[q]), kid)
    assert torch.equal(ba[2], kfm) and torch.equal(ba[3], kid) and torch.equal(ba[0][:, 1], loc[:, 1])
    assert torch.allclose(ba[0][:, 0], loc[:, 0], rtol=1e-6, atol=0) and torch.allclose(ba[1], glo, rtol=1e-6, atol=0)


@pytest.mark.gpu
def test_backward_update_equals_backward_then_update():
    """sb200_backward_update (backward blend + ONE per-surfel pass over the visible surfels) against the two-call form
    sb200_backward -> sb200_surfel_update(LOCAL) / sb200_dp_stage it replaces in the fused mapper, on the SAME
    accumulation records (R < 0 test hook: two backward blends would differ by atomic-order jitter).  What is left is
    fp32 contraction of the same backward-preprocess expressions inlined into another kernel: a few ulp."""
    mine = util.load_mine()
    mp, lib = mine.mapping, mine._load()
    scene, views = synthetic.make_scene("C3", P=30000, W=160, H=96)
    dev = torch.device("cuda")
    m = mapper_step.SurfelMapper(mine, scene, dev, supervision="rgb+depth+normal+dist", fused=True)
    gts = [{k: t.cuda() for k, t in synthetic.make_gt(scene, v, i).items()} for i, v in enumerate(views[:2])]
    vds = [mapper_step.view_to_device(v, dev) for v in views[:2]]
    m.train_view(vds[1], gts[1], kf_id=4)   # non-trivial moments and scores (through backward_update itself)
    assert float(m.optimizer.state[m._xyz]["exp_avg"].abs().sum()) > 0
    # the call hands the accumulation records back zeroed (only consumed ones are rewritten), so the next one needs
    # no memset: SB200_BWD_RECORDS_ZERO
    assert m._ws.grad_zero and not bool(m._ws.grad_ws.any())
    m.train_view(vds[0], gts[0], kf_id=5)   # ... and this view ran without the memset
    assert m._ws.grad_zero and not bool(m._ws.grad_ws.any())
    P, ws = m.P, m._ws
    prm = [getattr(m, k).data for k in mapper_step.GROUPS]
    lr = [mapper_step.LR[k] for k in mapper_step.GROUPS]
    g = torch.Generator(device="cpu").manual_seed(4)
    stable = (torch.rand(P, generator=g) < 0.1).cuda()
    opac, scales, rot = mp.activate(prm[2], prm[3], prm[4])
    vd, gt = vds[0], gts[0]
    ws.forward(prm[0], scales, rot, opac, prm[1], m.bg, vd["viewmatrix"], vd["projmatrix"])
    loss, d_color, d_allmap = mp.map_loss(ws.color, ws.allmap, gt["rgb"], gt["depth"], gt["depth_cov"], vd["w2c"], m.cx,
                                          m.cy, m.fx, m.fy, m.weights, ws=ws.loss_ws, out=ws.loss_out)
    ws.backward(d_color, d_allmap, prm[0], scales, rot, m.bg, vd["viewmatrix"], vd["projmatrix"], vd["tanfovx"], vd["tanfovy"])

    def state():
        st = [m.optimizer.state[getattr(m, k)] for k in mapper_step.GROUPS]
        return ([p.clone() for p in prm], [x["exp_avg"].clone() for x in st], [x["exp_avg_sq"].clone() for x in st],
                [m._local_scores.clone(), m._global_scores.clone(), m._globalkf_max_scores.clone(), m._globalkf_id.clone()])

    pa, ma, va, ba = state()
    mp.surfel_update(mp.UPDATE_LOCAL, pa, ma, va, lr, ws, stable, *ba, 7)
    pb, mb, vb, bb = state()
    mp.backward_update(ws, 0, d_color, d_allmap, prm[0], scales, rot, m.bg, vd["viewmatrix"], vd["projmatrix"],
                       vd["tanfovx"], vd["tanfovy"], pb, mb, vb, lr, stable, *bb, 7, reuse_records=True)
    torch.cuda.synchronize()
    for x, y in zip(ba, bb):      # the scores come straight out of the records: bookkeeping is exact
        assert torch.equal(x, y)
    for k, x, y in zip(mapper_step.GROUPS, ma + va, mb + vb):
        assert util.rel_err(y, x) < 1e-6, (k, util.rel_err(y, x))
    for k, x, y, p0 in zip(mapper_step.GROUPS, pa, pb, prm):
        assert float((x - y).abs().max()) <= 1e-4 * util.ADAM_RATIO_MAX * mapper_step.LR[k] + 2.5e-7 * float(x.abs().max()), k
        assert not torch.equal(x, p0) and torch.equal(y[stable], p0[stable])
        vis = ws.radii > 0
        assert torch.equal(y[~vis], p0[~vis])
    # mode 1: the multi-GPU stage rows, against sb200_dp_stage on the dense outputs (one-rank exchange handles)
    vp = ctypes.c_void_p
    st_ = vp(torch.cuda.current_stream().cuda_stream)
    nbytes = lib.sb200_dp_buffer_bytes(P, 1)
    buf_a, buf_b = torch.zeros(nbytes, dtype=torch.uint8, device=dev), torch.zeros(nbytes, dtype=torch.uint8, device=dev)
    buf_a[64 * P:65 * P] = 7
    buf_b[64 * P:65 * P] = 7      # visibility bytes start out as junk
    ha = lib.sb200_dp_create(0, 1, P, (vp * 1)(buf_a.data_ptr()))
    hb = lib.sb200_dp_create(0, 1, P, (vp * 1)(buf_b.data_ptr()))
    assert lib.sb200_dp_stage(vp(ha), P, vp(ws.d_means3D.data_ptr()), vp(ws.d_colors.data_ptr()), vp(ws.d_opacity.data_ptr()),
                              vp(ws.d_scales.data_ptr()), vp(ws.d_rot.data_ptr()), vp(ws.scores.data_ptr()),
                              vp(ws.radii.data_ptr()), vp(prm[2].data_ptr()), vp(prm[3].data_ptr()),
                              vp(prm[4].data_ptr()), st_) == 0
    before = state()
    mp.backward_update(ws, 1, d_color, d_allmap, prm[0], scales, rot, m.bg, vd["viewmatrix"], vd["projmatrix"],
                       vd["tanfovx"], vd["tanfovy"], prm, None, None, lr, None, None, None, None, None, 7,
                       dp=hb, reuse_records=True)
    torch.cuda.synchronize()
    stage_a, stage_b = buf_a[:64 * P].view(torch.float32), buf_b[:64 * P].view(torch.float32)
    vis_a, vis_b = buf_a[64 * P:65 * P], buf_b[64 * P:65 * P]   # layout: stage rows f32[P][16], then the vis bytes
    assert torch.equal(vis_a, vis_b) and torch.equal(vis_a.bool(), ws.radii > 0)
    seen = vis_a.bool()
    assert util.rel_err(stage_b.view(P, 16)[seen], stage_a.view(P, 16)[seen]) < 1e-6
    assert torch.equal(stage_b.view(P, 16)[:, 13:15][seen], ws.scores[seen])
    lib.sb200_dp_destroy(vp(ha)); lib.sb200_dp_destroy(vp(hb))
    for x, y in zip(before[0] + before[1] + before[2] + before[3], state()[0] + state()[1] + state()[2] + state()[3]):
        assert torch.equal(x, y)   # mode 1 updates nothing
