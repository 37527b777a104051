"""Host-side mapping step around the rasterizer -- the caller the drop-in sits behind.

Mirrors one iteration of VINGS-Mono's ``train_once_gaussian`` loop
(scripts/gaussian/gaussian_base.py:343-404) with the rasterizer module as a parameter, so the very
same step can be driven through this package *or* through the unmodified reference build:

    render_raw      gaussian_base.py:142-224   activations, settings, rasterizer call, depth/alpha, normals
    get_loss        loss_utils.py:110-153      0.8 L1 + 0.2 (1-SSIM), normal, alpha, 1/cov-weighted depth, dist
    add_records     gaussian_model.py:163-168  score bookkeeping
    weighting_grad  gaussian_utils.py:56-62    per-surfel gradient re-weighting by importance scores
    optimizer.step  gaussian_base.py:400-403   SparseGaussianAdam.step(radii > 0, P)

Multi-GPU (new; the reference has no multi-GPU code): the views of the keyframe
window are partitioned over the ranks, surfels are replicated.  Every optimizer step takes ONE view per rank:
per surfel the per-view-weighted gradients are summed, the score bookkeeping (add_records) is applied for every
view IN RANK ORDER -- so every replica ends up with the bits a single process taking the same views in that order
would have, `_globalkf_id` included -- and the identical masked Adam step follows.  The fused path exchanges only
the rows of surfels a view saw, through kernels over NVLink peer memory (``mapping.PeerExchange``:
stage -> reduce-scatter by peer loads -> all-gather fused into the update); where peer memory is unavailable it
falls back to NCCL (a dense ``[16,P]`` sum all-reduce plus a max all-reduce for the order-dependent bookkeeping).

PyTorch here is plumbing (parameters, autograd graph, loss glue); every rasterizer stage and the Adam
update run in the module handed in.
"""
import math
import os
import sys

import torch
import torch.nn.functional as F

LOSS_WEIGHTS = {
    # configs/*.yaml training_args.loss_weights; C3 adds normal + distortion supervision (BASELINE.json)
    "rgb+depth": dict(rgb_loss=1.0, depth_loss=1.0, alpha_loss=0.5, normal_loss=0.0, dist_loss=0.0),
    "rgb+depth+normal+dist": dict(rgb_loss=1.0, depth_loss=1.0, alpha_loss=0.5, normal_loss=0.05, dist_loss=100.0),
    "rgb": dict(rgb_loss=1.0, depth_loss=0.0, alpha_loss=0.5, normal_loss=0.0, dist_loss=0.0),
}
LR = dict(_xyz=0.00002, _rgb=0.0001, _opacity=0.02, _scaling=0.001, _rotation=0.001)  # configs/rtg/hotel.yaml:70-75
GROUPS = ("_xyz", "_rgb", "_opacity", "_scaling", "_rotation")
GROUP_COLS = dict(_xyz=3, _rgb=3, _opacity=1, _scaling=2, _rotation=4)


def _ssim_window(channel, device):
    """loss_utils.py:26-34 (11x11 Gaussian window, sigma 1.5)"""
    g = torch.tensor([math.exp(-(x - 5) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)])
    g = (g / g.sum()).unsqueeze(1)
    w2 = g.mm(g.t()).float().unsqueeze(0).unsqueeze(0)
    return w2.expand(channel, 1, 11, 11).contiguous().to(device)


def ssim_map(img1, img2, window):
    """loss_utils.py:36-55"""
    ch = img1.size(-3)
    mu1 = F.conv2d(img1, window, padding=5, groups=ch)
    mu2 = F.conv2d(img2, window, padding=5, groups=ch)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1 = F.conv2d(img1 * img1, window, padding=5, groups=ch) - mu1_sq
    s2 = F.conv2d(img2 * img2, window, padding=5, groups=ch) - mu2_sq
    s12 = F.conv2d(img1 * img2, window, padding=5, groups=ch) - mu1_mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    return ((2 * mu1_mu2 + C1) * (2 * s12 + C2)) / ((mu1_sq + mu2_sq + C1) * (s1 + s2 + C2))


def get_loss(weights, pred, gt, window):
    """loss_utils.py:110-153 with use_sky False"""
    sky_mask = gt["rgb"].sum(dim=0) == 0.0
    valid = torch.bitwise_and(~sky_mask, gt["depth"].sum(dim=0) > 0.0)
    Ll1 = torch.abs((pred["rgb"] - gt["rgb"])[:, valid]).mean()
    rgb_loss = 0.8 * Ll1 + 0.2 * (1.0 - ssim_map(pred["rgb"], gt["rgb"], window)[:, valid].mean())
    normal_loss = (1 - (pred["normal"] * pred["surf_normal"]).sum(dim=0)[valid]).mean()
    alpha_loss = (pred["accum"][:, sky_mask]).mean()
    weight = 1.0 / gt["depth_cov"]
    depth_loss = (torch.abs((pred["depth"] - gt["depth"])[:, valid]) * (weight[:, valid])).mean()
    dist_loss = (pred["dist"][:, valid]).mean()
    return (weights["rgb_loss"] * rgb_loss + weights["normal_loss"] * normal_loss + weights["alpha_loss"] * alpha_loss +
            weights["depth_loss"] * depth_loss + weights["dist_loss"] * dist_loss)


def dist_cuda2_reference(points):
    """distCUDA2 of the reference, verbatim semantics (gaussian_utils.py:12-16): scipy KD-tree on the host."""
    from scipy.spatial import KDTree
    points_np = points.detach().cpu().float().numpy()
    dists, _ = KDTree(points_np).query(points_np, k=4)
    return torch.tensor((dists[:, 1:] ** 2).mean(1), dtype=points.dtype, device=points.device)


def depth_propagate_normal(depth, fx, fy, cx, cy):
    """normal_utils.py:6-21 + tf.py:133-146 (depth -> camera points -> central-difference cross product)"""
    H, W = depth.shape
    v, u = torch.meshgrid(torch.arange(H, device=depth.device, dtype=depth.dtype),
                          torch.arange(W, device=depth.device, dtype=depth.dtype), indexing="ij")
    pc = torch.stack(((u - cx) / fx * depth, (v - cy) / fy * depth, depth), dim=-1)
    l2r = pc[1:H - 1, 2:W, :] - pc[1:H - 1, 0:W - 2, :]
    b2t = pc[0:H - 2, 1:W - 1, :] - pc[2:H, 1:W - 1, :]
    n = F.normalize(torch.cross(l2r, b2t, dim=-1), p=2, dim=-1)
    return F.pad(n, pad=(0, 0, 1, 1, 1, 1), mode="constant")


class SurfelMapper:
    """Surfel parameters + optimizer + the per-view training iteration, for a given rasterizer module."""

    def __init__(self, module, scene, device, supervision="rgb+depth", world_size=1, process_group=None, fused=None,
                 rank=0, exchange=None):
        """fused: run the per-view iteration through the module's fused mapping blocks (``module.mapping``:
        activate / RasterWorkspace / map_loss / surfel_update, SURVEY.md section 8f N1+N2) instead of torch
        autograd + torch loss.  Default: on when the module has them (the reference build has not)."""
        self.mod = module
        self.fused = hasattr(module, "mapping") if fused is None else bool(fused)
        if self.fused and not hasattr(module, "mapping"):
            raise RuntimeError("fused=True needs a rasterizer module with the fused mapping blocks")
        self.device = device
        self.weights = LOSS_WEIGHTS[supervision]
        self.W, self.H = scene["W"], scene["H"]
        self.fx = self.fy = 0.8 * self.W
        self.cx, self.cy = self.W / 2.0, self.H / 2.0
        self.world_size = world_size
        self.rank = rank
        self.pg = process_group
        # fused multi-GPU exchange: "p2p" (kernels over NVLink peer memory) or "nccl"; default p2p with a loud fallback
        self.exchange = exchange or os.environ.get("SB200_DP_EXCHANGE", "p2p")
        self._dpx = None
        self.bg = torch.zeros(3, device=device)
        self.window = _ssim_window(3, device)
        self.set_surfels(scene)

    # ---- surfel storage (raw, pre-activation parameters as gaussian_base.py:44-48,55-62) ----
    BOOK = dict(_local_scores=(2, torch.float32), _global_scores=(2, torch.float32), _stable_mask=(1, torch.bool),
                _globalkf_id=(1, torch.int64), _globalkf_max_scores=(1, torch.float32))

    def set_surfels(self, scene):
        """scene: ACTIVATED surfels (synthetic.make_scene): means3D, colors, opacities, scales, rotations."""
        self.set_raw(dict(_xyz=scene["means3D"], _rgb=scene["colors"],
                          _opacity=torch.logit(scene["opacities"].clamp(1e-6, 1 - 1e-6)),
                          _scaling=torch.log(scene["scales"]), _rotation=scene["rotations"]))

    def set_raw(self, raw, book=None):
        """raw: the five RAW parameter tensors (what checkpoints and PLY files hold, formats.py); book: optional score
        bookkeeping tensors (load_pt_ckpt).  A fresh optimizer either way, like load_*_ckpt's setup_optimizer()."""
        dev = self.device
        raw = {k: raw[k].detach().to(dev).float().contiguous() for k in GROUPS}
        P = raw["_xyz"].shape[0]
        raw["_opacity"] = raw["_opacity"].reshape(P, 1)
        zeros = dict(_local_scores=torch.zeros(P, 2, device=dev), _global_scores=torch.zeros(P, 2, device=dev),
                     _stable_mask=torch.zeros(P, dtype=torch.bool, device=dev),
                     _globalkf_id=torch.zeros(P, dtype=torch.long, device=dev),
                     _globalkf_max_scores=torch.zeros(P, device=dev))
        if book is not None:
            zeros = {k: book[k].detach().to(dev).to(zeros[k].dtype).reshape(zeros[k].shape).contiguous() for k in zeros}
        if self.fused:
            # every per-surfel array lives in one capacity-doubling store (diff_surfel_rasterization/store.py)
            spec = {k: (GROUP_COLS[k], torch.float32) for k in GROUPS}
            spec.update({"m" + k: (GROUP_COLS[k], torch.float32) for k in GROUPS})
            spec.update({"v" + k: (GROUP_COLS[k], torch.float32) for k in GROUPS})
            spec.update(self.BOOK)
            self.store = self.mod.store.SurfelStore(spec, dev, capacity=max(1 << 16, int(P * 1.25)))
            rows = dict(raw)
            rows.update(zeros)
            self.store.append(rows)
            self._ws = self.mod.mapping.RasterWorkspace(P, self.W, self.H, dev, capacity=self.store.capacity)
            self._bind_store()
        else:
            self._bind_tensors({k: raw[k].clone() for k in GROUPS}, None, zeros)

    def _bind_tensors(self, params, moments, book):
        """(Re)create parameters, optimizer and bookkeeping from tensors -- what setup_optimizer /
        prune_tensors_from_optimizer leave behind (gaussian_base.py:52-62,96-112).  moments: {name: (m, v)} or None
        for a fresh optimizer."""
        self.P = params["_xyz"].shape[0]
        for k in GROUPS:
            setattr(self, k, torch.nn.Parameter(params[k].requires_grad_(True)))
        self.optimizer = self.mod.SparseGaussianAdam(
            [{"params": [getattr(self, k)], "lr": LR[k], "name": k} for k in GROUPS], lr=0.0, eps=1e-15)
        if moments is not None:
            for k in GROUPS:
                self.optimizer.state[getattr(self, k)] = dict(step=torch.tensor(0.0), exp_avg=moments[k][0],
                                                              exp_avg_sq=moments[k][1])
        for name, t in book.items():
            setattr(self, name, t)
        self._bind_dp()

    def _bind_store(self):
        st = self.store
        P = st.P
        self._ws.set_P(P)
        self._act = (self._ws.act_opacity, self._ws.act_scales, self._ws.act_rot)
        self._bind_tensors({k: st.view(k) for k in GROUPS}, {k: (st.view("m" + k), st.view("v" + k)) for k in GROUPS},
                           dict(_local_scores=st.view("_local_scores"), _global_scores=st.view("_global_scores"),
                                _stable_mask=st.view("_stable_mask").reshape(P),
                                _globalkf_id=st.view("_globalkf_id").reshape(P),
                                _globalkf_max_scores=st.view("_globalkf_max_scores").reshape(P)))

    def _bind_dp(self):
        if self.world_size <= 1:
            return
        P, dev = self.P, self.device
        if self.fused and self.exchange == "p2p":
            cap = self.store.capacity
            if self._dpx is None or self._dpx.capacity < cap:
                try:
                    self._dpx = self.mod.mapping.PeerExchange(cap, self.pg, dev)
                except Exception as e:  # noqa: BLE001 -- no peer memory on this box: say so and use NCCL
                    print("[surfel_b200] peer-memory exchange unavailable (%r): falling back to NCCL all-reduce" % (e,),
                          file=sys.stderr, flush=True)
                    self.exchange = "nccl"
            if self.exchange == "p2p":
                return
        # NCCL / gloo paths: parameter .grad tensors are column blocks of one allocation so that a single all-reduce
        # moves gradients, the score sum and the visibility count
        self._fused = torch.zeros(16 * P, device=dev)
        self._grad_views, off = {}, 0
        for k in GROUPS:
            n = GROUP_COLS[k] * P
            self._grad_views[k] = self._fused[off:off + n].view(P, GROUP_COLS[k])
            off += n
        self._score_view = self._fused[off:off + 2 * P].view(P, 2)
        self._vis_view = self._fused[off + 2 * P:off + 3 * P]

    def parallelism(self):
        if self.world_size <= 1:
            return "single GPU"
        how = ("kernels over NVLink peer memory (stage rows of visible surfels, reduce-scatter by peer loads, all-gather "
               "fused into the Adam update; flag hand-offs between them, no collective call)" if (self.fused and self.exchange == "p2p")
               else "dense [16,P] sum all-reduce + [2,P] max all-reduce (NCCL)" if self.fused
               else "torch: fused [16,P] all-reduce + all-gather of the view scores")
        return "window views partitioned over %d ranks, surfels replicated; exchange = %s" % (self.world_size, how)

    def _all_kf_ids(self, kf_id, kf_ids):
        """The keyframe ids of every rank's view this step, in rank order (host ints)."""
        if kf_ids is not None:
            assert len(kf_ids) == self.world_size and int(kf_ids[self.rank]) == int(kf_id)
            return [int(k) for k in kf_ids]
        out = [None] * self.world_size   # slow path (host round trip): callers that know the schedule pass kf_ids
        torch.distributed.all_gather_object(out, int(kf_id), group=self.pg)
        return out

    # ---- churn: gaussian_model.py:92-161 (add_new_frame), gaussian_base.py:74-112 ----
    @torch.no_grad()
    def prune(self, delete_mask):
        """x = x[~delete_mask] for parameters, Adam moments and score bookkeeping
        (prune_tensors_from_optimizer + update_records("prune")).  Returns the new P."""
        if self.fused:
            self.store.prune(delete_mask)
            self._bind_store()
            return self.P
        keep = ~delete_mask
        moments = None
        if all(len(self.optimizer.state[getattr(self, k)]) for k in GROUPS):
            moments = {k: (self.optimizer.state[getattr(self, k)]["exp_avg"][keep],
                           self.optimizer.state[getattr(self, k)]["exp_avg_sq"][keep]) for k in GROUPS}
        self._bind_tensors({k: getattr(self, k).data[keep] for k in GROUPS}, moments,
                           {n: getattr(self, n)[keep] for n in self.BOOK})
        return self.P

    @torch.no_grad()
    def add_surfels(self, xyz, rgb, unnorm_rots):
        """The tail of add_new_frame (gaussian_model.py:136-161): isotropic log-scale from the mean squared distance
        to the 3 nearest new points, opacity 0.8, concatenation, zero bookkeeping rows, and a FRESH optimizer for
        every surfel (the reference ends with setup_optimizer(), which drops all Adam moments).  Returns the new P."""
        n = xyz.shape[0]
        if self.fused:
            dist2 = torch.clamp_min(self.mod.store.knn_mean_dist2(xyz), 0.0000001)
        else:
            dist2 = torch.clamp_min(dist_cuda2_reference(xyz), 0.0000001)
        new = dict(_xyz=xyz, _rgb=rgb, _scaling=torch.log(1.0 * torch.sqrt(dist2))[..., None].repeat(1, 2),
                   _rotation=unnorm_rots,
                   _opacity=torch.log(torch.full((n, 1), 0.8, device=xyz.device) / (1 - 0.8)).to(torch.float))
        if self.fused:
            self.store.append(new)
            for k in GROUPS:
                self.store.view("m" + k).zero_()
                self.store.view("v" + k).zero_()
            self._bind_store()
            return self.P
        book = {name: torch.cat((getattr(self, name), torch.zeros((n,) + tuple(getattr(self, name).shape[1:]),
                                                                    dtype=getattr(self, name).dtype, device=self.device)))
                for name in self.BOOK}
        self._bind_tensors({k: torch.cat((getattr(self, k).data, new[k]), dim=0) for k in GROUPS}, None, book)
        return self.P

    def params(self):
        return [getattr(self, k) for k in GROUPS]

    # ---- gaussian_base.py:142-224 ----
    def render(self, view):
        """view: dict(viewmatrix, projmatrix, campos on device; tanfovx, tanfovy; w2c [4,4] on device)"""
        m = self.mod
        P, H, W = self.P, self.H, self.W
        screenspace_points = torch.zeros_like(self._xyz, requires_grad=True) + 0
        try:
            screenspace_points.retain_grad()
        except Exception:
            pass
        pixel_mask = torch.ones(H * W, dtype=torch.bool, device=self.device)
        settings = m.GaussianRasterizationSettings(
            image_height=H, image_width=W, tanfovx=view["tanfovx"], tanfovy=view["tanfovy"], bg=self.bg,
            scale_modifier=1.0, viewmatrix=view["viewmatrix"], projmatrix=view["projmatrix"], sh_degree=0,
            campos=view["campos"], prefiltered=False, debug=False, pixel_mask=pixel_mask)
        rasterizer = m.GaussianRasterizer(raster_settings=settings)
        self._zeros = torch.zeros_like(self._xyz[:, :2]).contiguous().requires_grad_(True)
        rendered_image, radii, allmap = rasterizer(
            means3D=self._xyz, means2D=screenspace_points, shs=None, colors_precomp=self._rgb,
            opacities=torch.sigmoid(self._opacity), scales=torch.exp(self._scaling),
            rotations=F.normalize(self._rotation), scores=self._zeros, cov3D_precomp=None)
        w2c = view["w2c"]
        render_alpha = allmap[1:2]
        depth = torch.nan_to_num(allmap[0:1] / render_alpha, 0, 0)
        rets = dict(radii=radii, accum=render_alpha, rgb=rendered_image, depth=depth, dist=allmap[6:7])
        rets["normal"] = (allmap[2:5].permute(1, 2, 0) @ (w2c[:3, :3])).permute(2, 0, 1)
        sn = depth_propagate_normal(depth.squeeze(0), self.fx, self.fy, self.cx, self.cy).permute(2, 0, 1)
        rets["surf_normal"] = (sn.permute(1, 2, 0) @ (w2c[:3, :3])).permute(2, 0, 1)
        return rets

    # ---- gaussian_base.py:353-404, one iteration ----
    def train_view(self, view, gt, kf_id=0, kf_ids=None):
        """render -> loss -> backward -> score bookkeeping -> weighting_grad -> masked Adam.  Returns the loss
        tensor (device).  With world_size > 1 every rank calls this with its own view of the window; kf_ids = the
        keyframe ids of ALL ranks' views of this step in rank order (gathered if not given)."""
        if self.fused:
            return self._train_view_fused(view, gt, kf_id, kf_ids)
        dp = self.world_size > 1
        if dp:
            self._fused.zero_()
            for k in GROUPS:
                getattr(self, k).grad = self._grad_views[k]
        pred = self.render(view)
        total_loss = get_loss(self.weights, pred, gt, self.window)
        total_loss.backward()
        cur = self._zeros.grad.detach()
        radii = pred["radii"]
        if dp:
            # per-view weighting with this view's scores, applied locally before the reduce (section 8e)
            self._weighting_grad(cur, torch.clamp(self._global_scores[:, 0] + cur[:, 0], 0, 1e4))
            self._score_view.copy_(cur)
            self._vis_view.copy_(radii > 0)
            torch.distributed.all_reduce(self._fused, group=self.pg)
            # add_records is order dependent (max of the second score, first arg-max of the first): every rank applies
            # it for every view in rank order, so the replicas -- _globalkf_id included -- stay identical
            curs = [torch.empty_like(cur) for _ in range(self.world_size)]
            torch.distributed.all_gather(curs, cur.contiguous(), group=self.pg)
            for q, kf in enumerate(self._all_kf_ids(kf_id, kf_ids)):
                self._add_records(curs[q], kf)
            visible = self._vis_view > 0
        else:
            self._add_records(cur, kf_id)
            self._weighting_grad(cur, self._global_scores[:, 0])
            visible = radii > 0
        visible = visible & ~self._stable_mask   # radii[_stable_mask] = 0  (gaussian_base.py:401)
        self.optimizer.step(visible, self.P)
        if dp:
            pass  # grads live in the fused buffer, zeroed at the next call
        else:
            self.optimizer.zero_grad()
        return total_loss.detach()

    @torch.no_grad()
    def _train_view_fused(self, view, gt, kf_id, kf_ids=None):
        """The same iteration in 21 launches of the library: activate -> rasterizer forward -> fused loss -> backward blend
        -> one per-surfel kernel (backward preprocess + update); with world_size > 1 that kernel stages the view's rows
        and the reduce / post kernels of the peer-memory exchange follow."""
        mp, ws = self.mod.mapping, self._ws
        prm = [getattr(self, k).data for k in GROUPS]
        st = [self.optimizer.state[getattr(self, k)] for k in GROUPS]
        m, v = [x["exp_avg"] for x in st], [x["exp_avg_sq"] for x in st]
        lr = [g["lr"] for g in self.optimizer.param_groups]
        eps = self.optimizer.param_groups[0]["eps"]
        opac, scales, rot = mp.activate(prm[2], prm[3], prm[4], out=self._act)
        raster = ws.forward_async
        while True:
            raster(prm[0], scales, rot, opac, prm[1], self.bg, view["viewmatrix"], view["projmatrix"])
            loss, d_color, d_allmap = mp.map_loss(ws.color, ws.allmap, gt["rgb"], gt["depth"], gt["depth_cov"],
                                                  view["w2c"], self.cx, self.cy, self.fx, self.fy, self.weights,
                                                  ws=ws.loss_ws, out=ws.loss_out)
            # the only host wait of the view: the pair count, copied back right behind the preprocess; binning, forward
            # blend and loss are already queued, so the GPU has work while the host confirms that the kept scratch was
            # large enough.  False = it was not (it has been grown): run the forward again before anything is updated.
            if ws.confirm():
                break
            raster = ws.forward
        book = (self._stable_mask, self._local_scores, self._global_scores, self._globalkf_max_scores, self._globalkf_id)
        common = (d_color, d_allmap, prm[0], scales, rot, self.bg, view["viewmatrix"], view["projmatrix"],
                  view["tanfovx"], view["tanfovy"], prm, m, v, lr)
        if self.world_size > 1 and self.exchange == "p2p":
            mp.backward_update(ws, 1, *common, *book, kf_id, dp=self._dpx.handle, eps=eps)
            self._dpx.step(ws, prm, m, v, lr, *book, self._all_kf_ids(kf_id, kf_ids), eps=eps, staged=True)
        elif self.world_size > 1:
            ws.backward(d_color, d_allmap, prm[0], scales, rot, self.bg, view["viewmatrix"], view["projmatrix"],
                        view["tanfovx"], view["tanfovy"])
            self._nccl_step(mp, ws, prm, m, v, lr, book, self._all_kf_ids(kf_id, kf_ids), eps)
        else:
            # backward blend + ONE per-surfel pass (backward preprocess, activations' backward, records, weighting, Adam)
            mp.backward_update(ws, 0, *common, *book, kf_id, eps=eps)
        self.last_radii = ws.radii
        return loss[0]

    def _nccl_step(self, mp, ws, prm, m, v, lr, book, kf_ids, eps):
        """Fallback exchange without peer memory: DP_PRE -> dense sum all-reduce -> DP_POST, with add_records' two
        order-dependent quantities (max of score 1; first arg-max of score 0 and its keyframe id) from one max
        all-reduce of order-preserving integer keys (non-negative floats compare like their bit patterns)."""
        dist = torch.distributed
        N, r, P = self.world_size, self.rank, self.P
        mp.surfel_update(mp.UPDATE_DP_PRE, prm, m, v, lr, ws, *book, kf_ids[r], fused=self._fused, eps=eps)
        cur = ws.scores
        keys = torch.stack(((cur[:, 0].contiguous().view(torch.int32).to(torch.int64) << 32) | (N - 1 - r),
                            cur[:, 1].contiguous().view(torch.int32).to(torch.int64)))
        dist.all_reduce(self._fused, group=self.pg)
        dist.all_reduce(keys, op=dist.ReduceOp.MAX, group=self.pg)
        self._score_view[:, 1] = (keys[1] & 0xffffffff).to(torch.int32).view(torch.float32)   # max, not sum
        kfm0, kid0 = self._globalkf_max_scores.clone(), self._globalkf_id.clone()
        mp.surfel_update(mp.UPDATE_DP_POST, prm, m, v, lr, ws, *book, kf_ids[r], fused=self._fused, eps=eps)
        s0_max = (keys[0] >> 32).to(torch.int32).view(torch.float32)
        arg = (N - 1) - (keys[0] & 0xffffffff)                 # lowest rank among equal maxima = first in rank order
        kf_arg = torch.tensor(kf_ids, dtype=torch.int64, device=self.device)[arg]
        replace = kfm0 < s0_max
        self._globalkf_max_scores.copy_(torch.where(replace, s0_max, kfm0))
        self._globalkf_id.copy_(torch.where(replace, kf_arg, kid0))

    def _add_records(self, cur, kf_id):
        """gaussian_model.py:163-168 and gaussian_base.py:388-392"""
        self._local_scores[:, 0] += cur[:, 0]
        self._global_scores[:, 0] += cur[:, 0]
        large = cur[:, 1] > self._local_scores[:, 1]
        self._local_scores[:, 1] = torch.where(large, cur[:, 1], self._local_scores[:, 1])
        self._global_scores = torch.clamp(self._global_scores, 0, 1e4)
        replace = self._globalkf_max_scores < cur[:, 0]
        self._globalkf_max_scores = torch.where(replace, cur[:, 0], self._globalkf_max_scores)
        self._globalkf_id = torch.where(replace, torch.full_like(self._globalkf_id, kf_id), self._globalkf_id)

    def _weighting_grad(self, cur, global_s0):
        """gaussian_utils.py:56-62; global_s0 already contains this view's score (add_records runs first)"""
        gw = (cur[:, 0] / (global_s0 + 1e-6 + cur[:, 0])).unsqueeze(1)
        for k in GROUPS:
            getattr(self, k).grad.mul_(gw)


def view_to_device(view, device, non_blocking=False):
    """synthetic.View (CPU) -> the dict render() takes; 3 small H2D copies + the w2c pose."""
    w2c = torch.linalg.inv(view.c2w).float()
    return dict(viewmatrix=view.viewmatrix.to(device, non_blocking=non_blocking),
                projmatrix=view.projmatrix.to(device, non_blocking=non_blocking),
                campos=view.campos.to(device, non_blocking=non_blocking), w2c=w2c.to(device, non_blocking=non_blocking),
                tanfovx=view.tanfovx, tanfovy=view.tanfovy)
