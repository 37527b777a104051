"""Fused mapping-step building blocks around the rasterizer.

The drop-in names (``GaussianRasterizer`` ...) keep the reference's autograd contract.  A mapper that wants the
whole per-view iteration of ``train_once_gaussian`` (scripts/gaussian/gaussian_base.py:353-404) on the GPU
without ~400 torch launches uses these instead:

* ``activate``        raw parameters -> rasterizer inputs          (gaussian_base.py:160-166)
* ``RasterWorkspace`` the rasterizer's forward / backward on persistent, caller-owned buffers (no autograd)
* ``map_loss``        ``get_loss`` + ``render_raw`` post-processing + their backward (loss_utils.py:110-153)
* ``surfel_update``   activation backward + ``add_records`` + ``weighting_grad`` + masked Adam, one kernel

All of them are thin ctypes calls into ``libsurfel_b200.so``; there is no CPU path.
"""
import ctypes
import os

import torch

from . import _check, _f32c, _load, _ptr, _stream, _vp, last_counters

UPDATE_LOCAL, UPDATE_DP_PRE, UPDATE_DP_POST = 0, 1, 2
GROUP_COLS = (3, 3, 1, 2, 4)  # xyz, rgb, opacity, scaling, rotation


def activate(opacity_raw, scaling_raw, rotation_raw, out=None):
    """(sigmoid(opacity_raw) [P,1], exp(scaling_raw) [P,2], normalize(rotation_raw) [P,4])"""
    lib = _load()
    o, s, r = _f32c(opacity_raw, "opacity_raw"), _f32c(scaling_raw, "scaling_raw"), _f32c(rotation_raw, "rotation_raw")
    P = o.shape[0]
    if out is None:
        out = (torch.empty_like(o), torch.empty_like(s), torch.empty_like(r))
    _check(lib.sb200_activate(P, _ptr(o), _ptr(s), _ptr(r), _ptr(out[0]), _ptr(out[1]), _ptr(out[2]),
                              _stream(o.device)), "sb200_activate")
    return out


def map_loss(color, allmap, gt_rgb, gt_depth, gt_depth_cov, w2c, cx, cy, fx, fy, weights, ws=None, out=None):
    """Returns (loss[7], dL_dcolor[3,H,W], dL_dallmap[7,H,W]); loss[0] is the total (device tensor, no sync).

    weights: dict with rgb_loss, normal_loss, alpha_loss, depth_loss, dist_loss (training_args.loss_weights).
    (cx, cy, fx, fy): column/row principal point and focal lengths = the reference's (cv, cu, fv, fu)."""
    lib = _load()
    color, allmap = _f32c(color, "color"), _f32c(allmap, "allmap")
    gt_rgb, gt_depth, gt_cov = _f32c(gt_rgb, "gt_rgb"), _f32c(gt_depth, "gt_depth"), _f32c(gt_depth_cov, "gt_depth_cov")
    w2c = _f32c(w2c, "w2c")
    H, W = color.shape[-2], color.shape[-1]
    if allmap.shape != (7, H, W) or gt_rgb.shape != (3, H, W) or gt_depth.numel() != H * W or gt_cov.numel() != H * W:
        raise RuntimeError("map_loss: shape mismatch")
    if w2c.numel() != 16:
        raise RuntimeError("map_loss: w2c must be a 4x4 pose")
    dev = color.device
    if ws is None:
        ws = torch.empty(lib.sb200_map_loss_bytes(W, H), dtype=torch.uint8, device=dev)
    if out is None:
        out = (torch.empty(7, dtype=torch.float32, device=dev), torch.empty_like(color), torch.empty_like(allmap))
    loss, d_color, d_allmap = out
    _check(lib.sb200_map_loss(W, H, _ptr(color), _ptr(allmap), _ptr(gt_rgb), _ptr(gt_depth), _ptr(gt_cov), _ptr(w2c),
                              float(cx), float(cy), float(fx), float(fy), float(weights["rgb_loss"]),
                              float(weights["normal_loss"]), float(weights["alpha_loss"]),
                              float(weights["depth_loss"]), float(weights["dist_loss"]), _ptr(ws), _ptr(loss),
                              _ptr(d_color), _ptr(d_allmap), _stream(dev)), "sb200_map_loss")
    return loss, d_color, d_allmap


class RasterWorkspace:
    """Persistent buffers for one (P, W, H): the rasterizer's forward and backward without autograd and without
    per-call allocation.  Only the sort scratch and the sorted id list depend on the data (R); they grow
    geometrically and are kept."""

    def __init__(self, P, W, H, device, capacity=None):
        lib = _load()
        self.W, self.H, self.device = W, H, device
        u8 = lambda n: torch.empty(n, dtype=torch.uint8, device=device)
        f = lambda *s: torch.empty(s, dtype=torch.float32, device=device)
        self.image = u8(lib.sb200_image_bytes(W, H))
        self.loss_ws = u8(lib.sb200_map_loss_bytes(W, H))
        self.binning = u8(0)
        self.point_list = torch.empty(0, dtype=torch.uint8, device=device)   # sorted ids + hit masks
        self.pair_capacity = 0   # pairs the binning scratch and point_list hold
        self._pending = None     # capacity of an unconfirmed forward_async
        # hand-off state of the sync-free forward, owned by THIS workspace: a pinned landing pad for the pair count
        # and the event recorded behind its copy (forwards of other workspaces / streams cannot disturb them)
        self._count = torch.zeros(1, dtype=torch.int64).pin_memory()
        self._sync = lib.sb200_sync_create()
        if not self._sync:
            raise RuntimeError("sb200_sync_create failed: %s" % lib.sb200_last_error().decode())
        self.color, self.allmap = f(3, H, W), f(7, H, W)
        self.loss_out = (f(7), f(3, H, W), f(7, H, W))
        self.pixel_mask = torch.ones(H * W, dtype=torch.bool, device=device)
        self.R = 0
        self.R_seen = 0          # pair count of the last confirmed forward (diagnostics)
        self.capacity = 0
        self._alloc_rows(max(int(capacity or 0), P, 1))
        self.set_P(P)

    def __del__(self):
        try:
            if getattr(self, "_sync", None):
                _load().sb200_sync_destroy(self._sync)
                self._sync = None
        except Exception:
            pass

    def _alloc_rows(self, cap):
        """P-sized buffers for up to ``cap`` surfels (the layouts are monotone in P, so any P <= cap fits)."""
        lib = _load()
        dev = self.device
        f = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
        self.capacity = cap
        self.geom = torch.empty(lib.sb200_geom_bytes(cap), dtype=torch.uint8, device=dev)
        self.grad_ws = torch.zeros(lib.sb200_grad_bytes(cap), dtype=torch.uint8, device=dev)
        self.grad_zero = True    # True while ALL of grad_ws is known to be zero (backward_update leaves it so)
        self._rows = dict(radii=torch.empty(cap, dtype=torch.int32, device=dev), d_means3D=f(cap, 3), d_means2D=f(cap, 3),
                          d_colors=f(cap, 3), d_opacity=f(cap, 1), d_scales=f(cap, 2), d_rot=f(cap, 4), scores=f(cap, 2),
                          act_opacity=f(cap, 1), act_scales=f(cap, 2), act_rot=f(cap, 4))

    def set_P(self, P):
        """Surfel count of the next forward (changes with every prune / densify); grows geometrically, never shrinks."""
        if P > self.capacity:
            cap = self.capacity
            while cap < P:
                cap *= 2
            self._alloc_rows(cap)
        self.P = P
        for name, buf in self._rows.items():
            setattr(self, name, buf[:P])

    def forward(self, means3D, scales, rotations, opacities, colors, bg, viewmatrix, projmatrix):
        lib = _load()
        P, W, H, dev = self.P, self.W, self.H, self.device
        stream = _stream(dev)
        count = self._count
        _check(lib.sb200_forward_begin(P, W, H, _ptr(means3D), _ptr(scales), _ptr(rotations), _ptr(opacities),
                                       _ptr(colors), _ptr(viewmatrix), _ptr(projmatrix), _ptr(self.geom),
                                       _ptr(self.radii), _vp(count.data_ptr()), stream), "sb200_forward_begin")
        R = int(count[0])
        self.R = self.R_seen = R
        last_counters.update(P=P, R=R, W=W, H=H)
        if R > self.pair_capacity:
            self._grow_pairs(R)
        _check(lib.sb200_forward_finish(P, W, H, R, _ptr(bg), _ptr(self.geom), _vp(self.binning.data_ptr()),
                                        _vp(self.point_list.data_ptr()), _ptr(self.image), _ptr(self.color),
                                        _ptr(self.allmap), stream), "sb200_forward_finish")
        return self.color, self.radii, self.allmap

    def forward_async(self, means3D, scales, rotations, opacities, colors, bg, viewmatrix, projmatrix):
        """The same forward without a host synchronisation: the binning scratch kept from earlier views is used as
        is and the pair count stays on the device.  ``confirm()`` must be called before the view's results are
        committed; it returns False when the scratch was too small (then it has been grown: run ``forward`` again).
        The first call (no scratch yet) falls back to the synchronising ``forward``."""
        lib = _load()
        P, W, H, dev = self.P, self.W, self.H, self.device
        cap = self.pair_capacity
        if cap < 1:
            self._pending = None
            return self.forward(means3D, scales, rotations, opacities, colors, bg, viewmatrix, projmatrix)
        _check(lib.sb200_forward_async(P, W, H, _ptr(means3D), _ptr(scales), _ptr(rotations), _ptr(opacities),
                                       _ptr(colors), _ptr(viewmatrix), _ptr(projmatrix), _ptr(bg), _ptr(self.geom),
                                       _ptr(self.radii), _vp(self.binning.data_ptr()), cap,
                                       _vp(self.point_list.data_ptr()), _ptr(self.image), _ptr(self.color),
                                       _ptr(self.allmap), _vp(self._count.data_ptr()), _vp(self._sync), _stream(dev)),
               "sb200_forward_async")
        self.R = cap  # what sb200_backward is told; the live count stays on the device
        self._pending = cap
        return self.color, self.radii, self.allmap

    def confirm(self):
        """Wait for the pair count of the last ``forward_async`` (long since copied when called late in the view) and
        check it against the scratch capacity."""
        if self._pending is None:
            return True
        if os.environ.get("SB200_DIAG_NO_CONFIRM"):   # diagnostic only: never wait for the pair count
            self._pending = None
            return True
        cap = self._pending
        self._pending = None
        _check(_load().sb200_forward_wait(_vp(self._sync)), "sb200_forward_wait")
        R = int(self._count[0])
        self.R_seen = R
        last_counters.update(P=self.P, R=R, W=self.W, H=self.H)
        if R <= cap:
            return True
        self._grow_pairs(R)
        return False

    def _grow_pairs(self, R):
        lib = _load()
        want = int(R * 1.25) + 4096
        self.binning = torch.empty(lib.sb200_binning_bytes(want), dtype=torch.uint8, device=self.device)
        self.point_list = torch.empty(lib.sb200_pair_list_bytes(want, self.W, self.H), dtype=torch.uint8,
                                      device=self.device)
        self.pair_capacity = want

    def sorted_ids(self, R=None):
        """The first R entries of the pair list: surfel ids sorted by (tile, depth)."""
        R = self.R_seen if R is None else R
        return self.point_list[:4 * R].view(torch.int32)

    def backward(self, d_color, d_allmap, means3D, scales, rotations, bg, viewmatrix, projmatrix, tanfovx, tanfovy,
                 pixel_mask=None):
        lib = _load()
        mask = self.pixel_mask if pixel_mask is None else pixel_mask
        self.grad_zero = False   # sb200_backward leaves the accumulation records in place
        _check(lib.sb200_backward(self.P, self.W, self.H, self.R, _ptr(bg), float(tanfovx), float(tanfovy),
                                  _ptr(means3D), _ptr(scales), _ptr(rotations), _ptr(viewmatrix), _ptr(projmatrix),
                                  _ptr(self.radii), _ptr(self.geom), _vp(self.point_list.data_ptr()), _ptr(self.image),
                                  _ptr(self.color), _ptr(self.allmap), _ptr(d_color), _ptr(d_allmap), _ptr(mask),
                                  _ptr(self.grad_ws), _ptr(self.d_means3D), _ptr(self.d_means2D), _ptr(self.d_colors),
                                  _ptr(self.d_opacity), _ptr(self.d_scales), _ptr(self.d_rot), _vp(None),
                                  _ptr(self.scores), _stream(self.device)), "sb200_backward")


def backward_update(ws, mode, d_color, d_allmap, means3D, scales, rotations, bg, viewmatrix, projmatrix, tanfovx, tanfovy,
                    params, exp_avg, exp_avg_sq, lr, stable_mask, local_scores, global_scores, kf_max_scores, kf_id, kf,
                    dp=None, pixel_mask=None, betas=(0.9, 0.999), eps=1e-15, reuse_records=False):
    """sb200_backward_update (include/surfel_b200.h): the rasterizer's backward and the per-surfel update of the
    mapping step as one call, with no dense gradient arrays in between.  mode 0 = single-GPU update, mode 1 = stage
    rows + signal of the multi-GPU exchange (dp: the rank's exchange handle, PeerExchange.handle)."""
    lib = _load()
    mask = ws.pixel_mask if pixel_mask is None else pixel_mask
    lr_arr = (ctypes.c_float * 5)(*[float(x) for x in lr])
    null = _vp(None)
    if not reuse_records:
        # the call hands grad_ws back all zero, so the next one needs no memset (SB200_BWD_RECORDS_ZERO); after anything
        # else wrote records (sb200_backward leaves them in place) the WHOLE capacity is cleared once: rows beyond the
        # current P come into use when the map grows
        if not ws.grad_zero:
            ws.grad_ws.zero_()
        mode = int(mode) | 4
        ws.grad_zero = True
    _check(lib.sb200_backward_update(
        ws.P, ws.W, ws.H, -1 if reuse_records else ws.R, _ptr(bg), float(tanfovx), float(tanfovy), _ptr(means3D),
        _ptr(scales), _ptr(rotations), _ptr(viewmatrix), _ptr(projmatrix), _ptr(ws.radii), _ptr(ws.geom),
        _vp(ws.point_list.data_ptr()), _ptr(ws.image), _ptr(ws.color), _ptr(ws.allmap), _ptr(d_color), _ptr(d_allmap),
        _ptr(mask), _ptr(ws.grad_ws), int(mode), _ptr_array(params),
        _ptr_array(exp_avg) if exp_avg is not None else null, _ptr_array(exp_avg_sq) if exp_avg_sq is not None else null,
        _ptr(stable_mask), _ptr(local_scores), _ptr(global_scores), _ptr(kf_max_scores), _ptr(kf_id), int(kf), lr_arr,
        float(betas[0]), float(betas[1]), float(eps), _vp(dp) if dp else null, _stream(ws.device)),
        "sb200_backward_update")


def _ptr_array(tensors):
    arr = (ctypes.c_void_p * 5)()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr()
    return arr


def surfel_update(mode, params, exp_avg, exp_avg_sq, lr, ws, stable_mask, local_scores, global_scores, kf_max_scores,
                  kf_id, kf, fused=None, betas=(0.9, 0.999), eps=1e-15):
    """One fused per-surfel pass after ``RasterWorkspace.backward`` (see include/surfel_b200.h).

    params / exp_avg / exp_avg_sq: 5 contiguous CUDA tensors (xyz, rgb, opacity, scaling, rotation; raw values);
    ws: the RasterWorkspace holding this view's gradients, scores and radii."""
    lib = _load()
    P = params[0].shape[0]
    for grp in (params, exp_avg, exp_avg_sq):
        for t, nc in zip(grp, GROUP_COLS):
            if not (t.is_cuda and t.is_contiguous() and t.dtype == torch.float32 and t.numel() == P * nc):
                raise RuntimeError("surfel_update needs contiguous float32 CUDA tensors of the five groups")
    if kf_id.dtype != torch.int64:
        raise RuntimeError("kf_id must be int64")
    lr_arr = (ctypes.c_float * 5)(*[float(x) for x in lr])
    _check(lib.sb200_surfel_update(P, int(mode), _ptr_array(params), _ptr_array(exp_avg), _ptr_array(exp_avg_sq),
                                   _ptr(ws.d_means3D), _ptr(ws.d_colors), _ptr(ws.d_opacity), _ptr(ws.d_scales),
                                   _ptr(ws.d_rot), _ptr(ws.scores), _ptr(ws.radii), _ptr(stable_mask),
                                   _ptr(local_scores), _ptr(global_scores), _ptr(kf_max_scores), _ptr(kf_id),
                                   int(kf), _ptr(fused) if fused is not None else _vp(None), lr_arr,
                                   float(betas[0]), float(betas[1]), float(eps), _stream(params[0].device)),
           "sb200_surfel_update")


class PeerExchange:
    """The multi-GPU exchange of one optimizer step as three kernels over NVLink peer memory (csrc/dp_exchange.cu):
    every rank stages the rows of the surfels ITS view saw, reduces the slice of surfels it owns by loading its
    peers' rows, and every rank then applies bookkeeping + Adam from the owners' result rows.  The kernels hand over to
    each other through flags inside the peer-mapped buffers; PyTorch's symmetric memory only provides the allocation
    (one buffer per rank, mapped into every process) -- no barrier and no collective is called per step."""

    def __init__(self, capacity, group, device):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        lib = _load()
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        if self.world > 16:
            raise RuntimeError("PeerExchange supports up to 16 ranks")
        self.capacity, self.device = int(capacity), device
        total = lib.sb200_dp_buffer_bytes(self.capacity, self.world)
        self.buf = symm_mem.empty(total, dtype=torch.uint8, device=device)
        self.buf.zero_()
        self.hdl = symm_mem.rendezvous(self.buf, self.group)
        bases = [int(p) for p in self.hdl.buffer_ptrs]
        if len(bases) != self.world or not all(bases):
            raise RuntimeError("symmetric memory rendezvous returned no peer pointers")
        self.handle = lib.sb200_dp_create(self.rank, self.world, self.capacity, (ctypes.c_void_p * self.world)(*bases))
        if not self.handle:
            raise RuntimeError("sb200_dp_create failed: %s" % lib.sb200_last_error().decode())
        torch.cuda.synchronize(device)
        self.hdl.barrier(channel=0)   # once: every rank has zeroed its buffer (flags included) before anyone signals

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                _load().sb200_dp_destroy(self.handle)
                self.handle = None
        except Exception:
            pass

    def slice_rows(self, P):
        """Rows per owner for P surfels (the same on every rank)."""
        return int(_load().sb200_dp_slice(_vp(self.handle), int(P)))

    def wait_ms(self):
        """(ms waited before reduce, ms waited before post) by this rank's first CTAs since the last call: rank skew."""
        out = (ctypes.c_uint64 * 2)()
        _check(_load().sb200_dp_wait_ns(_vp(self.handle), out, _stream(self.device)), "sb200_dp_wait_ns")
        return out[0] * 1e-6, out[1] * 1e-6

    def check(self):
        """Raises if a kernel of this rank ever gave up waiting for a peer (synchronises the stream)."""
        _check(_load().sb200_dp_error(_vp(self.handle), _stream(self.device)), "multi-GPU exchange")

    def step(self, ws, params, exp_avg, exp_avg_sq, lr, stable_mask, local_scores, global_scores, kf_max_scores, kf_id,
             kf_ids, betas=(0.9, 0.999), eps=1e-15, staged=False):
        """staged=True: this view's stage rows were already written (and signalled) by backward_update(mode=1)."""
        lib = _load()
        P, dev = params[0].shape[0], self.device
        st = _stream(dev)
        h = _vp(self.handle)
        if not staged:
            _check(lib.sb200_dp_stage(h, P, _ptr(ws.d_means3D), _ptr(ws.d_colors), _ptr(ws.d_opacity), _ptr(ws.d_scales),
                                      _ptr(ws.d_rot), _ptr(ws.scores), _ptr(ws.radii), _ptr(params[2]), _ptr(params[3]),
                                      _ptr(params[4]), st), "sb200_dp_stage")
        kf_arr = (ctypes.c_int64 * self.world)(*[int(k) for k in kf_ids])
        _check(lib.sb200_dp_reduce(h, P, _ptr(global_scores), st), "sb200_dp_reduce")
        lr_arr = (ctypes.c_float * 5)(*[float(x) for x in lr])
        _check(lib.sb200_dp_post(h, P, kf_arr, _ptr_array(params), _ptr_array(exp_avg), _ptr_array(exp_avg_sq),
                                 _ptr(stable_mask), _ptr(local_scores), _ptr(global_scores), _ptr(kf_max_scores),
                                 _ptr(kf_id), lr_arr, float(betas[0]), float(betas[1]), float(eps), st), "sb200_dp_post")
