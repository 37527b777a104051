"""B200-native drop-in for VINGS-Mono's ``diff_surfel_rasterization`` module.

Same public names and call contract as the reference wrapper
(submodules/diff-surfel-rasterization/diff_surfel_rasterization/__init__.py):

* ``GaussianRasterizationSettings``  (:171-184)  -- 13 required fields, no defaults
* ``GaussianRasterizer``             (:186-237)  -- ``forward(...) -> (color, radii, allmap)``, ``markVisible``
* ``SparseGaussianAdam``             (:241-268)  -- visibility-masked Adam, ``step(visibility, N)``
* ``rasterize_gaussians`` / ``_RasterizeGaussians`` (:21-169) -- autograd contract; ``scores`` is delivered
  through the gradient slot of the ``scores`` input.

PyTorch is plumbing only (device memory, streams, autograd graph).  All compute is hand-written
sm_100a CUDA in ``libsurfel_b200.so`` reached through the C ABI of ``include/surfel_b200.h`` with
ctypes.  There is no CPU path and no fallback: importing works without a GPU (so host logic can be
tested), but any compute call without the library or without CUDA tensors raises.
"""
import ctypes
import os
from typing import NamedTuple

import torch
import torch.nn as nn

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.environ.get("SB200_LIB") or os.path.join(_HERE, "libsurfel_b200.so")  # SB200_LIB: debug variants
_lib = None

_vp = ctypes.c_void_p
_i32 = ctypes.c_int
_i64 = ctypes.c_int64
_f32 = ctypes.c_float
_u32 = ctypes.c_uint32
_sz = ctypes.c_size_t


def _load():
    """Load the C-ABI library once and declare every prototype of include/surfel_b200.h."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise RuntimeError(
            "libsurfel_b200.so is missing (%s): build it with `python vings-mono_b200/build.py`; "
            "there is no CPU fallback" % _LIB_PATH)
    lib = ctypes.CDLL(_LIB_PATH)
    lib.sb200_abi_version.restype = _i32
    lib.sb200_last_error.restype = ctypes.c_char_p
    for name in ("sb200_geom_bytes", "sb200_binning_bytes", "sb200_grad_bytes"):
        getattr(lib, name).restype = _sz
        getattr(lib, name).argtypes = [_i64]
    lib.sb200_image_bytes.restype = _sz
    lib.sb200_image_bytes.argtypes = [_i32, _i32]
    lib.sb200_pair_list_bytes.restype = _sz
    lib.sb200_pair_list_bytes.argtypes = [_i64, _i32, _i32]
    lib.sb200_geom_layout.argtypes = [_i64, ctypes.POINTER(_sz)]
    lib.sb200_geom_layout.restype = None
    lib.sb200_image_layout.argtypes = [_i32, _i32, ctypes.POINTER(_sz)]
    lib.sb200_image_layout.restype = None
    lib.sb200_binning_layout.argtypes = [_i64, ctypes.POINTER(_sz)]
    lib.sb200_binning_layout.restype = None
    lib.sb200_forward_begin.restype = _i32
    lib.sb200_forward_begin.argtypes = [_i32, _i32, _i32] + [_vp] * 7 + [_vp, _vp, _vp, _vp]
    lib.sb200_forward_finish.restype = _i32
    lib.sb200_forward_finish.argtypes = [_i32, _i32, _i32, _i64] + [_vp] * 8
    lib.sb200_forward_async.restype = _i32
    lib.sb200_forward_async.argtypes = [_i32, _i32, _i32] + [_vp] * 8 + [_vp, _vp, _vp, _i64] + [_vp] * 7
    lib.sb200_forward_wait.restype = _i32
    lib.sb200_forward_wait.argtypes = [_vp]
    lib.sb200_sync_create.restype = _vp
    lib.sb200_sync_create.argtypes = []
    lib.sb200_sync_destroy.restype = None
    lib.sb200_sync_destroy.argtypes = [_vp]
    lib.sb200_forward_begin_precomp.restype = _i32
    lib.sb200_forward_begin_precomp.argtypes = [_i32, _i32, _i32] + [_vp] * 6 + [_vp, _vp, _vp, _vp]
    lib.sb200_backward_precomp.restype = _i32
    lib.sb200_backward_precomp.argtypes = [_i32, _i32, _i32, _i64, _vp, _f32, _f32] + [_vp] * 22
    lib.sb200_sh_forward.restype = _i32
    lib.sb200_sh_forward.argtypes = [_i32, _i32, _i32] + [_vp] * 6
    lib.sb200_sh_backward.restype = _i32
    lib.sb200_sh_backward.argtypes = [_i32, _i32, _i32] + [_vp] * 9
    lib.sb200_backward.restype = _i32
    lib.sb200_backward.argtypes = [_i32, _i32, _i32, _i64, _vp, _f32, _f32] + [_vp] * 24
    lib.sb200_mark_visible.restype = _i32
    lib.sb200_mark_visible.argtypes = [_i32, _vp, _vp, _vp, _vp, _vp]
    lib.sb200_adam_update.restype = _i32
    lib.sb200_adam_update.argtypes = [_vp, _vp, _vp, _vp, _vp, _f32, _f32, _f32, _f32, _u32, _u32, _vp]
    lib.sb200_sort_pairs.restype = _i32
    lib.sb200_sort_pairs.argtypes = [_vp, _i64, _i32, _i32, _vp, _vp, _vp]
    lib.sb200_debug_walk_cap.restype = _i32
    lib.sb200_debug_walk_cap.argtypes = [_vp]
    lib.sb200_launch_count.restype = ctypes.c_longlong
    lib.sb200_profile_enable.restype = _i32
    lib.sb200_profile_enable.argtypes = [_i32]
    lib.sb200_profile_read.restype = _i32
    lib.sb200_profile_read.argtypes = [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_longlong)]
    lib.sb200_map_loss_bytes.restype = _sz
    lib.sb200_map_loss_bytes.argtypes = [_i32, _i32]
    lib.sb200_map_loss.restype = _i32
    lib.sb200_map_loss.argtypes = [_i32, _i32] + [_vp] * 6 + [_f32] * 9 + [_vp] * 5
    lib.sb200_activate.restype = _i32
    lib.sb200_activate.argtypes = [_i32] + [_vp] * 7
    lib.sb200_surfel_update.restype = _i32
    lib.sb200_surfel_update.argtypes = ([_i32, _i32] + [_vp] * 3 + [_vp] * 5 + [_vp] * 7 + [_i64, _vp, _vp] +
                                        [_f32] * 3 + [_vp])
    lib.sb200_backward_update.restype = _i32
    lib.sb200_backward_update.argtypes = ([_i32, _i32, _i32, _i64, _vp, _f32, _f32] + [_vp] * 15 + [_i32] + [_vp] * 8 +
                                          [_i64, _vp] + [_f32] * 3 + [_vp] * 2)
    lib.sb200_dp_buffer_bytes.restype = _sz
    lib.sb200_dp_buffer_bytes.argtypes = [_i64, _i32]
    lib.sb200_dp_create.restype = _vp
    lib.sb200_dp_create.argtypes = [_i32, _i32, _i64, _vp]
    lib.sb200_dp_destroy.restype = None
    lib.sb200_dp_destroy.argtypes = [_vp]
    lib.sb200_dp_slice.restype = _i32
    lib.sb200_dp_slice.argtypes = [_vp, _i32]
    lib.sb200_dp_error.restype = _i32
    lib.sb200_dp_error.argtypes = [_vp, _vp]
    lib.sb200_dp_wait_ns.restype = _i32
    lib.sb200_dp_wait_ns.argtypes = [_vp, ctypes.POINTER(ctypes.c_uint64), _vp]
    lib.sb200_dp_stage.restype = _i32
    lib.sb200_dp_stage.argtypes = [_vp, _i32] + [_vp] * 11
    lib.sb200_dp_reduce.restype = _i32
    lib.sb200_dp_reduce.argtypes = [_vp, _i32, _vp, _vp]
    lib.sb200_dp_post.restype = _i32
    lib.sb200_dp_post.argtypes = [_vp, _i32] + [_vp] * 10 + [_f32] * 3 + [_vp]
    lib.sb200_knn_mean_dist2.restype = _i32
    lib.sb200_knn_mean_dist2.argtypes = [_i32, _vp, _vp, _vp]
    lib.sb200_prune_scratch_bytes.restype = _sz
    lib.sb200_prune_scratch_bytes.argtypes = [_i64]
    lib.sb200_prune_rows.restype = _i32
    lib.sb200_prune_rows.argtypes = [_i32, _vp, _i32, _vp, _vp, _vp, _vp, _vp, _vp]
    lib.sb200_prune_rows_static.restype = _i32   # no caller scratch, device-side count (include/surfel_b200.h)
    lib.sb200_prune_rows_static.argtypes = [_i32, _vp, _i32, _vp, _vp, _vp, _vp, _vp]
    if lib.sb200_abi_version() != 7:
        raise RuntimeError("libsurfel_b200.so ABI mismatch")
    _lib = lib
    return lib


def _check(rc, what):
    if rc != 0:
        raise RuntimeError("%s failed: %s" % (what, _load().sb200_last_error().decode()))


def _ptr(t):
    return _vp(t.data_ptr()) if t is not None and t.numel() > 0 else _vp(None)


def _stream(device):
    return _vp(torch.cuda.current_stream(device).cuda_stream)


def _f32c(t, name):
    """The reference only asserts is_cuda (rasterize_points.cu:28-30) and calls .contiguous()."""
    if not t.is_cuda:
        raise RuntimeError("%s must be a CUDA tensor" % name)
    if t.dtype != torch.float32:
        raise RuntimeError("%s must be float32" % name)
    return t.contiguous()


class _Workspace:
    """Buffers one forward -> backward pair of the drop-in path needs besides its outputs: geometry, image and
    binning workspaces, the sorted id list, the gradient accumulation scratch, and the hand-off state of the
    sync-free forward (pinned pair-count pad + event).  Pooled per (device, stream) and leased to an autograd node
    for its lifetime, so a training loop allocates nothing per call after its second iteration (the reference
    resizes four byte tensors per call through its allocator callbacks, rasterize_points.cu:32-38)."""

    def __init__(self, device):
        lib = _load()
        self.device = device
        self.geom = self.image = self.grad_ws = self.binning = None
        self.point_list = None
        self.geom_P = self.grad_P = 0
        self.image_WH = None
        self.pair_capacity = 0
        self.count = torch.zeros(1, dtype=torch.int64).pin_memory()
        self.sync = lib.sb200_sync_create()
        if not self.sync:
            raise RuntimeError("sb200_sync_create failed: %s" % lib.sb200_last_error().decode())
        self.busy = False

    def __del__(self):
        try:
            if self.sync:
                _load().sb200_sync_destroy(self.sync)
                self.sync = None
        except Exception:
            pass

    def fit(self, P, W, H):
        lib = _load()
        u8 = lambda n: torch.empty(n, dtype=torch.uint8, device=self.device)
        if P > self.geom_P:   # the layouts are monotone in P: a larger buffer serves any smaller P
            cap = max(P, int(self.geom_P * 1.25))
            self.geom, self.geom_P = u8(lib.sb200_geom_bytes(cap)), cap
        if self.image_WH != (W, H):
            self.image, self.image_WH = u8(lib.sb200_image_bytes(W, H)), (W, H)
            self.pair_capacity = 0   # the pair list holds per-tile hit masks: sized for one image size

    def fit_grad(self, P):
        if P > self.grad_P:
            cap = max(P, int(self.grad_P * 1.25))
            self.grad_ws = torch.empty(_load().sb200_grad_bytes(cap), dtype=torch.uint8, device=self.device)
            self.grad_P = cap

    def grow_pairs(self, R):
        """Binning scratch and pair list (sorted ids + hit masks) for up to 1.25 R pairs at the current image size."""
        lib = _load()
        want = int(R * 1.25) + 4096
        W, H = self.image_WH
        self.binning = torch.empty(lib.sb200_binning_bytes(want), dtype=torch.uint8, device=self.device)
        self.point_list = torch.empty(lib.sb200_pair_list_bytes(want, W, H), dtype=torch.uint8, device=self.device)
        self.pair_capacity = want
        self.pair_WH = (W, H)


class _Lease:
    """Ties a pooled workspace to the autograd node that needs it in its backward; the workspace returns to the pool
    when the node dies (not at the first backward: retain_graph=True may run the backward again)."""

    def __init__(self, ws):
        self.ws = ws
        ws.busy = True

    def __del__(self):
        self.ws.busy = False


_pool = {}   # (device index, stream handle) -> [workspaces]


def _acquire(device, P, W, H):
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    lst = _pool.setdefault(key, [])
    ws = next((w for w in lst if not w.busy), None)
    if ws is None:
        ws = _Workspace(device)
        lst.append(ws)
    ws.fit(P, W, H)
    return _Lease(ws)


def release_workspaces():
    """Drop every pooled workspace that is not leased (frees their device memory)."""
    for key in list(_pool):
        _pool[key] = [w for w in _pool[key] if w.busy]


STAGES = ("preprocess", "duplicate", "sort", "ranges", "blend_fwd", "blend_bwd", "preprocess_bwd", "adam",
          "map_loss", "activate", "update", "churn", "surfel_sort", "dp_reduce", "dp_post")


def profile_enable(on=True):
    """Bracket every stage with CUDA events on its launching stream (include/surfel_b200.h)."""
    _check(_load().sb200_profile_enable(1 if on else 0), "sb200_profile_enable")


def profile_read():
    """{stage: (total_ms, launches)} since the previous read; waits for the recorded events."""
    ms = (ctypes.c_double * len(STAGES))()
    cnt = (ctypes.c_longlong * len(STAGES))()
    _check(_load().sb200_profile_read(ms, cnt), "sb200_profile_read")
    return {name: (ms[i], cnt[i]) for i, name in enumerate(STAGES)}


_walk_cap_keepalive = None


def debug_walk_cap(per_tile_cap):
    """Test hook (include/surfel_b200.h, sb200_debug_walk_cap): int32/uint32 CUDA tensor [tiles] or None."""
    global _walk_cap_keepalive
    _walk_cap_keepalive = per_tile_cap
    _check(_load().sb200_debug_walk_cap(_ptr(per_tile_cap) if per_tile_cap is not None else _vp(None)),
           "sb200_debug_walk_cap")


def launch_count():
    return int(_load().sb200_launch_count())


# counters of the most recent forward on this process: P, R (tile-surfel pairs); V is (radii > 0).sum()
last_counters = {}


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, scores, cov3Ds_precomp,
                raster_settings):
        lib = _load()
        rs = raster_settings
        if means3D.ndim != 2 or means3D.size(1) != 3:
            raise RuntimeError("means3D must have dimensions (num_points, 3)")  # rasterize_points.cu:63-65
        use_sh = sh is not None and sh.numel() != 0
        if use_sh and (sh.ndim != 3 or sh.size(0) != means3D.size(0) or sh.size(2) != 3):
            raise RuntimeError("shs must have dimensions (num_points, coefficients, 3)")
        use_pre = cov3Ds_precomp is not None and cov3Ds_precomp.numel() != 0
        if use_pre and cov3Ds_precomp.numel() != 9 * means3D.size(0):
            raise RuntimeError("cov3D_precomp (transMat_precomp) must have dimensions (num_points, 9)")
        P = means3D.size(0)
        H, W = int(rs.image_height), int(rs.image_width)
        dev = means3D.device
        ctx.in_shapes = [tuple(t.shape) if isinstance(t, torch.Tensor) else None
                         for t in (means3D, means2D, sh, colors_precomp, opacities, scales, rotations, scores,
                                   cov3Ds_precomp)] + [None]
        means3D = _f32c(means3D, "means3D")
        ctx.use_sh = use_sh and P > 0
        sh_c = clamped = campos = None
        if ctx.use_sh:
            # SH -> RGB as a stage in front of the rasterizer (reference: inside preprocessCUDA, forward.cu:237-242)
            sh_c = _f32c(sh, "shs")
            campos = _f32c(rs.campos, "campos")
            colors = torch.empty((P, 3), dtype=torch.float32, device=dev)
            clamped = torch.empty((P, 3), dtype=torch.uint8, device=dev)
            _check(lib.sb200_sh_forward(P, int(rs.sh_degree), sh_c.size(1), _ptr(means3D), _ptr(campos), _ptr(sh_c),
                                        _ptr(colors), _ptr(clamped), _stream(dev)), "sb200_sh_forward")
        else:
            colors = _f32c(colors_precomp, "colors_precomp")
        opac = _f32c(opacities, "opacities")
        ctx.use_pre = use_pre and P > 0
        # with a precomputed transform the reference's scales / rotations are empty tensors (__init__.py:216-222)
        tm_c = _f32c(cov3Ds_precomp, "cov3D_precomp") if ctx.use_pre else None
        scales_c = _f32c(scales, "scales")
        rot_c = _f32c(rotations, "rotations")
        bg = _f32c(rs.bg, "bg")
        view = _f32c(rs.viewmatrix, "viewmatrix")
        proj = _f32c(rs.projmatrix, "projmatrix")

        color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
        others = torch.empty((7, H, W), dtype=torch.float32, device=dev)
        radii = torch.empty((P,), dtype=torch.int32, device=dev)
        if P == 0:  # rasterize_points.cu:106 -- outputs stay zero, nothing is rendered
            color.zero_()
            others.zero_()
            ctx.empty = True
            ctx.raster_settings = rs
            ctx.P = 0
            return color, radii, others

        stream = _stream(dev)
        lease = _acquire(dev, P, W, H)
        ws = lease.ws

        def two_stage():
            """Waits for the pair count after the preprocess, sizes the scratch, then bins and blends."""
            if ctx.use_pre:
                _check(lib.sb200_forward_begin_precomp(P, W, H, _ptr(means3D), _ptr(tm_c), _ptr(opac), _ptr(colors),
                                                       _ptr(view), _ptr(proj), _ptr(ws.geom), _ptr(radii),
                                                       _vp(ws.count.data_ptr()), stream), "sb200_forward_begin_precomp")
            else:
                _check(lib.sb200_forward_begin(P, W, H, _ptr(means3D), _ptr(scales_c), _ptr(rot_c), _ptr(opac),
                                               _ptr(colors), _ptr(view), _ptr(proj), _ptr(ws.geom), _ptr(radii),
                                               _vp(ws.count.data_ptr()), stream), "sb200_forward_begin")
            R = int(ws.count[0])
            if R > ws.pair_capacity:
                ws.grow_pairs(R)
            _check(lib.sb200_forward_finish(P, W, H, R, _ptr(bg), _ptr(ws.geom), _vp(ws.binning.data_ptr()),
                                            _vp(ws.point_list.data_ptr()), _ptr(ws.image), _ptr(color), _ptr(others),
                                            stream), "sb200_forward_finish")
            return R

        if ctx.use_pre or ws.pair_capacity < 1:
            R = R_arg = two_stage()
        else:
            # sync-free forward on the scratch kept from earlier calls: the whole forward is queued at once and the
            # host only waits for the pair count (copied back right behind the preprocess) to confirm that the
            # scratch was large enough -- the GPU never idles while the host sizes buffers
            cap = ws.pair_capacity
            _check(lib.sb200_forward_async(P, W, H, _ptr(means3D), _ptr(scales_c), _ptr(rot_c), _ptr(opac), _ptr(colors),
                                           _ptr(view), _ptr(proj), _ptr(bg), _ptr(ws.geom), _ptr(radii),
                                           _vp(ws.binning.data_ptr()), cap, _vp(ws.point_list.data_ptr()),
                                           _ptr(ws.image), _ptr(color), _ptr(others), _vp(ws.count.data_ptr()),
                                           _vp(ws.sync), stream), "sb200_forward_async")
            _check(lib.sb200_forward_wait(_vp(ws.sync)), "sb200_forward_wait")
            R, R_arg = int(ws.count[0]), cap   # the backward locates the hit masks from the value the forward was given
            if R > cap:   # outgrown: only the first `cap` pairs were binned -- size up and run again
                ws.grow_pairs(R)
                R = R_arg = two_stage()
        last_counters.update(P=P, R=R, W=W, H=H)
        ctx.empty = False
        ctx.raster_settings = rs
        ctx.P = P
        ctx.num_rendered = R_arg
        ctx.sh_degree = int(rs.sh_degree)
        ctx.lease = lease   # geometry / image workspaces and the sorted id list live in the leased workspace
        ctx.save_for_backward(means3D, tm_c if ctx.use_pre else scales_c, rot_c, radii, color, others,
                              *((sh_c, clamped, campos) if ctx.use_sh else ()))
        ctx.mark_non_differentiable(radii)
        return color, radii, others

    @staticmethod
    def backward(ctx, grad_out_color, grad_radii, grad_depth):
        lib = _load()
        rs = ctx.raster_settings
        P = ctx.P
        H, W = int(rs.image_height), int(rs.image_width)
        if ctx.empty:
            dev = grad_out_color.device
            return tuple(torch.zeros(sh_, dtype=torch.float32, device=dev)
                         if sh_ is not None and ctx.needs_input_grad[i] else None
                         for i, sh_ in enumerate(ctx.in_shapes))
        means3D, scales_c, rot_c, radii, color, others = ctx.saved_tensors[:6]
        dev = means3D.device
        ws = ctx.lease.ws
        ws.fit_grad(P)
        g_color = _f32c(grad_out_color, "grad_out_color")
        g_others = _f32c(grad_depth, "grad_allmap")
        mask = rs.pixel_mask
        if not mask.is_cuda:
            raise RuntimeError("pixel_mask must be a CUDA tensor")
        mask = mask.contiguous()
        if mask.dtype != torch.bool:
            mask = mask != 0
        if mask.numel() != H * W:
            raise RuntimeError("pixel_mask must have H*W elements")
        bg = rs.bg.contiguous()
        view = rs.viewmatrix.contiguous()
        proj = rs.projmatrix.contiguous()
        e = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
        dmeans3D, dmeans2D, dcolors, dopac = e(P, 3), e(P, 3), e(P, 3), e(P, 1)
        dscales, drot, scores = e(P, 2), e(P, 4), e(P, 2)
        dtrans = e(P, 9) if ctx.use_pre else None   # dL/dtransMat only exists for precomputed transforms
        tail = (_ptr(view), _ptr(proj), _ptr(radii), _ptr(ws.geom), _vp(ws.point_list.data_ptr()), _ptr(ws.image),
                _ptr(color), _ptr(others), _ptr(g_color), _ptr(g_others), _ptr(mask), _ptr(ws.grad_ws), _ptr(dmeans3D),
                _ptr(dmeans2D), _ptr(dcolors), _ptr(dopac), _ptr(dscales), _ptr(drot), _ptr(dtrans), _ptr(scores),
                _stream(dev))
        if ctx.use_pre:  # scales_c holds the precomputed transforms
            _check(lib.sb200_backward_precomp(P, W, H, ctx.num_rendered, _ptr(bg), float(rs.tanfovx),
                                              float(rs.tanfovy), _ptr(scales_c), *tail), "sb200_backward_precomp")
        else:
            _check(lib.sb200_backward(P, W, H, ctx.num_rendered, _ptr(bg), float(rs.tanfovx), float(rs.tanfovy),
                                      _ptr(means3D), _ptr(scales_c), _ptr(rot_c), *tail), "sb200_backward")
        # forward-argument order of the reference (__init__.py:156-167).  The reference hands a tensor to
        # every slot; autograd rejects that for inputs that were None, so those slots get None here.
        dsh = None
        if ctx.use_sh:
            # dL/dcolors -> dL/dsh, plus the view-direction term of dL/dmeans3D (backward.cu:1038-1039)
            sh_c, clamped, campos = ctx.saved_tensors[6:]
            dsh = torch.empty_like(sh_c)
            _check(lib.sb200_sh_backward(P, ctx.sh_degree, sh_c.size(1), _ptr(means3D), _ptr(campos), _ptr(sh_c),
                                         _ptr(clamped), _ptr(radii), _ptr(dcolors), _ptr(dmeans3D), _ptr(dsh),
                                         _stream(dev)), "sb200_sh_backward")
        # slot 8 = cov3Ds_precomp: the reference returns dL/dtransMat there (__init__.py:165)
        grads = [dmeans3D, dmeans2D, dsh, None if ctx.use_sh else dcolors, dopac, dscales, drot, scores,
                 dtrans if ctx.use_pre else None, None]
        for i, shape in enumerate(ctx.in_shapes):
            if shape is None or not ctx.needs_input_grad[i]:
                grads[i] = None
            elif grads[i] is not None and tuple(grads[i].shape) != shape:
                grads[i] = grads[i].reshape(shape)
        return tuple(grads)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, scores, cov3Ds_precomp,
                        raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, scores,
                                     cov3Ds_precomp, raster_settings)


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float  # accepted and never used, like the reference (forward.cu:96)
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    pixel_mask: torch.Tensor


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """bool[P]: view-space z > 0.2 (rasterizer_impl.cu:106-118, auxiliary.h:212)."""
        lib = _load()
        with torch.no_grad():
            rs = self.raster_settings
            pos = _f32c(positions, "positions")
            P = pos.size(0)
            present = torch.zeros((P,), dtype=torch.bool, device=pos.device)
            _check(lib.sb200_mark_visible(P, _ptr(pos), _ptr(rs.viewmatrix.contiguous()),
                                          _ptr(rs.projmatrix.contiguous()), _ptr(present), _stream(pos.device)),
                   "sb200_mark_visible")
        return present

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, scores=None):
        raster_settings = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        dev = means3D.device
        empty = lambda: torch.empty(0, dtype=torch.float32, device=dev)
        if shs is None:
            shs = empty()
        if colors_precomp is None:
            colors_precomp = empty()
        if scales is None:
            scales = empty()
        if rotations is None:
            rotations = empty()
        if cov3D_precomp is None:
            cov3D_precomp = empty()
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations, scores,
                                   cov3D_precomp, raster_settings)


class SparseGaussianAdam(torch.optim.Adam):
    """Visibility-masked Adam without bias correction (reference __init__.py:241-268, adam.cu:9-38)."""

    def __init__(self, params, lr, eps):
        super().__init__(params=params, lr=lr, eps=eps)

    @torch.no_grad()
    def step(self, visibility, N):
        lib = _load()
        for group in self.param_groups:
            lr = group["lr"]
            eps = group["eps"]
            assert len(group["params"]) == 1, "more than one tensor in group"
            param = group["params"][0]
            if param.grad is None:
                continue
            state = self.state[param]
            if len(state) == 0:
                state['step'] = torch.tensor(0.0, dtype=torch.float32)
                state['exp_avg'] = torch.zeros_like(param, memory_format=torch.preserve_format)
                state['exp_avg_sq'] = torch.zeros_like(param, memory_format=torch.preserve_format)
            exp_avg = state["exp_avg"]
            exp_avg_sq = state["exp_avg_sq"]
            M = param.numel() // N
            if not (param.is_cuda and param.is_contiguous() and param.grad.is_contiguous()):
                raise RuntimeError("SparseGaussianAdam needs contiguous CUDA parameters")
            vis = visibility.contiguous()
            if vis.dtype != torch.bool:
                vis = vis != 0
            _check(lib.sb200_adam_update(_ptr(param), _ptr(param.grad), _ptr(exp_avg), _ptr(exp_avg_sq), _ptr(vis),
                                         float(lr), 0.9, 0.999, float(eps), int(N), int(M), _stream(param.device)),
                   "sb200_adam_update")


from . import mapping  # noqa: E402  (fused mapping-step blocks: N1/N2 of SURVEY.md section 8f)
from . import store  # noqa: E402  (surfel add / prune churn: N3)

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "SparseGaussianAdam", "rasterize_gaussians",
           "mapping", "store"]
