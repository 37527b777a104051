"""Seeded synthetic surfel clouds and cameras of the shapes BASELINE.json names.

Everything is generated on the CPU from a ``torch.Generator`` so that every implementation (this
package, the reference build, the CPU emulation) sees identical bits.  The cloud is grown the way
VINGS-Mono's mapper grows it (scripts/gaussian/gaussian_model.py:136-156): <= 40 000 back-projected
points per *source* keyframe on a forward-motion trajectory, isotropic scale from the sample spacing;
the training window is the last K source keyframes.  Cameras follow scripts/gaussian/cameras.py:40-67
(get_camera: FoV from focal, 3DGS projection with znear .01 / zfar 100, row-major tensors holding the
transposed matrices).
"""
import math
from typing import NamedTuple

import torch

CONFIGS = {
    # id: P surfels, window K, W, H, outdoor?, seed, per-keyframe points, supervision
    "C1": dict(P=10_000, K=1, W=320, H=240, outdoor=True, seed=0, per_kf=10_000, sup="rgb"),
    "C2": dict(P=1_200_000, K=8, W=1241, H=376, outdoor=True, seed=1, per_kf=40_000, sup="rgb+depth"),
    "C3": dict(P=500_000, K=8, W=640, H=480, outdoor=False, seed=2, per_kf=40_000, sup="rgb+depth+normal+dist"),
    "C4": dict(P=5_000_000, K=16, W=1920, H=1080, outdoor=True, seed=3, per_kf=40_000, sup="rgb+depth"),
    "C5": dict(P=3_000_000, K=5, W=1920, H=1280, outdoor=True, seed=4, per_kf=40_000, sup="rgb+depth", rig=5),
}


class View(NamedTuple):
    viewmatrix: torch.Tensor   # [4,4] = w2c.T
    projmatrix: torch.Tensor   # [4,4] = w2c.T @ P.T
    campos: torch.Tensor       # [3]
    tanfovx: float
    tanfovy: float
    W: int
    H: int
    c2w: torch.Tensor


def projection_matrix(fovx, fovy, znear=0.01, zfar=100.0):
    """scripts/gaussian/cameras.py:10-28"""
    t, r = math.tan(fovy / 2) * znear, math.tan(fovx / 2) * znear
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (2 * r)
    P[1, 1] = 2.0 * znear / (2 * t)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def make_view(c2w, fx, fy, W, H):
    """scripts/gaussian/cameras.py:40-67 (get_camera)"""
    w2c = torch.linalg.inv(c2w)
    fovy = 2 * math.atan(H / (2 * fy))
    fovx = 2 * math.atan(W / (2 * fx))
    wvt = w2c.t().contiguous()
    proj = projection_matrix(fovx, fovy).t().contiguous()
    full = (wvt.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0).contiguous()
    campos = torch.linalg.inv(wvt)[3, :3].contiguous()
    return View(wvt.float(), full.float(), campos.float(), math.tan(fovx * 0.5), math.tan(fovy * 0.5), W, H, c2w)


def _yaw(deg):
    a = math.radians(deg)
    R = torch.eye(4)
    R[0, 0], R[0, 2], R[2, 0], R[2, 2] = math.cos(a), math.sin(a), -math.sin(a), math.cos(a)
    return R


def _depth_field(dirs, outdoor, gen):
    """Smooth synthetic depth along unit-z rays: ground plane + two side walls + low-frequency noise."""
    if outdoor:
        cam_h, wall, zmin, zmax = 1.5, 6.0, 2.0, 40.0
    else:
        cam_h, wall, zmin, zmax = 1.2, 2.5, 0.5, 6.0
    dx, dy = dirs[:, 0], dirs[:, 1]
    big = torch.full_like(dx, 1e9)
    ground = torch.where(dy > 1e-6, cam_h / dy.clamp(min=1e-6), big)
    walls = torch.where(dx.abs() > 1e-6, wall / dx.abs().clamp(min=1e-6), big)
    z = torch.minimum(torch.minimum(ground, walls), torch.full_like(dx, zmax))
    ph = torch.rand(3, generator=gen) * 6.28
    noise = 0.05 * (torch.sin(3.0 * dx + ph[0]) + torch.sin(4.0 * dy + ph[1]) + torch.sin(2.0 * (dx + dy) + ph[2]))
    return (z * (1.0 + noise)).clamp(zmin, zmax)


def make_scene(name, P=None, K=None, W=None, H=None):
    """Returns dict(means3D, scales, rotations, opacities, colors) as CPU fp32 tensors (already the
    *activated* values the mapper hands to the rasterizer) and the list of window views."""
    cfg = dict(CONFIGS[name])
    if P is not None:
        cfg["P"] = P
    if K is not None:
        cfg["K"] = K
    if W is not None:
        cfg["W"] = W
    if H is not None:
        cfg["H"] = H
    P, W, H = cfg["P"], cfg["W"], cfg["H"]
    gen = torch.Generator(device="cpu").manual_seed(cfg["seed"])
    fx = fy = 0.8 * W
    cx, cy = W / 2.0, H / 2.0
    per_kf = min(cfg["per_kf"], P)
    S = max(1, math.ceil(P / per_kf))
    spacing = 1.0 if cfg["outdoor"] else 0.3
    rig = cfg.get("rig", 1)
    rig_yaws = [0.0, 45.0, -45.0, 90.0, -90.0][:rig]

    poses = []
    for s in range(S):
        c2w = torch.eye(4)
        c2w[2, 3] = s * spacing
        jitter = float(torch.randn(1, generator=gen)) * 2.0
        poses.append(c2w @ _yaw(jitter))

    means, scales = [], []
    left = P
    for s in range(S):
        n = min(per_kf, left)
        left -= n
        u = torch.rand(n, generator=gen) * W
        v = torch.rand(n, generator=gen) * H
        dirs = torch.stack([(u - cx) / fx, (v - cy) / fy, torch.ones(n)], -1)
        # for the rig config the source keyframes cycle through the rig's cameras
        c2w = poses[s] @ _yaw(rig_yaws[s % rig])
        z = _depth_field(dirs, cfg["outdoor"], gen)
        pc = dirs * z[:, None]
        pw = pc @ c2w[:3, :3].t() + c2w[:3, 3]
        means.append(pw)
        sc = (z / fx) * math.sqrt(W * H / per_kf) * torch.exp(0.3 * torch.randn(n, generator=gen))
        scales.append(torch.stack([sc, sc], -1))
    means3D = torch.cat(means).float().contiguous()
    scales = torch.cat(scales).float().contiguous()
    q = torch.randn(P, 4, generator=gen)
    rotations = (q / q.norm(dim=-1, keepdim=True)).float().contiguous()
    opacities = torch.sigmoid(1.5 * torch.randn(P, 1, generator=gen)).float().contiguous()
    colors = torch.rand(P, 3, generator=gen).float().contiguous()

    views = []
    if rig > 1:
        for yaw in rig_yaws[:cfg["K"]]:
            views.append(make_view(poses[S - 1] @ _yaw(yaw), fx, fy, W, H))
    else:
        for s in range(max(0, S - cfg["K"]), S):
            views.append(make_view(poses[s], fx, fy, W, H))
    while len(views) < cfg["K"]:  # tiny clouds: repeat the last pose with a small offset
        c2w = poses[-1].clone()
        c2w[0, 3] += 0.05 * len(views)
        views.append(make_view(c2w, fx, fy, W, H))
    scene = dict(means3D=means3D, scales=scales, rotations=rotations, opacities=opacities, colors=colors,
                 bg=torch.zeros(3), W=W, H=H, name=name, cfg=cfg)
    return scene, views


def upstream_grads(W, H, seed):
    """dL/dcolor ~ N(0,1)/HW, dL/dallmap ~ N(0,1)/HW on channels 0-4 and 6 (channel 5 has no effect, Q3)."""
    gen = torch.Generator(device="cpu").manual_seed(1000 + seed)
    g_color = torch.randn(3, H, W, generator=gen) / (W * H)
    g_allmap = torch.randn(7, H, W, generator=gen) / (W * H)
    g_allmap[5] = 0
    return g_color.float().contiguous(), g_allmap.float().contiguous()


def view_inputs(scene, view):
    """Flat dict of CPU tensors for one (scene, view) -- the shape the oracle's forward_backward() takes."""
    d = {k: scene[k] for k in ("means3D", "scales", "rotations", "opacities", "colors", "bg", "W", "H")}
    d.update(viewmatrix=view.viewmatrix, projmatrix=view.projmatrix, campos=view.campos, tanfovx=view.tanfovx,
             tanfovy=view.tanfovy)
    return d


def make_gt(scene, view, index):
    """Ground truth of one window view for the mapping step: box-blurred uniform rgb,
    the synthetic depth field seen from that pose, unit depth covariance.  The top rows are zeroed so that
    the loss's sky mask (rgb == 0, loss_utils.py:112) is not empty -- its mean would be NaN otherwise."""
    cfg = scene["cfg"]
    W, H = scene["W"], scene["H"]
    gen = torch.Generator(device="cpu").manual_seed(7000 + 31 * cfg["seed"] + index)
    rgb = torch.rand(1, 3, H, W, generator=gen)
    rgb = torch.nn.functional.avg_pool2d(rgb, 5, stride=1, padding=2, count_include_pad=False)[0]
    sky_rows = max(2, int((0.15 if cfg["outdoor"] else 0.01) * H))
    rgb[:, :sky_rows, :] = 0.0
    fx = fy = 0.8 * W
    v, u = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    dirs = torch.stack([(u.reshape(-1) - W / 2.0) / fx, (v.reshape(-1) - H / 2.0) / fy, torch.ones(H * W)], -1)
    depth = _depth_field(dirs, cfg["outdoor"], gen).reshape(1, H, W)
    cov = torch.ones(1, H, W)
    return dict(rgb=rgb.float().contiguous(), depth=depth.float().contiguous(), depth_cov=cov)
