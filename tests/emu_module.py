"""TEST INFRASTRUCTURE: a CPU stand-in for the rasterizer module (same three public names) built on the
oracle, so that host-side logic -- the mapping step and its multi-rank gradient exchange -- can be tested
without a GPU (gloo, world_size 2).  Never imported by the product."""
import os
import sys
from typing import NamedTuple

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import surfel_emulation as emu  # noqa: E402


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    pixel_mask: torch.Tensor


class _Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, colors, opacities, scales, rotations, scores, rs):
        W, H = rs.image_width, rs.image_height
        pre, bins, fwd = emu.rasterize_forward(means3D, colors, opacities, scales, rotations, rs.viewmatrix,
                                               rs.projmatrix, rs.bg, W, H)
        ctx.state = (pre, bins, fwd, rs)
        ctx.save_for_backward(means3D, colors, opacities, scales, rotations)
        radii = pre["radii"]
        ctx.mark_non_differentiable(radii)
        return fwd["color"], radii, fwd["allmap"]

    @staticmethod
    def backward(ctx, g_color, _g_radii, g_allmap):
        pre, bins, fwd, rs = ctx.state
        means3D, colors, opacities, scales, rotations = ctx.saved_tensors
        g = emu.rasterize_backward(pre, bins, fwd, means3D, colors, opacities, scales, rotations, rs.viewmatrix,
                                   rs.projmatrix, rs.bg, rs.image_width, rs.image_height, rs.tanfovx, rs.tanfovy,
                                   g_color, g_allmap, rs.pixel_mask)
        return (g["dL_dmeans3D"], g["dL_dmeans2D"], g["dL_dcolors"], g["dL_dopacity"], g["dL_dscales"],
                g["dL_drotations"], g["scores"], None)


class GaussianRasterizer(torch.nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, scores=None):
        return _Fn.apply(means3D, means2D, colors_precomp, opacities, scales, rotations, scores, self.raster_settings)


class SparseGaussianAdam(torch.optim.Adam):
    """adam.cu:26-37 in torch: visibility-masked, no bias correction."""

    def __init__(self, params, lr, eps):
        super().__init__(params=params, lr=lr, eps=eps)

    @torch.no_grad()
    def step(self, visibility, N):
        for group in self.param_groups:
            p = group["params"][0]
            if p.grad is None:
                continue
            st = self.state[p]
            if len(st) == 0:
                st["exp_avg"] = torch.zeros_like(p)
                st["exp_avg_sq"] = torch.zeros_like(p)
            vis = visibility.reshape(-1, *([1] * (p.dim() - 1)))
            m = 0.9 * st["exp_avg"] + 0.1 * p.grad
            v = 0.999 * st["exp_avg_sq"] + 0.001 * p.grad * p.grad
            p.copy_(torch.where(vis, p - group["lr"] * m / (v.sqrt() + group["eps"]), p))
            st["exp_avg"].copy_(torch.where(vis, m, st["exp_avg"]))
            st["exp_avg_sq"].copy_(torch.where(vis, v, st["exp_avg_sq"]))
