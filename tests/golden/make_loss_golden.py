"""Golden vectors for the fused mapping loss (sb200_map_loss), produced by the REFERENCE'S OWN Python:

    python tests/golden/make_loss_golden.py          # in the build container (needs /root/reference)

imports scripts/gaussian/loss_utils.get_loss, scripts/gaussian/normal_utils.depth_propagate_normal and
scripts/gaussian/tf.TFer from /root/reference (CPU, fp32), applies render_raw's post-processing exactly as
scripts/gaussian/gaussian_base.py:205-221 spells it, and stores inputs, the loss and autograd's gradients with
respect to the rasterizer outputs (color, allmap) in tests/golden/loss_*.npz.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF_SCRIPTS = "/root/reference/scripts"

CASES = {
    # name: (W, H, loss weights, seed)
    "loss_rgb_depth": (56, 40, dict(rgb_loss=1.0, depth_loss=1.0, alpha_loss=0.5, normal_loss=0.0, dist_loss=0.0), 11),
    "loss_normal_dist": (75, 37, dict(rgb_loss=1.0, depth_loss=1.0, alpha_loss=0.5, normal_loss=0.05, dist_loss=100.0), 12),
}


def case_inputs(name):
    """Seeded synthetic rasterizer outputs and ground truth with every special case of the loss: sky rows
    (gt rgb == 0), pixels without depth (gt depth == 0), pixels no surfel covers (allmap[1] == 0)."""
    W, H, weights, seed = CASES[name]
    g = torch.Generator(device="cpu").manual_seed(seed)
    r = lambda *s: torch.rand(*s, generator=g)
    yy, xx = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    alpha = (0.3 + 0.7 * r(H, W)).clamp(0, 0.999)
    depth = 2.0 + 0.05 * xx + 0.03 * yy + 0.2 * torch.sin(xx / 5.0) * torch.cos(yy / 4.0) + 0.05 * r(H, W)
    uncovered = r(H, W) < 0.03
    alpha[uncovered] = 0.0
    nrm = torch.nn.functional.normalize(torch.randn(3, H, W, generator=g), dim=0)
    allmap = torch.zeros(7, H, W)
    allmap[0] = depth * alpha
    allmap[1] = alpha
    allmap[2:5] = nrm * alpha
    allmap[5] = depth
    allmap[6] = 0.01 * r(H, W) * alpha
    color = r(3, H, W) * alpha
    gt_rgb = (0.6 * color + 0.4 * r(3, H, W)).clamp(0, 1)
    gt_rgb[:, :4, :] = 0.0                                # sky
    gt_depth = (depth + 0.1 * torch.randn(H, W, generator=g)).clamp(min=0.1).unsqueeze(0)
    gt_depth[0][r(H, W) < 0.1] = 0.0                      # no depth measurement
    gt_cov = (0.5 + r(1, H, W))
    a = 0.3
    Rz = torch.tensor([[np.cos(a), -np.sin(a), 0.0], [np.sin(a), np.cos(a), 0.0], [0.0, 0.0, 1.0]], dtype=torch.float32)
    Rx = torch.tensor([[1.0, 0.0, 0.0], [0.0, np.cos(0.2), -np.sin(0.2)], [0.0, np.sin(0.2), np.cos(0.2)]], dtype=torch.float32)
    w2c = torch.eye(4)
    w2c[:3, :3] = Rz @ Rx
    w2c[:3, 3] = torch.tensor([0.1, -0.2, 0.3])
    intr = dict(fu=0.8 * W, fv=0.8 * W, cu=H / 2.0 - 0.5, cv=W / 2.0 + 0.25, H=H, W=W)  # u = row, v = column
    return dict(color=color, allmap=allmap, gt_rgb=gt_rgb, gt_depth=gt_depth, gt_cov=gt_cov, w2c=w2c, intr=intr,
                weights=weights)


def reference_loss(c):
    sys.path.insert(0, REF_SCRIPTS)
    from gaussian.loss_utils import get_loss
    from gaussian.normal_utils import depth_propagate_normal
    from gaussian.tf import TFer
    cfg = {"intrinsic": c["intr"], "device": {"mapper": "cpu"}, "use_sky": False,
           "training_args": {"loss_weights": c["weights"]}}
    tfer = TFer(cfg)
    color = c["color"].clone().requires_grad_(True)
    allmap = c["allmap"].clone().requires_grad_(True)
    w2c = c["w2c"]
    # ---- scripts/gaussian/gaussian_base.py:205-221, verbatim semantics ----
    render_alpha = allmap[1:2]
    render_depth_expected = allmap[0:1]
    render_depth_expected = (render_depth_expected / render_alpha)
    render_depth_expected = torch.nan_to_num(render_depth_expected, 0, 0)
    rets = {}
    rets["accum"] = render_alpha
    rets["rgb"] = color
    rets["depth"] = render_depth_expected
    rets["normal"] = (allmap[2:5].permute(1, 2, 0) @ (w2c[:3, :3])).permute(2, 0, 1)
    rets["dist"] = allmap[6:7]
    rets["surf_normal"] = depth_propagate_normal(rets["depth"].squeeze(0), tfer).permute(2, 0, 1)
    rets["surf_normal"] = (rets["surf_normal"].permute(1, 2, 0) @ (w2c[:3, :3])).permute(2, 0, 1)
    gt = {"rgb": c["gt_rgb"], "depth": c["gt_depth"], "depth_cov": c["gt_cov"]}
    loss = get_loss(cfg, rets, gt)
    loss.backward()
    return loss.detach(), color.grad, allmap.grad


def main():
    for name in CASES:
        c = case_inputs(name)
        loss, g_color, g_allmap = reference_loss(c)
        out = os.path.join(ROOT, "tests", "golden", name + ".npz")
        np.savez_compressed(out, color=c["color"].numpy(), allmap=c["allmap"].numpy(), gt_rgb=c["gt_rgb"].numpy(),
                            gt_depth=c["gt_depth"].numpy(), gt_cov=c["gt_cov"].numpy(), w2c=c["w2c"].numpy(),
                            intr=np.array([c["intr"][k] for k in ("fu", "fv", "cu", "cv")], dtype=np.float64),
                            weights=np.array([c["weights"][k] for k in ("rgb_loss", "normal_loss", "alpha_loss",
                                                                          "depth_loss", "dist_loss")]),
                            loss=loss.numpy(), g_color=g_color.numpy(), g_allmap=g_allmap.numpy())
        nan = int(torch.isnan(g_allmap).sum())
        print(name, "loss", float(loss), "NaN grads (uncovered pixels)", nan, "->", out)


if __name__ == "__main__":
    main()
