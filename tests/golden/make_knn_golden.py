"""Golden vector for sb200_knn_mean_dist2 from the REFERENCE'S OWN distCUDA2 (scripts/gaussian/gaussian_utils.py:12-16,
scipy KD-tree), run on CPU in the build container:  python tests/golden/make_knn_golden.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, "/root/reference/scripts")
import types  # noqa: E402
for missing in ("pytz",):  # imported by gaussian/general_utils.py for log file names only; not installed here
    sys.modules.setdefault(missing, types.ModuleType(missing))
from gaussian.gaussian_utils import distCUDA2  # noqa: E402


def main():
    g = torch.Generator(device="cpu").manual_seed(21)
    # a back-projected depth map patch (the mapper's use) + clustered points + exact duplicates
    u, v = torch.meshgrid(torch.arange(40.0), torch.arange(30.0), indexing="ij")
    z = 3.0 + 0.02 * u + 0.5 * torch.sin(v / 4.0) + 0.01 * torch.rand(40, 30, generator=g)
    grid = torch.stack(((u - 20) / 30 * z, (v - 15) / 30 * z, z), -1).reshape(-1, 3)
    cluster = 0.05 * torch.randn(500, 3, generator=g) + torch.tensor([1.0, -2.0, 4.0])
    pts = torch.cat((grid, cluster, cluster[:25], grid[:10])).float()
    out = distCUDA2(pts)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "knn_dist2.npz"), points=pts.numpy(), dist2=out.numpy())
    print(pts.shape, float(out.min()), float(out.max()), int((out == 0).sum()))


if __name__ == "__main__":
    main()
