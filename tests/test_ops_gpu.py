"""GPU tests of the building blocks behind the C ABI: the onesweep radix sort, masked Adam, mark_visible,
degenerate inputs, and size-independent properties at BASELINE.json's full sizes."""
import ctypes
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vings-mono_b200"), os.path.join(ROOT, "tests")]
import synthetic  # noqa: E402
import util  # noqa: E402

pytestmark = pytest.mark.gpu
vp = ctypes.c_void_p


@pytest.fixture(scope="module")
def mine():
    return util.load_mine()


def _sort(lib, keys, vals, end_bit, mode=0):
    """keys int64 holding u32 values; vals None = the surfel-sort form (implicit indices, 0xFFFFFFFF keys dropped)"""
    n = keys.numel()
    ws = torch.zeros(max(lib.sb200_binning_bytes(n), 16), dtype=torch.uint8, device="cuda")
    off = (ctypes.c_size_t * 4)()
    lib.sb200_binning_layout(n, off)
    if n:
        ws[off[0]:off[0] + 4 * n].view(torch.int32).copy_(keys.to(torch.int32) if keys.dtype != torch.int32 else keys)
        if vals is not None:
            ws[off[2]:off[2] + 4 * n].view(torch.int32).copy_(vals)
    vo = torch.full((max(n, 1),), -1, dtype=torch.int32, device="cuda")
    nv = torch.zeros(1, dtype=torch.int32, device="cuda")
    rc = lib.sb200_sort_pairs(vp(ws.data_ptr()), n, end_bit, mode | (2 if vals is None else 0), vp(vo.data_ptr()),
                              vp(nv.data_ptr()), vp(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.sb200_last_error()
    torch.cuda.synchronize()
    return vo[:n], int(nv[0])


def _u32_as_i32(x):
    """int64 tensor of values < 2^32 -> the int32 tensor with the same bits"""
    return torch.where(x >= (1 << 31), x - (1 << 32), x).to(torch.int32)


@pytest.mark.parametrize("n", [0, 1, 31, 2047, 2048, 4095, 4096, 4097, 100_000, 3_000_001])
@pytest.mark.parametrize("end_bit,small", [(1, 0), (8, 1), (11, 0), (13, 0), (16, 1), (17, 0), (32, 1), (32, 0)])
def test_radix_sort_is_a_stable_lsd_sort(mine, n, end_bit, small):
    """Same contract as cub::DeviceRadixSort::SortPairs(bits [0,end_bit)) at rasterizer_impl.cu:410-418, on the 32-bit
    keys this pipeline sorts (tile ids: 11-16 bits in 2 passes; depth words: 32 bits in 4): stable, only the low
    end_bit bits participate."""
    lib = mine._load()
    gen = torch.Generator(device="cpu").manual_seed(n + end_bit)
    # few distinct values -> long runs of equal keys exercise stability; junk above end_bit must be ignored
    lo = torch.randint(0, 1 << min(end_bit, 9), (n,), generator=gen, dtype=torch.int64) << max(0, end_bit - 9)
    junk = torch.randint(0, 1 << 6, (n,), generator=gen, dtype=torch.int64) << end_bit if end_bit <= 26 else 0
    keys = lo | junk
    vals = torch.arange(n, dtype=torch.int32, device="cuda")
    vo, _ = _sort(lib, _u32_as_i32(keys).cuda(), vals, end_bit, mode=small)
    masked = keys & ((1 << end_bit) - 1)
    _, order = torch.sort(masked, stable=True)
    assert torch.equal(vo.long().cpu(), order)


@pytest.mark.parametrize("n,frac", [(1, 0.0), (1, 1.0), (2048, 0.5), (5000, 1.0), (300_000, 0.6), (1_200_000, 0.59),
                                    (1_200_000, 0.999)])
def test_surfel_sort_drops_invisible_and_keeps_index_order_on_ties(mine, n, frac):
    """The first pass of the surfel sort compacts: keys 0xFFFFFFFF (surfels touching no tile) vanish, the rest come
    out as indices in stable (depth word, index) order."""
    lib = mine._load()
    gen = torch.Generator(device="cpu").manual_seed(n)
    depth = (0.2 + 80 * torch.rand(n, generator=gen)).view(torch.int32).long()
    depth[torch.rand(n, generator=gen) < 0.3] = depth[0]          # many exact ties
    dead = torch.rand(n, generator=gen) < frac
    keys = torch.where(dead, torch.full_like(depth, 0xFFFFFFFF), depth)
    vo, kept = _sort(lib, _u32_as_i32(keys).cuda(), None, 32, mode=1)
    live = torch.nonzero(~dead).squeeze(1)
    assert kept == live.numel()
    _, o = torch.sort(depth[live], stable=True)
    assert torch.equal(vo[:kept].long().cpu(), live[o])


def test_adam_matches_reference_kernel(mine):
    ref = util.load_ref()
    gen = torch.Generator(device="cpu").manual_seed(3)
    N, M = 100_003, 3
    p0 = torch.randn(N, M, generator=gen).cuda()
    g = torch.randn(N, M, generator=gen).cuda()
    m0 = torch.randn(N, M, generator=gen).cuda() * 0.1
    v0 = torch.rand(N, M, generator=gen).cuda() * 0.01
    vis = (torch.rand(N, generator=gen) < 0.6).cuda()
    lib = mine._load()
    p1, m1, v1 = p0.clone(), m0.clone(), v0.clone()
    rc = lib.sb200_adam_update(vp(p1.data_ptr()), vp(g.data_ptr()), vp(m1.data_ptr()), vp(v1.data_ptr()),
                               vp(vis.data_ptr()), 1e-3, 0.9, 0.999, 1e-15, N, M, vp(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    # adam.cu:26-37 restated
    b1, b2 = torch.tensor(0.9, device="cuda"), torch.tensor(0.999, device="cuda")   # fp32 constants: 1 - 0.999f != 0.001
    m2 = torch.where(vis[:, None], b1 * m0 + (1.0 - b1) * g, m0)
    v2 = torch.where(vis[:, None], b2 * v0 + (1.0 - b2) * g * g, v0)
    p2 = torch.where(vis[:, None], p0 - 1e-3 * m2 / (v2.sqrt() + 1e-15), p0)
    # torch evaluates the same formulas with different FMA contraction -> a few ulp
    assert torch.allclose(m1, m2, rtol=1e-5, atol=1e-7) and torch.allclose(v1, v2, rtol=1e-5, atol=1e-9)
    assert torch.allclose(p1, p2, rtol=1e-5, atol=1e-6)
    assert torch.equal(p1[~vis], p0[~vis]) and torch.equal(m1[~vis], m0[~vis])   # invisible rows untouched
    if ref is not None:
        p3, m3, v3 = p0.clone(), m0.clone(), v0.clone()
        ref._C.adamUpdate(p3, g, m3, v3, vis, 1e-3, 0.9, 0.999, 1e-15, N, M)
        torch.cuda.synchronize()
        assert torch.equal(p1, p3) and torch.equal(m1, m3) and torch.equal(v1, v3)   # bit-exact


def test_sparse_adam_optimizer_class(mine):
    p = torch.nn.Parameter(torch.ones(10, 3, device="cuda"))
    opt = mine.SparseGaussianAdam([{"params": [p], "lr": 0.1, "name": "x"}], lr=0.0, eps=1e-15)
    p.grad = torch.ones_like(p)
    vis = torch.zeros(10, dtype=torch.bool, device="cuda")
    vis[:4] = True
    opt.step(vis, 10)
    torch.cuda.synchronize()
    assert torch.all(p[4:] == 1.0) and torch.all(p[:4] < 1.0)
    # no bias correction: m = 0.1, v = 0.001 -> step = lr * 0.1 / sqrt(0.001)
    assert torch.allclose(p[:4], torch.full((4, 3), 1.0 - 0.1 * 0.1 / (0.001 ** 0.5), device="cuda"), rtol=1e-5)


def test_mark_visible(mine):
    ref = util.load_ref()
    scene, views = synthetic.make_scene("C3", P=50_000)
    inp = util.cuda_inputs(synthetic.view_inputs(scene, views[0]))
    inp["means3D"][::3, 2] *= -1.0     # a third of the cloud behind the camera
    rast = mine.GaussianRasterizer(util.make_settings(mine, inp))
    got = rast.markVisible(inp["means3D"])
    vm = inp["viewmatrix"].reshape(16)
    p = inp["means3D"]
    pz = vm[2] * p[:, 0] + vm[6] * p[:, 1] + vm[10] * p[:, 2] + vm[14]
    assert got.dtype == torch.bool and got.shape == (50_000,)
    assert int((got != (pz > 0.2)).sum()) <= 2   # ulp-level differences of the dot product at the threshold
    assert 0 < int(got.sum()) < 50_000
    if ref is not None:
        want = ref.GaussianRasterizer(util.make_settings(ref, inp)).markVisible(inp["means3D"])
        assert torch.equal(got, want)


def test_empty_cloud_and_fully_culled_cloud(mine):
    scene, views = synthetic.make_scene("C1", P=1000)
    inp = util.cuda_inputs(synthetic.view_inputs(scene, views[0]))
    gc, ga = synthetic.upstream_grads(inp["W"], inp["H"], 0)
    # P == 0 (rasterize_points.cu:106): zero images, empty radii
    empty = dict(inp)
    for k in ("means3D", "scales", "rotations", "opacities", "colors"):
        empty[k] = inp[k][:0].contiguous()
    rast = mine.GaussianRasterizer(util.make_settings(mine, empty))
    color, radii, allmap = rast(means3D=empty["means3D"], means2D=None, opacities=empty["opacities"],
                                colors_precomp=empty["colors"], scales=empty["scales"], rotations=empty["rotations"],
                                scores=None)
    assert radii.numel() == 0 and float(color.abs().max()) == 0 and float(allmap.abs().max()) == 0
    # every surfel behind the camera: R == 0, background only, all gradients exactly zero
    behind = dict(inp)
    behind["means3D"] = (inp["means3D"] * torch.tensor([1.0, 1.0, -1.0], device="cuda")).contiguous()
    behind["bg"] = torch.tensor([0.25, 0.5, 0.75], device="cuda")
    out = util.run_impl(mine, behind, gc, ga)
    assert int(out["radii"].abs().sum()) == 0
    assert torch.allclose(out["color"][1], torch.full_like(out["color"][1], 0.5))
    assert float(out["allmap"].abs().max()) == 0
    for k in out:
        if k.startswith("d_"):
            assert float(out[k].abs().max()) == 0


def test_api_errors(mine):
    scene, views = synthetic.make_scene("C1", P=100)
    inp = util.cuda_inputs(synthetic.view_inputs(scene, views[0]))
    rast = mine.GaussianRasterizer(util.make_settings(mine, inp))
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        rast(means3D=inp["means3D"], means2D=None, opacities=inp["opacities"], scales=inp["scales"],
             rotations=inp["rotations"])
    with pytest.raises(Exception, match="scale/rotation pair"):
        rast(means3D=inp["means3D"], means2D=None, opacities=inp["opacities"], colors_precomp=inp["colors"])
    with pytest.raises(RuntimeError, match="num_points, 3"):
        rast(means3D=inp["means3D"].reshape(-1), means2D=None, opacities=inp["opacities"],
             colors_precomp=inp["colors"], scales=inp["scales"], rotations=inp["rotations"])
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        rast(means3D=inp["means3D"].cpu(), means2D=None, opacities=inp["opacities"], colors_precomp=inp["colors"],
             scales=inp["scales"], rotations=inp["rotations"])


@pytest.mark.parametrize("name", ["C2", "C3", "C4"])
def test_full_size_properties(mine, name):
    """BASELINE.json sizes (1.2 M / 500 k / 5 M surfels at 1080p): properties that need no oracle."""
    scene, views = synthetic.make_scene(name)
    inp = util.cuda_inputs(synthetic.view_inputs(scene, views[-1]))
    W, H = inp["W"], inp["H"]
    mi = util.mine_intermediates(mine, inp)
    R = mi["R"]
    assert R == int(mi["tiles_touched"].sum()) and R > 0
    # the surfel order is the stable (depth word, id) order of the visible surfels; pairs are emitted in that order,
    # row-major over each surfel's tile rect
    vis_ids = np.nonzero(mi["radii"] > 0)[0]
    dbits = mi["depths"].view(np.uint32)
    want = vis_ids[np.argsort(dbits[vis_ids], kind="stable")]
    assert np.array_equal(mi["surfel_order"], want.astype(np.uint32))
    assert np.array_equal(mi["emitted_ids"], np.repeat(want, mi["tiles_touched"][want]).astype(np.uint32))
    starts = np.cumsum(mi["tiles_touched"][want].astype(np.int64)) - mi["tiles_touched"][want]
    gx = (W + 15) // 16
    for k in np.linspace(0, len(want) - 1, 200).astype(np.int64):
        r = mi["rect"][want[k]].astype(np.int64)
        tiles = (np.arange(r[1], r[3])[:, None] * gx + np.arange(r[0], r[2])[None, :]).reshape(-1)
        assert np.array_equal(mi["emitted_tiles"][starts[k]:starts[k] + len(tiles)].astype(np.int64), tiles)
    # rect area == tiles_touched, radius > 0 <=> touches tiles
    rect = mi["rect"].astype(np.int64)
    area = (rect[:, 2] - rect[:, 0]) * (rect[:, 3] - rect[:, 1])
    vis = mi["radii"] > 0
    assert np.array_equal(area[vis], mi["tiles_touched"][vis].astype(np.int64)) and not mi["tiles_touched"][~vis].any()
    # sortedness on the significant bits, ranges partition the list by tile id
    tiles_n = ((W + 15) // 16) * ((H + 15) // 16)
    bits = 32 + util._higher_msb(tiles_n)
    masked = mi["keys"] & np.uint64((1 << bits) - 1)
    assert (np.diff(masked.astype(np.uint64)) >= 0).all() if R > 1 else True
    tid = (mi["keys"] >> np.uint64(32)).astype(np.int64)
    ranges = mi["ranges"].astype(np.int64)
    assert int((ranges[:, 1] - ranges[:, 0]).sum()) == R
    nz = np.nonzero(ranges[:, 1] > ranges[:, 0])[0]
    assert (tid[ranges[nz, 0]] == nz).all() and (tid[ranges[nz, 1] - 1] == nz).all()
    # the sorted id list is a permutation of the emitted pairs
    assert np.array_equal(np.bincount(mi["point_list"], minlength=len(vis)), mi["tiles_touched"].astype(np.int64))
    # blend invariants: alpha in [0,1], T = 1 - alpha, contributor counts within the tile list
    alpha = mi["allmap"][1].reshape(-1).numpy()
    assert alpha.min() >= 0 and alpha.max() <= 1.0
    assert np.allclose(1.0 - alpha, mi["final_T"][0], atol=1e-6)
    # forward is deterministic (no atomics): a second run is bit-identical
    mi2 = util.mine_intermediates(mine, inp)
    assert torch.equal(mi["color"], mi2["color"]) and torch.equal(mi["allmap"], mi2["allmap"])
    assert np.array_equal(mi["point_list"], mi2["point_list"])
    # linearity of the backward in the upstream gradient: grads(2g) == 2 grads(g) up to atomics order
    gc, ga = synthetic.upstream_grads(W, H, 11)
    o1 = util.run_impl(mine, inp, gc, ga)
    o2 = util.run_impl(mine, inp, 2 * gc, 2 * ga)
    for k in ("d_means3D", "d_colors", "d_opacities", "d_scales", "d_rotations"):
        assert util.rel_err(o2[k], 2 * o1[k]) < 1e-5, k
    assert util.rel_err(o2["d_scores"][:, 0], o1["d_scores"][:, 0]) < 1e-5   # scores[:,0] ignores the gradient
