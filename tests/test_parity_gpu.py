"""GPU parity tests: this package (through its C ABI / public Python API) against
  (1) the reference's own rasterizer built for sm_100 (oracle/_ref) on the same seeded inputs,
  (2) the committed golden vectors (tests/golden, produced by that reference build),
  (3) size-independent properties at BASELINE.json's full sizes.
Bars: bit-exact for radii / keys / sorted order / tile ranges / contributor counts -- and, since the blend
arithmetic is pinned to the reference's operation order, for the rendered maps as well; gradients within
GRAD_TOL relative L1, justified below."""
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

# Gradients are sums of up to thousands of fp32 terms per surfel.  The reference adds them with atomics in
# nondeterministic order (its own run-to-run jitter measured on these configs: 1e-9 .. 2e-6 relative L1);
# this implementation reduces 32 pixels by shuffle, 8 warps through shared memory and tiles by atomics -- a
# different, also nondeterministic, association.  Measured agreement: 5e-7 (C1) .. 7e-5 (C2) relative L1.
GRAD_TOL = 5e-4
TAINT_TILES = 0.15   # share of the tiles whose stored max_contrib quirk Q17 may corrupt (measured 0 .. 7 %)
GOLDEN = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="module")
def mine():
    return util.load_mine()


@pytest.fixture(scope="module")
def ref():
    r = util.load_ref()
    if r is None:
        pytest.skip("oracle/_ref not built")
    return r


def _case(name, P=None, W=None, H=None, view=-1, bg=None):
    scene, views = synthetic.make_scene(name, P=P, W=W, H=H)
    inp = synthetic.view_inputs(scene, views[view])
    if bg is not None:
        inp["bg"] = torch.tensor(bg, dtype=torch.float32)
    gc, ga = synthetic.upstream_grads(inp["W"], inp["H"], synthetic.CONFIGS[name]["seed"])
    return util.cuda_inputs(inp), gc, ga


CASES = [
    ("C1", dict()),                                        # BASELINE config 0 in full: 10k surfels, 320x240
    ("C1", dict(P=3000, W=200, H=150)),                    # ragged image (not a multiple of the 16-px tile)
    ("C3", dict(P=60000, bg=(0.3, 0.1, 0.7))),             # indoor, non-zero background (quirk Q6)
    ("C2", dict(P=150000)),                                # KITTI-shaped wide image
    ("C1", dict(P=4000, W=4112, H=2600)),                  # 41 891 tiles: 16 tile bits, tile counts by global REDs
]


@pytest.mark.parametrize("name,kw", CASES)
def test_integer_stages_bit_exact(mine, ref, name, kw):
    inp, gc, ga = _case(name, **kw)
    mi = util.mine_intermediates(mine, inp)
    ri = util.ref_intermediates(ref, inp)
    assert mi["R"] == ri["R"]
    assert np.array_equal(mi["radii"], ri["radii"])
    assert np.array_equal(mi["tiles_touched"], ri["tiles_touched"])
    vis = ri["radii"] > 0
    for k in ("depths", "means2D", "transMat", "normal_opacity"):
        assert np.array_equal(mi[k][vis].view(np.uint32), ri[k][vis].view(np.uint32)), k
    assert np.array_equal(mi["keys"], ri["keys"])
    assert np.array_equal(mi["point_list"], ri["point_list"])
    assert np.array_equal(mi["ranges"], ri["ranges"])
    assert np.array_equal(mi["n_contrib"], ri["n_contrib"])     # last and median contributor per pixel


@pytest.mark.parametrize("name,kw", CASES)
def test_rendered_maps_bit_exact(mine, ref, name, kw):
    inp, gc, ga = _case(name, **kw)
    mi = util.mine_intermediates(mine, inp)
    ri = util.ref_intermediates(ref, inp)
    # tolerance 0: colour, depth, alpha, normal, median depth, distortion, and the saved T / M1 / M2
    assert torch.equal(mi["color"], ri["color"])
    assert torch.equal(mi["allmap"], ri["allmap"])
    assert np.array_equal(mi["final_T"].view(np.uint32), ri["final_T"].view(np.uint32))


@pytest.mark.parametrize("name,kw", CASES)
def test_gradients_match_reference(mine, ref, name, kw):
    inp, gc, ga = _case(name, **kw)
    gen = torch.Generator(device="cpu").manual_seed(5)
    mask = (torch.rand(inp["W"] * inp["H"], generator=gen) < 0.8).cuda()   # exercises pixel_mask (Q14)
    om = util.run_impl(mine, inp, gc, ga, mask)
    orf = util.run_impl(ref, inp, gc, ga, mask, ref_state=True)
    # reference UB, quirk Q17: read off the buffers of the same forward whose backward produced orf's gradients
    ids, nbad = util.tainted_surfels(orf["state"], inp["W"], inp["H"])
    keep = torch.ones(om["d_scores"].shape[0], dtype=torch.bool)
    keep[torch.from_numpy(ids)] = False
    # the exclusion set (reference UB, Q17) is bounded; test_parity_full_gpu.py proves it is exactly right
    tiles_n = ((inp["W"] + 15) // 16) * ((inp["H"] + 15) // 16)
    assert nbad <= TAINT_TILES * tiles_n, (nbad, tiles_n, len(ids))
    assert torch.equal(om["radii"], orf["radii"])
    for k in ("d_means3D", "d_means2D", "d_colors", "d_opacities", "d_scales", "d_rotations", "d_scores"):
        assert om[k].shape == orf[k].shape
        assert torch.isfinite(om[k]).all()
        rel = util.rel_err(om[k][keep], orf[k][keep])
        assert rel < GRAD_TOL, (k, rel)
    # culled surfels get exact zeros
    culled = orf["radii"] == 0
    if culled.any():
        assert float(om["d_means3D"][culled].abs().max()) == 0.0 and float(om["d_scores"][culled].abs().max()) == 0.0


def test_scores_with_zero_upstream_gradient(mine, ref):
    """Q13: scripts/loop/loop_model.py:180-186 backpropagates a zero loss only to read scores[:,0]."""
    inp, gc, ga = _case("C1", P=3000)
    om = util.run_impl(mine, inp, torch.zeros_like(gc), torch.zeros_like(ga))
    orf = util.run_impl(ref, inp, torch.zeros_like(gc), torch.zeros_like(ga), ref_state=True)
    ids, _ = util.tainted_surfels(orf["state"], inp["W"], inp["H"])
    keep = torch.ones(om["d_scores"].shape[0], dtype=torch.bool)
    keep[torch.from_numpy(ids)] = False
    assert float(om["d_scores"][:, 0].sum()) > 0
    assert util.rel_err(om["d_scores"][keep, 0], orf["d_scores"][keep, 0]) < GRAD_TOL
    assert float(om["d_colors"].abs().max()) == 0.0 and float(om["d_scores"][:, 1].abs().max()) == 0.0


@pytest.mark.parametrize("name", ["outdoor_small", "indoor_ragged"])
def test_against_committed_golden_vectors(mine, name):
    """No reference build needed: the fixtures hold the reference's outputs."""
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    inp = {k[3:]: torch.from_numpy(g[k]).cuda() for k in g.files if k.startswith("in_")}
    inp.update(W=int(g["W"]), H=int(g["H"]), tanfovx=float(g["tanfovx"]), tanfovy=float(g["tanfovy"]))
    mi = util.mine_intermediates(mine, inp)
    assert mi["R"] == int(g["R"])
    assert np.array_equal(mi["radii"], g["radii"])
    assert np.array_equal(mi["keys"], g["keys"])
    assert np.array_equal(mi["point_list"], g["point_list"])
    assert np.array_equal(mi["ranges"], g["ranges"])
    assert np.array_equal(mi["n_contrib"], g["n_contrib"])
    assert np.array_equal(mi["color"].numpy(), g["color"])
    assert np.array_equal(mi["allmap"].numpy(), g["allmap"])
    out = util.run_impl(mine, inp, torch.from_numpy(g["g_color"]), torch.from_numpy(g["g_allmap"]),
                        torch.from_numpy(g["mask"]).cuda())
    keep = np.ones(len(g["radii"]), bool)
    keep[g["tainted"]] = False
    for k in util.GRAD_NAMES:
        a, b = out["d_" + k].numpy()[keep], g["d_" + k][keep]
        rel = np.abs(a - b).sum() / max(np.abs(b).sum(), 1e-30)
        assert rel < GRAD_TOL, (k, rel)
