"""Full-size parity: this package against the reference's own sm_100 build (oracle/_ref) at BASELINE.json's sizes.

    C2  1.2 M surfels, 1241x376      (configs[1], one view of the 8-keyframe window)
    C3  500 k surfels, 640x480       (configs[2])
    C4  5 M surfels, 1920x1080       (configs[3], one view of the 16-keyframe window)
    C5  3 M surfels, 1920x1280       (configs[4], one camera of the 5-camera rig)

Bars (the same as the reduced cases of test_parity_gpu.py): radii, tiles_touched, depth bits, per-surfel transforms,
sorted surfel ids (`point_list`), tile ranges and last / median contributor bit-exact; colour, allmap, final T / M1 /
M2 tolerance 0; gradients <= GRAD_TOL relative L1 on the surfels the reference computes deterministically, with the
excluded set (reference quirk Q17) bounded and PROVEN to be the whole story: with the repo's backward walk capped at the
reference's stored `max_contrib` (sb200_debug_walk_cap) the excluded surfels agree as well.

The 64-bit keys are not an independent item here: this pipeline never materialises them (split sort, DESIGN.md
section 4); util.mine_intermediates rebuilds them from ranges + depth words + point_list, all compared below.

The reference needs ~13 KiB of SampleState per 32-surfel bucket (rasterizer_impl.cu:271-286): 14 GB for the C4 view,
which fits a 180 GB B200.
"""
import gc
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

GRAD_TOL = 5e-4          # relative L1, atomic-order jitter (justified in test_parity_gpu.py)
# Q17: share of the TILES whose stored max_contrib is not the tile maximum (the reference's backward drops buckets
# there).  Measured 4.4 % (C2) .. 7 % (C3); the surfels sitting in such buckets are 18 % of C2's visible surfels
# because the affected tiles are the deep ones.  The bound only guards against the exclusion swallowing the test;
# that the set is exactly right is what test_q17_explains_the_whole_exclusion_set proves.
TAINT_TILES = 0.15
FULL = ["C2", "C3", "C4", "C5"]
GRADS = ("d_means3D", "d_means2D", "d_colors", "d_opacities", "d_scales", "d_rotations", "d_scores")


@pytest.fixture(scope="module")
def mine():
    return util.load_mine()


@pytest.fixture(scope="module")
def ref():
    r = util.load_ref()
    if r is None:
        pytest.skip("oracle/_ref not built")
    return r


_cache = {}


def _run(mine, ref, name):
    """Everything one config needs, computed once (the scene alone is 5 M surfels at C4)."""
    if name in _cache:
        return _cache[name]
    _cache.clear()          # one config resident at a time
    gc.collect()
    torch.cuda.empty_cache()
    scene, views = synthetic.make_scene(name)
    inp = util.cuda_inputs(synthetic.view_inputs(scene, views[-1]))
    W, H = inp["W"], inp["H"]
    gc_, ga = synthetic.upstream_grads(W, H, synthetic.CONFIGS[name]["seed"])
    gen = torch.Generator(device="cpu").manual_seed(5)
    mask = (torch.rand(W * H, generator=gen) < 0.8).cuda()   # exercises pixel_mask (Q14)
    mi = util.mine_intermediates(mine, inp)
    ri = util.ref_intermediates(ref, inp)
    torch.cuda.empty_cache()
    # the reference's max_contrib depends on stale shared memory (Q17): read it from the buffers of the very forward
    # whose backward produces the gradients compared below, not from the separate forward above
    orf = util.run_impl(ref, inp, gc_, ga, mask, ref_state=True)
    rs = orf.pop("state")
    assert np.array_equal(rs["point_list"], ri["point_list"]) and np.array_equal(rs["n_contrib"], ri["n_contrib"])
    torch.cuda.empty_cache()
    om = util.run_impl(mine, inp, gc_, ga, mask)
    ids, nbad = util.tainted_surfels(rs, W, H)
    # the same backward with every tile's walk stopped where the reference's stored max_contrib stops its own
    cap = torch.from_numpy(rs["max_contrib"].astype(np.int64)).to(torch.int32).cuda()
    mine.debug_walk_cap(cap)
    try:
        oc = util.run_impl(mine, inp, gc_, ga, mask)
    finally:
        mine.debug_walk_cap(None)
    _cache[name] = dict(inp=inp, mi=mi, ri=ri, om=om, orf=orf, oc=oc, tainted=ids, bad_tiles=nbad)
    return _cache[name]


@pytest.mark.parametrize("name", FULL)
def test_full_size_integer_stages_bit_exact(mine, ref, name):
    c = _run(mine, ref, name)
    mi, ri = c["mi"], c["ri"]
    assert mi["R"] == ri["R"] > 0
    assert np.array_equal(mi["radii"], ri["radii"])
    assert np.array_equal(mi["tiles_touched"], ri["tiles_touched"])
    vis = ri["radii"] > 0
    for k in ("depths", "means2D", "transMat", "normal_opacity"):
        assert np.array_equal(mi[k][vis].view(np.uint32), ri[k][vis].view(np.uint32)), k
    assert np.array_equal(mi["point_list"], ri["point_list"])
    assert np.array_equal(mi["ranges"], ri["ranges"])
    assert np.array_equal(mi["keys"], ri["keys"])   # derived (see the module docstring); kept as a cross-check
    assert np.array_equal(mi["n_contrib"], ri["n_contrib"])
    print("%s: P=%d V=%d R=%d" % (name, len(vis), int(vis.sum()), mi["R"]))


@pytest.mark.parametrize("name", FULL)
def test_full_size_maps_bit_exact(mine, ref, name):
    c = _run(mine, ref, name)
    mi, ri = c["mi"], c["ri"]
    assert torch.equal(mi["color"], ri["color"])
    assert torch.equal(mi["allmap"], ri["allmap"])
    assert np.array_equal(mi["final_T"].view(np.uint32), ri["final_T"].view(np.uint32))


@pytest.mark.parametrize("name", FULL)
def test_full_size_gradients(mine, ref, name):
    c = _run(mine, ref, name)
    om, orf, ids = c["om"], c["orf"], c["tainted"]
    P = om["d_scores"].shape[0]
    V = int((orf["radii"] > 0).sum())
    frac = len(ids) / max(V, 1)
    print("%s: Q17 excludes %d of %d visible surfels (%.3f %%), %d tiles with a wrong max_contrib"
          % (name, len(ids), V, 100 * frac, c["bad_tiles"]))
    tiles_n = ((c["inp"]["W"] + 15) // 16) * ((c["inp"]["H"] + 15) // 16)
    assert c["bad_tiles"] <= TAINT_TILES * tiles_n, (c["bad_tiles"], tiles_n)
    assert frac < 0.5, frac
    keep = torch.ones(P, dtype=torch.bool)
    keep[torch.from_numpy(ids)] = False
    assert torch.equal(om["radii"], orf["radii"])
    for k in GRADS:
        assert torch.isfinite(om[k]).all(), k
        rel = util.rel_err(om[k][keep], orf[k][keep])
        assert rel < GRAD_TOL, (k, rel)
    culled = orf["radii"] == 0
    assert float(om["d_means3D"][culled].abs().max()) == 0.0 and float(om["d_scores"][culled].abs().max()) == 0.0


# The excluded surfels are, by construction, the deepest entries of the deepest tile lists: their blend weights are
# ~1e-3 of a front surfel's, while the prefix-minus-final accumulators they are differenced against (backward.cu:
# 629-648) carry the rounding of the whole list in front of them -- and the two implementations round that prefix
# differently (the reference restarts from its per-bucket checkpoints, this backward accumulates from the tile start).
# Measured on this subset with the walk capped: 1e-6 (scores) .. 5.3e-4 (means2D) relative L1; uncapped: 1e-3 .. 7e-2.
SUBSET_TOL = 2e-3


@pytest.mark.parametrize("name", FULL)
def test_q17_explains_the_whole_exclusion_set(mine, ref, name):
    """With the walk capped at the reference's stored max_contrib (sb200_debug_walk_cap) the repo's backward reproduces
    the reference's gradients on EVERY surfel -- the whole cloud to GRAD_TOL with nothing excluded, and the excluded
    set taken alone to SUBSET_TOL, several times closer than without the cap: the reference's dropped buckets account
    for all of the difference the exclusion hides."""
    c = _run(mine, ref, name)
    oc, om, orf, ids = c["oc"], c["om"], c["orf"], torch.from_numpy(c["tainted"])
    for k in GRADS:
        rel = util.rel_err(oc[k], orf[k])
        assert rel < GRAD_TOL, (k, rel)
    if len(ids) == 0:
        return
    report = {}
    for k in GRADS:
        capped, uncapped = util.rel_err(oc[k][ids], orf[k][ids]), util.rel_err(om[k][ids], orf[k][ids])
        report[k] = (capped, uncapped)
        assert capped < SUBSET_TOL, (k, "excluded set", capped, uncapped)
        assert capped < 0.5 * uncapped, (k, "the cap does not explain the difference", capped, uncapped)
    print("%s: excluded set, capped vs uncapped relative L1: %s" % (
        name, ", ".join("%s %.1e/%.1e" % (k[2:], a, b) for k, (a, b) in report.items())))
    # and the uncapped backward really does differ there: the exclusion is not vacuous
    assert report["d_means3D"][1] > 10 * GRAD_TOL
