"""The rasterizer's spherical-harmonics colour inputs (`shs`, `sh_degree`; SURVEY.md Q16 -- reference features VINGS-Mono
never uses) against the reference build: same public call with shs instead of colors_precomp, every degree."""
import pytest
import torch

import synthetic
import util

pytestmark = pytest.mark.gpu


def _run(mod, inp, shs, degree, gc, ga):
    P = inp["means3D"].shape[0]
    dev = inp["means3D"].device
    leaf = lambda t: t.detach().clone().requires_grad_(True)
    means3D, sh, opac = leaf(inp["means3D"]), leaf(shs), leaf(inp["opacities"])
    scales, rots = leaf(inp["scales"]), leaf(inp["rotations"])
    means2D = torch.zeros(P, 3, device=dev, requires_grad=True)
    scores = torch.zeros(P, 2, device=dev, requires_grad=True)
    st = util.make_settings(mod, inp, None)._replace(sh_degree=degree)
    color, radii, allmap = mod.GaussianRasterizer(raster_settings=st)(
        means3D=means3D, means2D=means2D, opacities=opac, shs=sh, colors_precomp=None, scales=scales, rotations=rots,
        cov3D_precomp=None, scores=scores)
    torch.autograd.backward([color, allmap], [gc.to(dev), ga.to(dev)])
    torch.cuda.synchronize()
    return dict(color=color.detach().cpu(), allmap=allmap.detach().cpu(), radii=radii.cpu(), d_sh=sh.grad.cpu(),
                d_means3D=means3D.grad.cpu(), d_opac=opac.grad.cpu(), d_scales=scales.grad.cpu(), d_rot=rots.grad.cpu(),
                scores=scores.grad.cpu())


@pytest.mark.parametrize("degree,n_coeffs", [(0, 1), (0, 16), (1, 4), (2, 9), (3, 16)])
def test_sh_colours_match_reference(degree, n_coeffs):
    mine, ref = util.load_mine(), util.load_ref()
    if ref is None:
        pytest.skip("oracle/_ref not built")
    scene, views = synthetic.make_scene("C1", P=6000, W=256, H=192)
    inp = util.cuda_inputs(synthetic.view_inputs(scene, views[-1]))
    gc, ga = synthetic.upstream_grads(inp["W"], inp["H"], 11)
    g = torch.Generator(device="cpu").manual_seed(degree * 31 + n_coeffs)
    # a wide spread around zero so that the clamp at 0 is hit by a good share of the colours
    shs = (torch.randn(inp["means3D"].shape[0], n_coeffs, 3, generator=g) * 0.8).cuda()
    a, b = _run(mine, inp, shs, degree, gc, ga), _run(ref, inp, shs, degree, gc, ga)
    assert torch.equal(a["radii"], b["radii"])
    # colours differ from the reference's by fp32 contraction order only (one SH polynomial, no gates)
    assert torch.allclose(a["color"], b["color"], rtol=0, atol=2e-5)
    assert torch.allclose(a["allmap"], b["allmap"], rtol=0, atol=1e-6)
    for k in ("d_sh", "d_means3D", "d_opac", "d_scales", "d_rot", "scores"):
        assert util.rel_err(a[k], b[k]) < 5e-4, (k, util.rel_err(a[k], b[k]))
    n = (degree + 1) ** 2
    assert float(a["d_sh"][:, n:].abs().sum()) == 0.0           # coefficients above the active degree
    assert float(a["d_sh"][a["radii"] <= 0].abs().sum()) == 0.0  # surfels that touched no tile
    assert float(a["d_sh"].abs().sum()) > 0.0


def test_sh_and_colors_precomp_are_exclusive():
    mine = util.load_mine()
    scene, views = synthetic.make_scene("C1", P=100, W=64, H=48)
    inp = util.cuda_inputs(synthetic.view_inputs(scene, views[-1]))
    rast = mine.GaussianRasterizer(raster_settings=util.make_settings(mine, inp, None))
    shs = torch.zeros(100, 1, 3, device="cuda")
    with pytest.raises(Exception, match="either SHs or precomputed colors"):
        rast(means3D=inp["means3D"], means2D=None, opacities=inp["opacities"], shs=shs, colors_precomp=inp["colors"],
             scales=inp["scales"], rotations=inp["rotations"])
    with pytest.raises(RuntimeError, match="num_points, coefficients, 3"):
        rast(means3D=inp["means3D"], means2D=None, opacities=inp["opacities"], shs=shs[:, 0], colors_precomp=None,
             scales=inp["scales"], rotations=inp["rotations"])


def _run_precomp(mod, inp, T, gc, ga):
    P = inp["means3D"].shape[0]
    dev = inp["means3D"].device
    leaf = lambda t: t.detach().clone().requires_grad_(True)
    means3D, colors, opac, tm = leaf(inp["means3D"]), leaf(inp["colors"]), leaf(inp["opacities"]), leaf(T)
    means2D = torch.zeros(P, 3, device=dev, requires_grad=True)
    scores = torch.zeros(P, 2, device=dev, requires_grad=True)
    color, radii, allmap = mod.GaussianRasterizer(raster_settings=util.make_settings(mod, inp, None))(
        means3D=means3D, means2D=means2D, opacities=opac, shs=None, colors_precomp=colors, scales=None, rotations=None,
        cov3D_precomp=tm, scores=scores)
    torch.autograd.backward([color, allmap], [gc.to(dev), ga.to(dev)])
    torch.cuda.synchronize()
    z = lambda t: t.grad.cpu() if t.grad is not None else torch.zeros_like(t).cpu()
    return dict(color=color.detach().cpu(), allmap=allmap.detach().cpu(), radii=radii.cpu(), d_T=z(tm), d_colors=z(colors),
                d_opac=z(opac), d_means2D=z(means2D), d_means3D=z(means3D), scores=z(scores))


def test_precomputed_transforms_match_reference():
    """cov3D_precomp = transMat_precomp [P,9] instead of scales + rotations (forward.cu:204-219, backward.cu:875-964):
    fed with the transforms a normal forward computes, the images are the reference's bit for bit (normal = (0,0,1))."""
    mine, ref = util.load_mine(), util.load_ref()
    if ref is None:
        pytest.skip("oracle/_ref not built")
    scene, views = synthetic.make_scene("C1", P=6000, W=256, H=192)
    inp = util.cuda_inputs(synthetic.view_inputs(scene, views[-1]))
    gc, ga = synthetic.upstream_grads(inp["W"], inp["H"], 12)
    T = torch.from_numpy(util.mine_intermediates(mine, inp)["transMat"].copy()).cuda()   # zero rows where culled
    a, b = _run_precomp(mine, inp, T, gc, ga), _run_precomp(ref, inp, T, gc, ga)
    assert int((a["radii"] > 0).sum()) > 1000
    assert torch.equal(a["radii"], b["radii"])
    assert torch.equal(a["color"], b["color"]) and torch.equal(a["allmap"], b["allmap"])
    for k in ("d_T", "d_colors", "d_opac", "d_means2D", "scores"):
        assert util.rel_err(a[k], b[k]) < 5e-4, (k, util.rel_err(a[k], b[k]))
    assert float(a["d_means3D"].abs().sum()) == 0.0   # no path from the loss to the means with a precomputed transform
