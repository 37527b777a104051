"""Shared helpers for the parity tests: drive this package, the reference build (oracle/_ref) and the
CPU emulation (oracle/surfel_emulation.py) on identical seeded inputs."""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref", "diff_surfel_rasterization")
GOLDEN = os.path.join(ROOT, "tests", "golden")


def load_mine():
    import diff_surfel_rasterization as m
    assert os.path.dirname(m.__file__).startswith(os.path.join(ROOT, "vings-mono_b200"))
    return m


def load_ref():
    """The unmodified reference extension built by oracle/build_ref.py, under an alias module name."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    try:
        import ref_loader
    finally:
        sys.path.pop(0)
    return ref_loader.load_ref()


def cuda_inputs(inp):
    return {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in inp.items()}


def make_settings(mod, inp, mask=None):
    W, H = inp["W"], inp["H"]
    dev = inp["means3D"].device
    if mask is None:
        mask = torch.ones(H * W, dtype=torch.bool, device=dev)
    return mod.GaussianRasterizationSettings(
        image_height=H, image_width=W, tanfovx=inp["tanfovx"], tanfovy=inp["tanfovy"], bg=inp["bg"],
        scale_modifier=1.0, viewmatrix=inp["viewmatrix"], projmatrix=inp["projmatrix"], sh_degree=0,
        campos=inp["campos"], prefiltered=False, debug=False, pixel_mask=mask)


GRAD_NAMES = ("means3D", "means2D", "colors", "opacities", "scales", "rotations", "scores")


def run_impl(mod, inp, g_color, g_allmap, mask=None, ref_state=False):
    """One fwd+bwd through the public API (the call VINGS-Mono's mapper makes, gaussian_base.py:193-203).
    Returns CPU tensors.  ref_state=True (reference module only): out["state"] = ref_intermediates() of the buffers
    THIS forward saved for its backward -- the reference's stored max_contrib depends on stale shared memory (quirk
    Q17), so it must be read from the same forward whose backward produced the gradients."""
    P = inp["means3D"].shape[0]
    dev = inp["means3D"].device
    leaf = lambda t: t.detach().clone().requires_grad_(True)
    means3D, colors, opac = leaf(inp["means3D"]), leaf(inp["colors"]), leaf(inp["opacities"])
    scales, rots = leaf(inp["scales"]), leaf(inp["rotations"])
    means2D = torch.zeros(P, 3, device=dev, requires_grad=True)
    scores = torch.zeros(P, 2, device=dev, requires_grad=True)
    rast = mod.GaussianRasterizer(raster_settings=make_settings(mod, inp, mask))
    color, radii, allmap = rast(means3D=means3D, means2D=means2D, opacities=opac, shs=None, colors_precomp=colors,
                                scales=scales, rotations=rots, cov3D_precomp=None, scores=scores)
    state = None
    if ref_state:
        fn = color.grad_fn   # the reference's ctx: saved (colors, means3D, scales, rotations, cov3D, radii, sh, geom,
        sv = fn.saved_tensors  # binning, img, sample) + num_rendered (DSR/diff_surfel_rasterization/__init__.py:101-104)
        state = parse_ref_buffers(P, inp["W"], inp["H"], int(fn.num_rendered), 0, color.detach(), allmap.detach(),
                                  radii, sv[7], sv[8], sv[9])
    torch.autograd.backward([color, allmap], [g_color.to(dev), g_allmap.to(dev)])
    torch.cuda.synchronize()
    out = dict(color=color.detach().cpu(), radii=radii.cpu(), allmap=allmap.detach().cpu())
    if state is not None:
        out["state"] = state
    for name, t in zip(GRAD_NAMES, (means3D, means2D, colors, opac, scales, rots, scores)):
        out["d_" + name] = t.grad.detach().cpu() if t.grad is not None else torch.zeros_like(t).cpu()
    return out


def _a128(x):
    return (x + 127) // 128 * 128


def ref_intermediates(ref, inp):
    """Calls the reference's _C.rasterize_gaussians and slices its opaque byte buffers with the fromChunk
    layouts (rasterizer_impl.cu:228-301; every field 128-B aligned) -- no reference source is modified."""
    P = inp["means3D"].shape[0]
    W, H = inp["W"], inp["H"]
    dev = inp["means3D"].device
    e = torch.empty(0, device=dev)
    R, B, color, others, radii, geom, binning, img, sample = ref._C.rasterize_gaussians(
        inp["bg"], inp["means3D"], inp["colors"], inp["opacities"], inp["scales"], inp["rotations"], 1.0, e,
        inp["viewmatrix"], inp["projmatrix"], inp["tanfovx"], inp["tanfovy"], H, W, e, 0, inp["campos"], False, False)
    return parse_ref_buffers(P, W, H, R, B, color, others, radii, geom, binning, img)


def parse_ref_buffers(P, W, H, R, B, color, others, radii, geom, binning, img):
    torch.cuda.synchronize()
    g = geom.cpu().numpy()
    off = 0
    def take(buf, off, nbytes, dtype):
        off = _a128(off)
        return np.frombuffer(buf[off:off + nbytes].tobytes(), dtype=dtype), off + nbytes
    depths, off = take(g, off, 4 * P, np.float32)
    _, off = take(g, off, 3 * P, np.uint8)
    _, off = take(g, off, 4 * P, np.int32)
    means2D, off = take(g, off, 8 * P, np.float32)
    transMat, off = take(g, off, 36 * P, np.float32)
    nrm_opa, off = take(g, off, 16 * P, np.float32)
    _, off = take(g, off, 12 * P, np.float32)
    tiles, off = take(g, off, 4 * P, np.uint32)
    b = binning.cpu().numpy()
    off = 0
    point_list, off = take(b, off, 4 * R, np.uint32)
    vals_unsorted, off = take(b, off, 4 * R, np.uint32)
    keys, off = take(b, off, 8 * R, np.uint64)
    keys_unsorted, off = take(b, off, 8 * R, np.uint64)
    im = img.cpu().numpy()
    N = W * H
    off = 0
    accum, off = take(im, off, 12 * N, np.float32)
    n_contrib, off = take(im, off, 8 * N, np.uint32)
    ranges, off = take(im, off, 8 * N, np.uint32)
    tiles_n = ((W + 15) // 16) * ((H + 15) // 16)
    # max_contrib sits one (128-B aligned) N-word array before pixel_colors, which is a copy of out_color
    # (rasterizer_impl.cu:258-259,474); the cub scan scratch in between has an unknown size, so search for it
    max_contrib = None
    want = np.frombuffer(color.cpu().numpy().tobytes(), dtype=np.uint8)
    nzs = np.flatnonzero(want)
    if len(nzs):
        # candidates are screened on a 4-KiB window around the first non-zero byte before the full comparison
        # (the full image is 128 MB at 4112x2600 and zero background matches scratch memory everywhere)
        p0 = int(nzs[0]) // 128 * 128
        probe = want[p0:p0 + 4096]
        o0 = _a128(off)
        # every 128-B aligned candidate offset at once: the 8 bytes at (candidate + p0 + first non-zero byte's 8-B word)
        w0 = (int(nzs[0]) - p0) // 8 * 8
        tail = im[o0 + p0 + w0:]
        rows = (len(tail) - 8) // 128 + 1 if len(tail) >= 8 else 0
        if rows > 0:
            first8 = np.lib.stride_tricks.as_strided(tail, shape=(rows, 8), strides=(128, 1))
            key = want[p0 + w0:p0 + w0 + 8]
            for r in np.flatnonzero((first8 == key).all(axis=1)):
                o = o0 + int(r) * 128
                if o + len(want) > len(im):
                    break
                if np.array_equal(im[o + p0:o + p0 + len(probe)], probe) and np.array_equal(im[o:o + len(want)], want):
                    o_mc = o - _a128(4 * N)
                    max_contrib = np.frombuffer(im[o_mc:o_mc + 4 * tiles_n].tobytes(), dtype=np.uint32)
                    break
    return dict(R=R, B=B, max_contrib=max_contrib, color=color.cpu(), allmap=others.cpu(), radii=radii.cpu().numpy(), depths=depths,
                means2D=means2D.reshape(P, 2), transMat=transMat.reshape(P, 9), normal_opacity=nrm_opa.reshape(P, 4),
                tiles_touched=tiles, point_list=point_list, vals_unsorted=vals_unsorted, keys=keys,
                keys_unsorted=keys_unsorted, final_T=accum.reshape(3, N), n_contrib=n_contrib.reshape(2, N),
                ranges=ranges.reshape(N, 2)[:tiles_n])


def mine_intermediates(mine, inp):
    """Drives the C ABI directly (forward_begin / forward_finish) and reads the workspaces."""
    lib = mine._load()
    P = inp["means3D"].shape[0]
    W, H = inp["W"], inp["H"]
    dev = inp["means3D"].device
    vp = ctypes.c_void_p
    ptr = lambda t: vp(t.data_ptr())
    st = vp(torch.cuda.current_stream().cuda_stream)
    geom = torch.zeros(lib.sb200_geom_bytes(P), dtype=torch.uint8, device=dev)
    image = torch.zeros(lib.sb200_image_bytes(W, H), dtype=torch.uint8, device=dev)
    radii = torch.zeros(P, dtype=torch.int32, device=dev)
    count = torch.zeros(1, dtype=torch.int64).pin_memory()
    rc = lib.sb200_forward_begin(P, W, H, ptr(inp["means3D"]), ptr(inp["scales"]), ptr(inp["rotations"]),
                                 ptr(inp["opacities"]), ptr(inp["colors"]), ptr(inp["viewmatrix"]),
                                 ptr(inp["projmatrix"]), ptr(geom), ptr(radii), vp(count.data_ptr()), st)
    assert rc == 0, lib.sb200_last_error()
    R = int(count[0])
    binning = torch.zeros(lib.sb200_binning_bytes(R), dtype=torch.uint8, device=dev)
    # pair list = sorted ids [R] + the forward's hit masks (include/surfel_b200.h)
    point_list = torch.zeros(lib.sb200_pair_list_bytes(R, W, H) // 4, dtype=torch.int32, device=dev)
    color = torch.zeros(3, H, W, device=dev)
    others = torch.zeros(7, H, W, device=dev)
    rc = lib.sb200_forward_finish(P, W, H, R, ptr(inp["bg"]), ptr(geom), ptr(binning), ptr(point_list), ptr(image),
                                  ptr(color), ptr(others), st)
    assert rc == 0, lib.sb200_last_error()
    torch.cuda.synchronize()
    go = (ctypes.c_size_t * 5)(); lib.sb200_geom_layout(P, go)
    io = (ctypes.c_size_t * 3)(); lib.sb200_image_layout(W, H, io)
    bo = (ctypes.c_size_t * 4)(); lib.sb200_binning_layout(R, bo)
    g = geom.cpu().numpy(); im = image.cpu().numpy(); b = binning.cpu().numpy()
    fb = lambda buf, off, n, dt: np.frombuffer(buf[off:off + n].tobytes(), dtype=dt)
    rec = fb(g, go[0], 128 * P, np.float32).reshape(P, 32)  # 8 quads, layout in csrc/common.cuh
    N = W * H
    tiles_n = ((W + 15) // 16) * ((H + 15) // 16)
    ranges = fb(im, io[0], 8 * tiles_n, np.uint32).reshape(tiles_n, 2)
    depths = fb(g, go[1], 4 * P, np.float32)
    pl = point_list.cpu().numpy()[:R].view(np.uint32)
    # the reference's sorted 64-bit keys, restated from what this pipeline keeps: the tile of a list position comes
    # from the ranges, the depth word from the surfel at that position
    tile_of = np.zeros(R, dtype=np.uint64)
    for t in np.nonzero(ranges[:, 1] > ranges[:, 0])[0]:
        tile_of[ranges[t, 0]:ranges[t, 1]] = t
    keys_sorted = (tile_of << np.uint64(32)) | depths.view(np.uint32)[pl].astype(np.uint64)
    V = int((radii > 0).sum())
    vis = radii.cpu().numpy() > 0
    return dict(R=R, color=color.cpu(), allmap=others.cpu(), radii=radii.cpu().numpy(), visible=vis,
                depths=depths, rect=fb(g, go[2], 8 * P, np.uint16).reshape(P, 4),
                tiles_touched=fb(g, go[3], 4 * P, np.uint32), surfel_order=fb(g, go[4], 4 * P, np.uint32)[:V],
                emitted_tiles=fb(b, bo[0], 4 * R, np.uint32), emitted_ids=fb(b, bo[2], 4 * R, np.uint32),
                means2D=np.stack([rec[:, 3], rec[:, 7]], -1),
                transMat=np.concatenate([rec[:, 12:15], rec[:, 16:19], rec[:, 20:23]], -1),
                normal_opacity=np.concatenate([rec[:, 24:27], rec[:, 15:16]], -1),
                cull_tau=rec[:, 11], bbox=rec[:, 28:32],
                keys=keys_sorted, point_list=pl, ranges=ranges,
                final_T=fb(im, io[1], 12 * N, np.float32).reshape(3, N),
                n_contrib=fb(im, io[2], 8 * N, np.uint32).reshape(2, N))


def _higher_msb(n):
    msb, step = 16, 16
    while step > 1:
        step //= 2
        msb = msb + step if n >> msb else msb - step
    if n >> msb:
        msb += 1
    return msb


def rel_err(a, b):
    a, b = a.double(), b.double()
    return float((a - b).abs().sum() / b.abs().sum().clamp(min=1e-30))


def tainted_surfels(ri, W, H):
    """Surfels whose reference gradient is garbage-dependent (reference quirk Q17): the
    reference's forward reduces `max_contrib` with cub::BlockReduce<uint32_t, 256> on a 16x16 block
    (forward.cu:496-502), which mis-indexes the warp aggregates, so the stored value is max(rows 0-1 of the
    tile, stale shared memory) rather than the tile maximum; its backward then drops every 32-surfel bucket
    starting at or beyond that value (backward.cu:503).  Returns (ids of the surfels sitting in dropped
    buckets that some pixel still needs, number of affected tiles)."""
    mc = ri["max_contrib"]
    if mc is None:
        raise RuntimeError("max_contrib not located in the reference image buffer")
    gx, gy = (W + 15) // 16, (H + 15) // 16
    last = np.zeros((gy * 16, gx * 16), np.int64)
    last[:H, :W] = ri["n_contrib"][0].reshape(H, W)
    true_max = last.reshape(gy, 16, gx, 16).max(axis=(1, 3)).reshape(-1)
    need = (true_max + 31) // 32           # buckets some pixel needs
    done = (mc.astype(np.int64) + 31) // 32  # buckets the reference processes
    ids = []
    bad = np.nonzero(done < need)[0]
    for t in bad:
        r0, r1 = ri["ranges"][t]
        ids.append(ri["point_list"][int(r0) + 32 * int(done[t]):int(r1)])
    ids = np.unique(np.concatenate(ids)) if ids else np.zeros(0, np.int64)
    return ids.astype(np.int64), len(bad)


def sync_mapper_state(dst, src):
    """Copy parameters, Adam moments and score bookkeeping of mapper `src` into mapper `dst` (same surfel count).

    Trajectory tests compare two implementations of the mapping step (fused kernels vs torch autograd).  Gradient
    sums differ by atomic-order jitter between ANY two runs, and Adam's m / sqrt(v) normalisation amplifies that
    from step to step, so a bound on "loss after k steps" is a bound on chaos, not on the code.  The tests therefore
    re-synchronise the two mappers after every step and compare ONE step from bit-identical state: the loss then
    depends only on the (deterministic, atomics-free) forward and on the loss kernel's summation order."""
    import mapper_step
    assert dst.P == src.P
    with torch.no_grad():
        for k in mapper_step.GROUPS:
            pd, ps = getattr(dst, k), getattr(src, k)
            pd.data.copy_(ps.data)
            sd, ss = dst.optimizer.state[pd], src.optimizer.state[ps]
            if len(ss) == 0:
                assert len(sd) == 0 or float(sd["exp_avg"].abs().sum()) == 0.0
                continue
            if len(sd) == 0:   # a fresh torch-side optimizer: the other side's moments must be zeros too
                assert float(ss["exp_avg"].abs().sum()) == 0.0 and float(ss["exp_avg_sq"].abs().sum()) == 0.0, k
                continue
            sd["exp_avg"].copy_(ss["exp_avg"])
            sd["exp_avg_sq"].copy_(ss["exp_avg_sq"])
        for name in dst.BOOK:
            cur = getattr(dst, name)
            new = getattr(src, name).detach().reshape(cur.shape).to(cur.dtype)
            if dst.fused:
                cur.copy_(new)   # views into the surfel store
            else:
                setattr(dst, name, new.clone())


# Cauchy-Schwarz over the two exponential windows: |m| / sqrt(v) <= sqrt(sum_j (1-b1)^2 b1^2j / ((1-b2) b2^j)) = 7.3
ADAM_RATIO_MAX = 7.3


def mapper_snapshot(m):
    """Parameters and moments of a mapper (clones), for assert_one_step_agrees."""
    import mapper_step
    out = {}
    for k in mapper_step.GROUPS:
        p = getattr(m, k)
        st = m.optimizer.state[p]
        z = torch.zeros_like(p.data)
        out[k] = (p.data.clone(), st["exp_avg"].clone() if len(st) else z, st["exp_avg_sq"].clone() if len(st) else z.clone())
    return out


def assert_one_step_agrees(a, b, before, ctx=""):
    """Mappers `a` (torch path) and `b` (fused path) took ONE training step from the bit-identical state `before`
    (mapper_snapshot after sync_mapper_state) through different implementations.  Checked separately:

      1. what is LINEAR in the gradient agrees across the paths: the Adam moments (relative L1 2e-3: the gradients
         agree to atomic-order jitter, <= GRAD_TOL, plus the loss's fp32 summation order, 2e-5) and the scores;
      2. the fused path's Adam ARITHMETIC, against its own new moments: p = p_old - lr * m / (sqrt(v) + eps)
         (adam.cu:26-37) on the rows whose moments moved, p = p_old elsewhere -- deterministic, a few ulp;
      3. across the paths a parameter can then only differ through m / sqrt(v), which is sign-like right after a
         densify (zero moments: exactly +-3.16 whatever |g| is) and flips wherever a gradient sum cancels below the
         jitter; no fixed fraction bounds that, only the size of a step does: |p_a - p_b| <= 2 * 7.3 * lr."""
    import mapper_step
    eps = 1e-15
    for k in mapper_step.GROUPS:
        pa, pb = getattr(a, k).data, getattr(b, k).data
        sa, sb = a.optimizer.state[getattr(a, k)], b.optimizer.state[getattr(b, k)]
        assert rel_err(sb["exp_avg"], sa["exp_avg"]) < 2e-3, (ctx, k, rel_err(sb["exp_avg"], sa["exp_avg"]))
        assert rel_err(sb["exp_avg_sq"], sa["exp_avg_sq"]) < 2e-3, (ctx, k, rel_err(sb["exp_avg_sq"], sa["exp_avg_sq"]))
        lr = mapper_step.LR[k]
        p0, m0, v0 = before[k]
        m1, v1 = sb["exp_avg"], sb["exp_avg_sq"]
        moved = ((m1 != m0) | (v1 != v0)).any(dim=1, keepdim=True)
        want = torch.where(moved, p0 - lr * m1 / (v1.sqrt() + eps), p0)
        tol = 4e-7 * (p0.abs() + lr * ADAM_RATIO_MAX)
        assert bool(((pb - want).abs() <= tol).all()), (ctx, k, float((pb - want).abs().max()))
        assert bool(moved.any()), (ctx, k)
        assert bool(((pa - pb).abs() <= 2.0 * ADAM_RATIO_MAX * lr + 2.5e-7 * pa.abs()).all()), (ctx, k)
    assert rel_err(b._local_scores, a._local_scores) < 1e-4, ctx
    assert rel_err(b._global_scores, a._global_scores) < 1e-4, ctx
    assert rel_err(b._globalkf_max_scores, a._globalkf_max_scores) < 1e-4, ctx
    assert float((a._globalkf_id != b._globalkf_id).float().mean()) < 1e-3, ctx
