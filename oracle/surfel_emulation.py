"""TEST INFRASTRUCTURE (oracle) -- pure-PyTorch fp32 CPU emulation of the reference surfel rasterizer.

NOT part of the product: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
reference legs may import this file.  The product path (vings-mono_b200/) never does.

Restates, function by function, the reference's CUDA (paths relative to
/root/reference/submodules/diff-surfel-rasterization/):
  preprocess()      cuda_rasterizer/forward.cu:83-157,160-249 + auxiliary.h:66-76,91-130,197-302
  binning()         cuda_rasterizer/rasterizer_impl.cu:123-196 (duplicateWithKeys, stable sort, identifyTileRanges)
  blend_forward()   cuda_rasterizer/forward.cu:254-504 (renderCUDA, including the per-32-splat checkpoints)
  blend_backward()  cuda_rasterizer/backward.cu:445-845 (PerGaussianRenderCUDA -- the LIVE backward, with its quirks)
  preprocess_backward()  cuda_rasterizer/backward.cu:849-1045
  rasterize() / EmulatedRasterizer   diff_surfel_rasterization/__init__.py:46-169 (argument and gradient order)

Pinning: the reference has no tests or golden vectors for this path, so this
emulation is pinned by running the reference's own CUDA build (oracle/_ref, built by
oracle/build_ref.py) on a B200 and committing its outputs as fixtures under tests/golden/ (see
tests/golden/make_golden.py); tests/test_oracle_golden.py checks this file against those fixtures.

The blend is evaluated in lock-step over the position j inside each tile's sorted range, vectorised
over all pixels, so a 320x240 view takes seconds.  fp32 throughout; CPU torch does not contract
FMAs, so values agree with the CUDA reference to rounding, not bit for bit.
"""
import math

import numpy as np
import torch

TILE = 16
NEAR = 0.2
FAR = 100.0
FILTER_INV_SQUARE = 2.0
F32 = torch.float32


def higher_msb(n: int) -> int:
    """rasterizer_impl.cu:35-50"""
    msb = 16
    step = msb
    while step > 1:
        step //= 2
        if n >> msb:
            msb += step
        else:
            msb -= step
    if n >> msb:
        msb += 1
    return msb


def _quat_to_rotmat(q):
    """auxiliary.h:225-245; q stored (w,x,y,z).  Returns R[P,3(col),3(row)] (column-major like glm)."""
    s = torch.rsqrt(q[:, 3] * q[:, 3] + q[:, 0] * q[:, 0] + q[:, 1] * q[:, 1] + q[:, 2] * q[:, 2])
    w, x, y, z = q[:, 0] * s, q[:, 1] * s, q[:, 2] * s, q[:, 3] * s
    c0 = torch.stack([1.0 - 2.0 * (y * y + z * z), 2.0 * (x * y + w * z), 2.0 * (x * z - w * y)], -1)
    c1 = torch.stack([2.0 * (x * y - w * z), 1.0 - 2.0 * (x * x + z * z), 2.0 * (y * z + w * x)], -1)
    c2 = torch.stack([2.0 * (x * z + w * y), 2.0 * (y * z - w * x), 1.0 - 2.0 * (x * x + y * y)], -1)
    return torch.stack([c0, c1, c2], 1)


def _transmat(means3D, scales, rotations, proj, W, H, backward_assoc=False):
    """forward.cu:83-126 (T = (splat2world^T * world2ndc) * ndc2pix) or, with backward_assoc, the
    association of backward.cu:909-910 (T = splat2world^T * (world2ndc * ndc2pix)).
    Returns T[P,3(col: Tu,Tv,Tw),3], L[P,3,3], R[P,3,3], Pm[3,4]."""
    P = means3D.shape[0]
    R = _quat_to_rotmat(rotations)
    L = R.clone()
    L[:, 0, :] = R[:, 0, :] * scales[:, 0:1]
    L[:, 1, :] = R[:, 1, :] * scales[:, 1:2]
    # splat2world^T as 4 columns of 3: col j = (L0[j], L1[j], p[j]) for j<3, col 3 = (0,0,1)
    s2wT = torch.zeros(P, 4, 3, dtype=F32)
    s2wT[:, 0:3, 0] = L[:, 0, :]
    s2wT[:, 0:3, 1] = L[:, 1, :]
    s2wT[:, 0:3, 2] = means3D
    s2wT[:, 3, 2] = 1.0
    pm = proj.reshape(16)
    w2n = pm.reshape(4, 4).t().contiguous()  # w2n[col][row] = pm[4*row+col]
    n2p = torch.tensor([[W / 2.0, 0, 0, (W - 1) / 2.0], [0, H / 2.0, 0, (H - 1) / 2.0], [0, 0, 0, 1.0]], dtype=F32)
    Pm = torch.zeros(3, 4, dtype=F32)
    for c in range(3):
        for r in range(4):
            Pm[c, r] = w2n[0, r] * n2p[c, 0] + w2n[1, r] * n2p[c, 1] + w2n[2, r] * n2p[c, 2] + w2n[3, r] * n2p[c, 3]
    T = torch.zeros(P, 3, 3, dtype=F32)
    if backward_assoc:
        for c in range(3):
            for r in range(3):
                T[:, c, r] = s2wT[:, 0, r] * Pm[c, 0] + s2wT[:, 1, r] * Pm[c, 1] + s2wT[:, 2, r] * Pm[c, 2] + \
                    s2wT[:, 3, r] * Pm[c, 3]
    else:
        A = torch.zeros(P, 4, 3, dtype=F32)
        for c in range(4):
            for r in range(3):
                A[:, c, r] = s2wT[:, 0, r] * w2n[c, 0] + s2wT[:, 1, r] * w2n[c, 1] + s2wT[:, 2, r] * w2n[c, 2] + \
                    s2wT[:, 3, r] * w2n[c, 3]
        for c in range(3):
            for r in range(3):
                T[:, c, r] = A[:, 0, r] * n2p[c, 0] + A[:, 1, r] * n2p[c, 1] + A[:, 2, r] * n2p[c, 2] + \
                    A[:, 3, r] * n2p[c, 3]
    return T, L, R, Pm


def _view_normal(L, view):
    vm = view.reshape(16)
    l2 = L[:, 2, :]
    n = torch.stack([vm[0] * l2[:, 0] + vm[4] * l2[:, 1] + vm[8] * l2[:, 2],
                     vm[1] * l2[:, 0] + vm[5] * l2[:, 1] + vm[9] * l2[:, 2],
                     vm[2] * l2[:, 0] + vm[6] * l2[:, 1] + vm[10] * l2[:, 2]], -1)
    mult = torch.where(n[:, 2] < 0, torch.ones(()), -torch.ones(()))
    return n * mult[:, None], mult


def _get_rect(xy, radius_i, grid_x, grid_y):
    """auxiliary.h:66-76 with an int radius."""
    r = radius_i.to(F32)
    def lo(p, g):
        return torch.clamp(((p - r) / TILE).to(torch.int64).clamp(min=0), max=g)
    def hi(p, g):
        return torch.clamp(((p + r + TILE - 1) / TILE).to(torch.int64).clamp(min=0), max=g)
    return lo(xy[:, 0], grid_x), lo(xy[:, 1], grid_y), hi(xy[:, 0], grid_x), hi(xy[:, 1], grid_y)


def _fma32(a, b, c):
    """fp32 fused multiply-add: the product is exact in float64, one rounding to float32 at the end."""
    return (a.double() * b.double() + c.double()).float()


def preprocess(means3D, scales, rotations, opacities, view, proj, W, H):
    P = means3D.shape[0]
    vm = view.reshape(16)
    # View depth with the FMA contraction of the CUDA build (FMUL on m6*y, then two FFMAs, then FADD): its bits
    # are the low word of the sort key, so ties between equal-depth surfels must break as on the GPU.
    pz = _fma32(means3D[:, 2], vm[10], _fma32(means3D[:, 0], vm[2], means3D[:, 1] * vm[6])) + vm[14]
    T, L, R, _ = _transmat(means3D, scales, rotations, proj, W, H)
    normal, _ = _view_normal(L, view)
    Tu, Tv, Tw = T[:, 0, :], T[:, 1, :], T[:, 2, :]
    sign = torch.tensor([1.0, 1.0, -1.0])
    dist = (Tw[:, 0] * Tw[:, 0] + Tw[:, 1] * Tw[:, 1]) + Tw[:, 2] * Tw[:, 2] * -1.0
    f = (1.0 / dist)[:, None] * sign[None, :]
    def s3(a):
        return (a[:, 0] + a[:, 1]) + a[:, 2]
    cx, cy = s3(f * Tu * Tw), s3(f * Tv * Tw)
    tx, ty = s3(f * Tu * Tu), s3(f * Tv * Tv)
    ex = torch.sqrt(torch.clamp(cx * cx - tx, min=1e-4))
    ey = torch.sqrt(torch.clamp(cy * cy - ty, min=1e-4))
    radius = 3.0 * torch.ceil(torch.maximum(ex, ey))
    xy = torch.stack([cx, cy], -1)
    grid_x, grid_y = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    ok = ~(pz <= 0.2) & (dist != 0)
    radius_i = torch.where(ok, radius, torch.zeros(())).nan_to_num(0.0, 2.0e9, 0.0).clamp(max=2.0e9).to(torch.int64)
    x0, y0, x1, y1 = _get_rect(torch.where(ok[:, None], xy, torch.zeros(())).nan_to_num(0.0, 1e9, -1e9), radius_i,
                               grid_x, grid_y)
    area = (x1 - x0) * (y1 - y0)
    vis = ok & (area > 0)
    return dict(radii=torch.where(vis, radius_i, torch.zeros((), dtype=torch.int64)).to(torch.int32), visible=vis,
                depth=pz, means2D=xy, T=T, normal=normal,
                tiles_touched=torch.where(vis, area, torch.zeros((), dtype=torch.int64)),
                rect=torch.stack([x0, y0, x1, y1], -1), grid=(grid_x, grid_y))


def binning(pre, W, H):
    """Keys (tile<<32 | depth bits) in emission order, stable sort on the reference's bit range, ranges."""
    grid_x, grid_y = pre["grid"]
    vis = pre["visible"].numpy()
    rect = pre["rect"].numpy()
    depth_bits = pre["depth"].numpy().view(np.uint32).astype(np.uint64)
    idxs = np.nonzero(vis)[0]
    keys, vals = [], []
    for i in idxs:
        x0, y0, x1, y1 = rect[i]
        ys, xs = np.meshgrid(np.arange(y0, y1), np.arange(x0, x1), indexing="ij")
        t = (ys * grid_x + xs).reshape(-1).astype(np.uint64)
        keys.append((t << np.uint64(32)) | depth_bits[i])
        vals.append(np.full(t.shape, i, dtype=np.uint32))
    keys = np.concatenate(keys) if keys else np.zeros(0, np.uint64)
    vals = np.concatenate(vals) if vals else np.zeros(0, np.uint32)
    bits = 32 + higher_msb(grid_x * grid_y)
    masked = keys & np.uint64((1 << bits) - 1)
    order = np.argsort(masked, kind="stable")
    skeys, svals = keys[order], vals[order]
    tiles = grid_x * grid_y
    ranges = np.zeros((tiles, 2), np.uint32)
    if len(skeys):
        tid = (skeys >> np.uint64(32)).astype(np.int64)
        change = np.nonzero(np.diff(tid))[0] + 1
        starts = np.concatenate([[0], change])
        ends = np.concatenate([change, [len(skeys)]])
        ranges[tid[starts], 0] = starts
        ranges[tid[starts], 1] = ends
    return dict(keys_unsorted=keys, vals_unsorted=vals, keys=skeys, point_list=svals, ranges=ranges, num_rendered=len(keys))


def _pixel_grid(W, H):
    grid_x = (W + TILE - 1) // TILE
    py, px = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    px, py = px.reshape(-1), py.reshape(-1)
    tile = (py // TILE) * grid_x + (px // TILE)
    return px.to(F32), py.to(F32), tile


def _pair_geometry(T, xy, g, pxf, pyf):
    """forward.cu:401-417 / backward.cu:673-686 for pixel vectors and gathered surfel ids g."""
    Tu, Tv, Tw = T[g, 0, :], T[g, 1, :], T[g, 2, :]
    k = pxf[:, None] * Tw - Tu
    l = pyf[:, None] * Tw - Tv
    p = torch.stack([k[:, 1] * l[:, 2] - k[:, 2] * l[:, 1], k[:, 2] * l[:, 0] - k[:, 0] * l[:, 2],
                     k[:, 0] * l[:, 1] - k[:, 1] * l[:, 0]], -1)
    sx, sy = p[:, 0] / p[:, 2], p[:, 1] / p[:, 2]
    rho3d = sx * sx + sy * sy
    dx, dy = xy[g, 0] - pxf, xy[g, 1] - pyf
    rho2d = FILTER_INV_SQUARE * (dx * dx + dy * dy)
    use3d = rho3d <= rho2d  # False for NaN, like the CUDA comparison
    depth = torch.where(use3d, (sx * Tw[:, 0] + sy * Tw[:, 1]) + Tw[:, 2], Tw[:, 2])
    rho = torch.where(torch.isnan(rho3d), rho2d, torch.minimum(rho3d, rho2d))  # fminf semantics
    return dict(k=k, l=l, p=p, sx=sx, sy=sy, rho3d=rho3d, rho2d=rho2d, use3d=use3d, depth=depth, rho=rho, dx=dx,
                dy=dy, Tw=Tw)


def blend_forward(pre, bins, colors, opacities, bg, W, H, keep_checkpoints=True):
    HW = W * H
    pxf, pyf, tile = _pixel_grid(W, H)
    ranges = torch.from_numpy(bins["ranges"].astype(np.int64))
    plist = torch.from_numpy(bins["point_list"].astype(np.int64))
    start, end = ranges[tile, 0], ranges[tile, 1]
    max_len = int((end - start).max()) if HW else 0
    T, xy, normal = pre["T"], pre["means2D"], pre["normal"]
    opa = opacities.reshape(-1)
    Tr = torch.ones(HW)
    C = torch.zeros(HW, 3)
    N = torch.zeros(HW, 3)
    D = torch.zeros(HW); M1 = torch.zeros(HW); M2 = torch.zeros(HW); dist = torch.zeros(HW)
    med_d = torch.zeros(HW); med_c = -torch.ones(HW)
    last = torch.zeros(HW, dtype=torch.int64)
    done = torch.zeros(HW, dtype=torch.bool)
    ckpts = []
    k_far = FAR / (FAR - NEAR)
    for j in range(max_len):
        if j % 32 == 0 and keep_checkpoints:  # forward.cu:373-394 (done pixels keep stale values; never read)
            ckpts.append(dict(T=Tr.clone(), C=C.clone(), D=D.clone(), A=(1 - Tr), N=N.clone(), reg=dist.clone(),
                              M1=M1.clone(), M2=M2.clone()))
        live = (~done) & (start + j < end)
        if not bool(live.any()):
            if bool(((start + j) >= end).all() | done.all()):
                break
            continue
        g = plist[torch.where(live, start + j, torch.zeros((), dtype=torch.int64)).clamp(max=max(len(plist) - 1, 0))]
        q = _pair_geometry(T, xy, g, pxf, pyf)
        ok = live & (q["p"][:, 2] != 0) & ~(q["depth"] < NEAR)
        power = -0.5 * q["rho"]
        ok = ok & ~(power > 0)
        alpha = torch.minimum(torch.tensor(0.99), opa[g] * torch.exp(power))
        ok = ok & ~(alpha < 1.0 / 255.0)
        test_T = Tr * (1 - alpha)
        stop = ok & (test_T < 0.0001)
        done = done | stop
        ok = ok & ~stop
        w = alpha * Tr
        A = 1 - Tr
        m = k_far * (1 - NEAR / q["depth"])
        okf = ok.to(F32)
        def acc(dst, val):
            return torch.where(ok, dst + val, dst)
        dist = acc(dist, (m * m * A + M2 - 2 * m * M1) * w)
        D = acc(D, q["depth"] * w)
        M1n = acc(M1, m * w)
        M2 = acc(M2, m * m * w)
        M1 = M1n
        med = ok & (Tr > 0.5)
        med_d = torch.where(med, q["depth"], med_d)
        med_c = torch.where(med, torch.full((), float(j + 1)), med_c)
        N = torch.where(ok[:, None], N + normal[g] * w[:, None], N)
        C = torch.where(ok[:, None], C + colors[g] * w[:, None], C)
        Tr = torch.where(ok, test_T, Tr)
        last = torch.where(ok, torch.full((), j + 1, dtype=torch.int64), last)
    color = (C + Tr[:, None] * bg[None, :]).t().reshape(3, H, W).contiguous()
    allmap = torch.stack([D, 1 - Tr, N[:, 0], N[:, 1], N[:, 2], med_d, dist], 0).reshape(7, H, W).contiguous()
    return dict(color=color, allmap=allmap, final_T=torch.stack([Tr, M1, M2], 0), n_contrib=last,
                median_contrib=med_c.clamp(min=0).to(torch.int64), checkpoints=ckpts, max_len=max_len)


def blend_backward(pre, bins, fwd, colors, opacities, bg, dL_dcolor, dL_dallmap, pixel_mask, W, H):
    """Live backward, restated with the reference's own structure: per-bucket reload of the forward
    checkpoint, 'ar' = checkpoint - final, T re-propagated forward, M1/M2 frozen per bucket."""
    HW = W * H
    P = colors.shape[0]
    pxf, pyf, tile = _pixel_grid(W, H)
    ranges = torch.from_numpy(bins["ranges"].astype(np.int64))
    plist = torch.from_numpy(bins["point_list"].astype(np.int64))
    start, end = ranges[tile, 0], ranges[tile, 1]
    T, xy, normal = pre["T"], pre["means2D"], pre["normal"]
    opa = opacities.reshape(-1)
    mask = pixel_mask.reshape(-1).bool()
    last = fwd["n_contrib"]
    T_final, M1_final = fwd["final_T"][0], fwd["final_T"][1]
    fin_C = fwd["color"].reshape(3, HW).t()
    fin = fwd["allmap"].reshape(7, HW)
    dC = dL_dcolor.reshape(3, HW).t()
    dO = dL_dallmap.reshape(7, HW)
    dD, dA, dN, dReg = dO[0], dO[1], dO[2:5].t(), dO[6]
    acc = torch.zeros(P, 20)
    k_far = FAR / (FAR - NEAR)
    max_walk = int(torch.where(mask, last, torch.zeros((), dtype=torch.int64)).max()) if HW else 0
    for j in range(max_walk):
        if j % 32 == 0:  # backward.cu:625-657
            ck = fwd["checkpoints"][j // 32]
            Tr = ck["T"].clone()
            ar = -fin_C + ck["C"]
            M1b, M2b = ck["M1"], ck["M2"]
            ar_d = -fin[0] + ck["D"]
            ar_a = -fin[1] + ck["A"]
            ar_n = -fin[2:5].t() + ck["N"]
            ar_reg = -fin[6] + ck["reg"]
        act = mask & (j < last) & (start + j < end)
        if not bool(act.any()):
            continue
        g = plist[torch.where(act, start + j, torch.zeros((), dtype=torch.int64)).clamp(max=max(len(plist) - 1, 0))]
        q = _pair_geometry(T, xy, g, pxf, pyf)
        c_d = q["depth"]
        m_d = k_far * (1 - NEAR / c_d)
        dmd_dd = (FAR * NEAR) / ((FAR - NEAR) * c_d * c_d)
        power = -0.5 * q["rho"]
        G = torch.exp(power)
        alpha = torch.minimum(torch.tensor(0.99), opa[g] * G)
        ok = act & ~(power > 0) & ~(alpha < 1.0 / 255.0)
        if not bool(ok.any()):
            continue
        w = alpha * Tr
        inv = 1.0 / (1.0 - alpha)
        col, nrm = colors[g], normal[g]
        ar = torch.where(ok[:, None], ar + (Tr * alpha)[:, None] * col, ar)
        g_col = w[:, None] * dC
        dL_dalpha = (((col * Tr[:, None]) - inv[:, None] * (-ar)) * dC).sum(-1)
        bg_dot = (bg[None, :] * dC).sum(-1)
        s1 = (w[:, None] * dC).abs().sum(-1)
        dL_dalpha = dL_dalpha + (-T_final / (1.0 - alpha)) * bg_dot
        ar_d = torch.where(ok, ar_d + Tr * alpha * c_d, ar_d)
        dL_dalpha = dL_dalpha + ((c_d * Tr) - inv * (-ar_d)) * dD
        dL_dz = alpha * Tr * dD
        ar_a = torch.where(ok, ar_a + Tr * alpha, ar_a)
        dL_dalpha = dL_dalpha + (Tr - inv * (-ar_a)) * dA
        ar_n = torch.where(ok[:, None], ar_n + (Tr * alpha)[:, None] * nrm, ar_n)
        g_nrm = w[:, None] * dN
        dL_dalpha = dL_dalpha + (((nrm * Tr[:, None]) - inv[:, None] * (-ar_n)) * dN).sum(-1)
        dist = m_d * m_d * (1 - Tr) + M2b - 2 * m_d * M1b
        ar_reg = torch.where(ok, ar_reg + Tr * alpha * dist, ar_reg)
        dL_dalpha = dL_dalpha + ((dist * Tr) - inv * (-ar_reg)) * dReg
        dL_dmd = 2.0 * (Tr * alpha) * (m_d * (1 - T_final) - M1_final) * dReg
        dL_dz = dL_dz + dL_dmd * dmd_dd
        Tr = torch.where(ok, Tr * (1.0 - alpha), Tr)
        dL_dG = opa[g] * dL_dalpha
        use3d = q["use3d"]
        sx, sy, p, k, l, Tw = q["sx"], q["sy"], q["p"], q["k"], q["l"], q["Tw"]
        dsx = dL_dG * -G * sx + dL_dz * Tw[:, 0]
        dsy = dL_dG * -G * sy + dL_dz * Tw[:, 1]
        dsx_pz, dsy_pz = dsx / p[:, 2], dsy / p[:, 2]
        dp = torch.stack([dsx_pz, dsy_pz, -(dsx_pz * sx + dsy_pz * sy)], -1)
        dk = torch.stack([l[:, 1] * dp[:, 2] - l[:, 2] * dp[:, 1], l[:, 2] * dp[:, 0] - l[:, 0] * dp[:, 2],
                          l[:, 0] * dp[:, 1] - l[:, 1] * dp[:, 0]], -1)
        dl = torch.stack([dp[:, 1] * k[:, 2] - dp[:, 2] * k[:, 1], dp[:, 2] * k[:, 0] - dp[:, 0] * k[:, 2],
                          dp[:, 0] * k[:, 1] - dp[:, 1] * k[:, 0]], -1)
        dTw = pxf[:, None] * dk + pyf[:, None] * dl + dL_dz[:, None] * torch.stack([sx, sy, torch.ones_like(sx)], -1)
        contrib = torch.zeros(HW, 20)
        m3 = (ok & use3d)[:, None]
        contrib[:, 0:3] = torch.where(m3, -dk, torch.zeros(()))
        contrib[:, 3:6] = torch.where(m3, -dl, torch.zeros(()))
        contrib[:, 6:9] = torch.where(m3, dTw, torch.zeros(()))
        m2 = ok & ~use3d
        contrib[:, 8] = torch.where(m2, dL_dz, contrib[:, 8])
        contrib[:, 9] = torch.where(m2, dL_dG * (-G * FILTER_INV_SQUARE * q["dx"]), torch.zeros(()))
        contrib[:, 10] = torch.where(m2, dL_dG * (-G * FILTER_INV_SQUARE * q["dy"]), torch.zeros(()))
        contrib[:, 11:14] = g_nrm
        contrib[:, 14] = G * dL_dalpha
        contrib[:, 15:18] = g_col
        contrib[:, 18] = w
        contrib[:, 19] = s1
        contrib = torch.where(ok[:, None], contrib, torch.zeros(()))
        sel = torch.nonzero(ok).reshape(-1)
        acc.index_add_(0, g[sel], contrib[sel])
    return acc


def _quat_vjp(q, vR):
    """auxiliary.h:248-292; vR[P,3(col),3(row)]"""
    s = torch.rsqrt(q[:, 3] * q[:, 3] + q[:, 0] * q[:, 0] + q[:, 1] * q[:, 1] + q[:, 2] * q[:, 2])
    w, x, y, z = q[:, 0] * s, q[:, 1] * s, q[:, 2] * s, q[:, 3] * s
    v = lambda c, r: vR[:, c, r]
    o0 = 2.0 * (x * (v(1, 2) - v(2, 1)) + y * (v(2, 0) - v(0, 2)) + z * (v(0, 1) - v(1, 0)))
    o1 = 2.0 * (-2.0 * x * (v(1, 1) + v(2, 2)) + y * (v(0, 1) + v(1, 0)) + z * (v(0, 2) + v(2, 0)) + w * (v(1, 2) - v(2, 1)))
    o2 = 2.0 * (x * (v(0, 1) + v(1, 0)) - 2.0 * y * (v(0, 0) + v(2, 2)) + z * (v(1, 2) + v(2, 1)) + w * (v(2, 0) - v(0, 2)))
    o3 = 2.0 * (x * (v(0, 2) + v(2, 0)) + y * (v(1, 2) + v(2, 1)) - 2.0 * z * (v(0, 0) + v(1, 1)) + w * (v(0, 1) - v(1, 0)))
    return torch.stack([o0, o1, o2, o3], -1)


def preprocess_backward(pre, acc, means3D, scales, rotations, view, proj, W, H, tanfovx, tanfovy):
    """backward.cu:849-1045.  acc[P,20] is the blend backward's accumulation (layout of SB_GRAD_F)."""
    P = means3D.shape[0]
    f32 = np.float32
    focal_x = f32(W) / (f32(2.0) * f32(tanfovx))
    focal_y = f32(H) / (f32(2.0) * f32(tanfovy))
    Wq = int(f32(focal_x * f32(tanfovx)) * f32(2))
    Hq = int(f32(focal_y * f32(tanfovy)) * f32(2))
    T, L, R, Pm = _transmat(means3D, scales, rotations, proj, Wq, Hq, backward_assoc=True)
    normal, mult = _view_normal(L, view)
    dL_dT = acc[:, 0:9].reshape(P, 3, 3).clone()
    g2 = acc[:, 9:11] + acc[:, 9:11]  # the two identical atomics (Q2)
    nz = (g2[:, 0] != 0) | (g2[:, 1] != 0)
    Tu, Tv, Tw = T[:, 0, :], T[:, 1, :], T[:, 2, :]
    distance = Tw[:, 0] * Tw[:, 0] + Tw[:, 1] * Tw[:, 1] - Tw[:, 2] * Tw[:, 2]
    f = 1 / distance
    row0 = torch.stack([f * Tw[:, 0], f * Tw[:, 1], -f * Tw[:, 2]], -1)
    fac = torch.stack([f - 2 * f * f * Tw[:, 0] * Tw[:, 0], f - 2 * f * f * Tw[:, 1] * Tw[:, 1],
                       -(f + 2 * f * f * Tw[:, 2] * Tw[:, 2])], -1)
    add = torch.zeros(P, 3, 3)
    add[:, 0, :] = g2[:, 0:1] * row0
    add[:, 1, :] = g2[:, 1:2] * row0
    add[:, 2, :] = g2[:, 0:1] * (Tu * fac) + g2[:, 1:2] * (Tv * fac)
    dL_dT = torch.where(nz[:, None, None], dL_dT + add, dL_dT)
    dL_dM = torch.zeros(P, 3, 4)
    for c in range(3):
        for r in range(4):
            dL_dM[:, c, r] = Pm[0, r] * dL_dT[:, 0, c] + Pm[1, r] * dL_dT[:, 1, c] + Pm[2, r] * dL_dT[:, 2, c]
    vm = view.reshape(16)
    dn = acc[:, 11:14]
    dtn = torch.stack([vm[0] * dn[:, 0] + vm[1] * dn[:, 1] + vm[2] * dn[:, 2],
                       vm[4] * dn[:, 0] + vm[5] * dn[:, 1] + vm[6] * dn[:, 2],
                       vm[8] * dn[:, 0] + vm[9] * dn[:, 1] + vm[10] * dn[:, 2]], -1) * mult[:, None]
    dR = torch.stack([dL_dM[:, 0, 0:3] * scales[:, 0:1], dL_dM[:, 1, 0:3] * scales[:, 1:2], dtn], 1)
    vis = (pre["radii"] > 0)
    z = lambda t: torch.where(vis.reshape([-1] + [1] * (t.dim() - 1)), t, torch.zeros(()))
    drot = _quat_vjp(rotations, dR)
    dscale = torch.stack([(dL_dM[:, 0, 0:3] * R[:, 0, :]).sum(-1), (dL_dM[:, 1, 0:3] * R[:, 1, :]).sum(-1)], -1)
    dmean3D = dL_dM[:, 2, 0:3]
    depth = pre["T"][:, 2, 2]
    dmean2D = torch.stack([acc[:, 2] * depth * 0.5 * float(Wq), acc[:, 5] * depth * 0.5 * float(Hq),
                           torch.zeros(P)], -1)
    return dict(dL_dmeans3D=z(dmean3D), dL_dmeans2D=z(dmean2D), dL_dcolors=z(acc[:, 15:18]),
                dL_dopacity=z(acc[:, 14:15]), dL_dscales=z(dscale), dL_drotations=z(drot),
                dL_dtransMat=z(acc[:, 0:9]), scores=z(acc[:, 18:20]))


def rasterize_forward(means3D, colors, opacities, scales, rotations, view, proj, bg, W, H):
    pre = preprocess(means3D, scales, rotations, opacities, view, proj, W, H)
    bins = binning(pre, W, H)
    fwd = blend_forward(pre, bins, colors, opacities, bg, W, H)
    return pre, bins, fwd


def rasterize_backward(pre, bins, fwd, means3D, colors, opacities, scales, rotations, view, proj, bg, W, H, tanfovx,
                       tanfovy, dL_dcolor, dL_dallmap, pixel_mask):
    acc = blend_backward(pre, bins, fwd, colors, opacities, bg, dL_dcolor, dL_dallmap, pixel_mask, W, H)
    return preprocess_backward(pre, acc, means3D, scales, rotations, view, proj, W, H, tanfovx, tanfovy)


def forward_backward(inp, dL_dcolor, dL_dallmap, pixel_mask=None):
    """One fwd+bwd iteration on a dict of CPU tensors (keys as produced by tests/scenes.py)."""
    W, H = inp["W"], inp["H"]
    if pixel_mask is None:
        pixel_mask = torch.ones(W * H, dtype=torch.bool)
    pre, bins, fwd = rasterize_forward(inp["means3D"], inp["colors"], inp["opacities"], inp["scales"],
                                       inp["rotations"], inp["viewmatrix"], inp["projmatrix"], inp["bg"], W, H)
    grads = rasterize_backward(pre, bins, fwd, inp["means3D"], inp["colors"], inp["opacities"], inp["scales"],
                               inp["rotations"], inp["viewmatrix"], inp["projmatrix"], inp["bg"], W, H,
                               inp["tanfovx"], inp["tanfovy"], dL_dcolor, dL_dallmap, pixel_mask)
    return pre, bins, fwd, grads
