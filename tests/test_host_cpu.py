"""CPU: the drop-in boundary without a GPU -- the C-ABI library loads and exports every symbol of
include/surfel_b200.h, workspace size/layout queries, the Python surface mirrors the reference module, the
product path refuses to run without CUDA (no fallback), and the synthetic generators are deterministic."""
import ctypes
import inspect
import os
import re
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vings-mono_b200"), os.path.join(ROOT, "tests")]
import diff_surfel_rasterization as dsr  # noqa: E402
import synthetic  # noqa: E402

HEADER = os.path.join(ROOT, "include", "surfel_b200.h")


def test_library_exports_every_declared_symbol():
    lib = dsr._load()
    names = set(re.findall(r"\b(sb200_\w+)\s*\(", open(HEADER).read()))
    assert len(names) >= 15
    for n in sorted(names):
        assert hasattr(lib, n), n
    assert lib.sb200_abi_version() == 7
    assert lib.sb200_last_error() == b""


def test_workspace_queries_need_no_gpu():
    lib = dsr._load()
    assert lib.sb200_geom_bytes(0) >= 0
    prev = 0
    for P in (1, 255, 256, 257, 10_000, 1_200_000, 5_000_000):
        b = lib.sb200_geom_bytes(P)
        assert b >= prev and b >= 128 * P + 32 * P   # 128-B record + depth/rect/tiles + 2 x (key, id) per surfel
        prev = b
        off = (ctypes.c_size_t * 5)()
        lib.sb200_geom_layout(P, off)
        assert off[0] == 0 and all(o % 256 == 0 for o in off) and list(off) == sorted(off)
    # 5 M surfels / 100 M pairs at 1080p stay far below 180 GB (the reference's SampleState alone would not)
    assert lib.sb200_geom_bytes(5_000_000) < 1 << 30
    assert lib.sb200_binning_bytes(100_000_000) < 4 << 30
    assert lib.sb200_grad_bytes(1000) >= 80 * 1000
    io = (ctypes.c_size_t * 3)()
    lib.sb200_image_layout(1241, 376, io)
    assert io[0] == 0 and io[1] >= 8 * 78 * 24 and io[2] >= io[1] + 12 * 1241 * 376
    assert lib.sb200_image_bytes(1241, 376) >= io[2] + 8 * 1241 * 376
    bo = (ctypes.c_size_t * 4)()
    lib.sb200_binning_layout(1000, bo)
    assert bo[1] - bo[0] >= 4000 and bo[3] - bo[2] >= 4000


def test_bad_arguments_return_errors_not_exceptions():
    lib = dsr._load()
    vp = ctypes.c_void_p
    cnt = ctypes.c_int64(7)
    rc = lib.sb200_forward_begin(-1, 64, 64, vp(), vp(), vp(), vp(), vp(), vp(), vp(), vp(), vp(),
                                 ctypes.cast(ctypes.pointer(cnt), vp), vp())
    assert rc != 0 and b"bad dimensions" in lib.sb200_last_error()
    rc = lib.sb200_forward_begin(10, 64, 64, vp(), vp(), vp(), vp(), vp(), vp(), vp(), vp(), vp(),
                                 ctypes.cast(ctypes.pointer(cnt), vp), vp())
    assert rc != 0 and b"NULL" in lib.sb200_last_error()
    # P == 0 is legal and touches nothing (reference: rasterize_points.cu:106)
    rc = lib.sb200_forward_begin(0, 64, 64, vp(), vp(), vp(), vp(), vp(), vp(), vp(), vp(), vp(),
                                 ctypes.cast(ctypes.pointer(cnt), vp), vp())
    assert rc == 0 and cnt.value == 0 and lib.sb200_last_error() == b""
    rc = lib.sb200_forward_begin(10, 16 * 70000, 64, vp(), vp(), vp(), vp(), vp(), vp(), vp(), vp(), vp(),
                                 ctypes.cast(ctypes.pointer(cnt), vp), vp())
    assert rc != 0 and b"65535" in lib.sb200_last_error()


def test_python_surface_mirrors_the_reference_module():
    """diff_surfel_rasterization/__init__.py:171-184 (settings), :202 (forward signature), :241-268 (optimizer)."""
    assert dsr.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug", "pixel_mask")
    assert not dsr.GaussianRasterizationSettings._field_defaults      # 13 required fields, no defaults
    sig = inspect.signature(dsr.GaussianRasterizer.forward)
    assert list(sig.parameters) == ["self", "means3D", "means2D", "opacities", "shs", "colors_precomp", "scales",
                                    "rotations", "cov3D_precomp", "scores"]
    assert issubclass(dsr.GaussianRasterizer, torch.nn.Module) and hasattr(dsr.GaussianRasterizer, "markVisible")
    assert issubclass(dsr.SparseGaussianAdam, torch.optim.Adam)
    assert list(inspect.signature(dsr.SparseGaussianAdam.step).parameters) == ["self", "visibility", "N"]
    ref_init = "/root/reference/submodules/diff-surfel-rasterization/diff_surfel_rasterization/__init__.py"
    if os.path.exists(ref_init):   # in the build container: compare with the reference source itself
        src = open(ref_init).read()
        fields = re.findall(r"^\s+(\w+)\s*:\s*[\w.]+\s*$", src[src.index("class GaussianRasterizationSettings"):
                                                              src.index("class GaussianRasterizer(")], re.M)
        assert tuple(fields) == dsr.GaussianRasterizationSettings._fields


def test_no_cpu_fallback():
    scene, views = synthetic.make_scene("C1", P=50, W=32, H=32)
    inp = synthetic.view_inputs(scene, views[0])
    st = dsr.GaussianRasterizationSettings(
        image_height=32, image_width=32, tanfovx=inp["tanfovx"], tanfovy=inp["tanfovy"], bg=inp["bg"],
        scale_modifier=1.0, viewmatrix=inp["viewmatrix"], projmatrix=inp["projmatrix"], sh_degree=0,
        campos=inp["campos"], prefiltered=False, debug=False, pixel_mask=torch.ones(1024, dtype=torch.bool))
    rast = dsr.GaussianRasterizer(st)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        rast(means3D=inp["means3D"], means2D=None, opacities=inp["opacities"], colors_precomp=inp["colors"],
             scales=inp["scales"], rotations=inp["rotations"])
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        rast(means3D=inp["means3D"], means2D=None, opacities=inp["opacities"], scales=inp["scales"],
             rotations=inp["rotations"])
    # the product never imports the oracle
    for root, _, files in os.walk(os.path.join(ROOT, "vings-mono_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                txt = open(os.path.join(root, f)).read()
                assert "surfel_emulation" not in txt and "import oracle" not in txt, f


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(dsr, "_lib", None)
    monkeypatch.setattr(dsr, "_LIB_PATH", "/nonexistent/libsurfel_b200.so")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        dsr._load()


def test_synthetic_generators_are_deterministic_and_shaped():
    a, va = synthetic.make_scene("C1")
    b, vb = synthetic.make_scene("C1")
    for k in ("means3D", "scales", "rotations", "opacities", "colors"):
        assert torch.equal(a[k], b[k])
    assert a["means3D"].shape == (10_000, 3) and a["scales"].shape == (10_000, 2) and len(va) == 1
    assert torch.equal(va[0].viewmatrix, vb[0].viewmatrix)
    assert torch.allclose(a["rotations"].norm(dim=1), torch.ones(10_000), atol=1e-5)
    for name, P, K, W, H in (("C2", 1_200_000, 8, 1241, 376), ("C3", 500_000, 8, 640, 480),
                             ("C4", 5_000_000, 16, 1920, 1080), ("C5", 3_000_000, 5, 1920, 1280)):
        c = synthetic.CONFIGS[name]
        assert (c["P"], c["K"], c["W"], c["H"]) == (P, K, W, H)
    s, v = synthetic.make_scene("C2", P=80_000)
    assert len(v) == 8 and s["W"] == 1241
    gt = synthetic.make_gt(s, v[0], 0)
    assert gt["rgb"].shape == (3, 376, 1241) and (gt["rgb"].sum(0) == 0).any() and (gt["depth"] > 0).all()
    # get_camera convention (scripts/gaussian/cameras.py:58-60): projmatrix = viewmatrix @ P^T, campos = c2w[:3,3]
    assert torch.allclose(v[0].campos, v[0].c2w[:3, 3].float(), atol=1e-5)


def _build_module():
    import importlib.util
    spec = importlib.util.spec_from_file_location("sb200_build_t", os.path.join(ROOT, "vings-mono_b200", "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_csrc_hash_ignores_comments_and_sees_code(tmp_path, monkeypatch):
    """The stamp of profiles/traffic.json (bench.py: roofline.traffic_stale): a comment or whitespace edit keeps an ncu
    capture valid, any change of code invalidates it."""
    import shutil
    b = _build_module()
    here = tmp_path / "pkg"
    shutil.copytree(os.path.join(ROOT, "vings-mono_b200", "csrc"), here / "csrc",
                    ignore=shutil.ignore_patterns("build*"))
    os.makedirs(tmp_path / "include")
    shutil.copy(os.path.join(ROOT, "include", "surfel_b200.h"), tmp_path / "include" / "surfel_b200.h")
    monkeypatch.setattr(b, "HERE", str(here))
    monkeypatch.setattr(b, "CSRC", str(here / "csrc"))
    h0 = b.csrc_hash()
    f = here / "csrc" / "ring.cuh"
    src = f.read_text()
    f.write_text("// a new comment\n" + src.replace("#pragma once", "#pragma once   /* spaced */\n"))
    assert b.csrc_hash() == h0
    f.write_text(src.replace("#define BB_RING_ARRIVALS 32", "#define BB_RING_ARRIVALS 31"))
    assert b.csrc_hash() != h0


def test_forward_walk_model_runs_and_finds_no_unsafe_cull(capsys, monkeypatch):
    """tools/forward_walk_model.py (the CPU model behind DESIGN.md section 5) on three tiles of C1."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("fwm", os.path.join(ROOT, "tools", "forward_walk_model.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    monkeypatch.setattr(sys, "argv", ["forward_walk_model.py", "C1", "3"])
    mod.main()
    out = capsys.readouterr().out
    assert "non-conservative culls: 0" in out and "model warp instructions per pair" in out
