"""Row N4 of SURVEY.md section 8f: PLY / .pt interchange (vings-mono_b200/formats.py), CPU only.  The reference writes
PLY through plyfile (not installed here), so the checks are against the PLY format itself: header text, record
layout, and the reference's load_ply semantics restated (vis_utils.py:254-306)."""
import os

import numpy as np
import pytest
import torch

import emu_module
import formats
import mapper_step
import synthetic


def _props(P=257, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return {"_xyz": torch.randn(P, 3, generator=g), "_rgb": torch.rand(P, 3, generator=g),
            "_opacity": torch.randn(P, 1, generator=g), "_scaling": torch.randn(P, 2, generator=g) - 3,
            "_rotation": torch.randn(P, 4, generator=g)}


@pytest.mark.parametrize("mode,nscale", [("2dgs", 2), ("3dgs", 3), ("sky", 3)])
def test_ply_layout_and_round_trip(tmp_path, mode, nscale):
    p = _props()
    path = str(tmp_path / ("a_%s.ply" % mode))
    formats.save_ply(p, path, save_mode=mode)
    raw = open(path, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    lines = head.decode("ascii").splitlines()
    names = formats.construct_list_of_attributes(mode)
    assert lines[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 257"]
    assert lines[3:] == ["property float %s" % n for n in names]
    assert len(names) == 3 + 3 + 3 + 45 + 1 + nscale + 4
    rec = np.frombuffer(body, dtype="<f4").reshape(257, len(names))
    assert np.array_equal(rec[:, 0:3], p["_xyz"].numpy()) and not rec[:, 3:6].any() and not rec[:, 9:54].any()
    assert np.allclose(rec[:, 6:9], (p["_rgb"].numpy() - 0.5) / formats.C0, rtol=1e-6)
    assert np.array_equal(rec[:, 54], p["_opacity"].numpy()[:, 0])
    if nscale == 3:
        assert np.all(rec[:, 57] == -10.0)
    d = formats.load_ply(path)
    assert np.array_equal(d["_xyz"], p["_xyz"].numpy()) and np.array_equal(d["_opacity"], p["_opacity"].numpy())
    assert d["_scaling"].shape == (257, 2) and np.array_equal(d["_scaling"], p["_scaling"].numpy())
    assert np.array_equal(d["_rotation"], p["_rotation"].numpy())
    assert np.allclose(d["_rgb"], p["_rgb"].numpy(), atol=1e-6)
    t = formats.ply_to_scene(path)
    assert all(t[k].dtype == torch.float32 for k in t) and t["_opacity"].shape == (257, 1)


def test_ply_reader_accepts_ascii_and_foreign_property_order(tmp_path):
    path = str(tmp_path / "b.ply")
    names = ["rot_3", "rot_2", "rot_1", "rot_0", "scale_1", "scale_0", "opacity", "f_dc_2", "f_dc_1", "f_dc_0", "z", "y", "x"]
    rows = np.arange(2 * len(names), dtype=np.float64).reshape(2, len(names))
    with open(path, "w") as f:
        f.write("ply\nformat ascii 1.0\ncomment made by hand\nelement vertex 2\n")
        f.write("".join("property float %s\n" % n for n in names))
        f.write("element face 0\nproperty list uchar int vertex_indices\nend_header\n")
        for r in rows:
            f.write(" ".join(repr(float(x)) for x in r) + "\n")
    d = formats.load_ply(path)
    assert np.array_equal(d["_xyz"], rows[:, [12, 11, 10]])
    assert np.array_equal(d["_rotation"], rows[:, [3, 2, 1, 0]]) and np.array_equal(d["_scaling"], rows[:, [5, 4]])


def test_empty_ply(tmp_path):
    path = str(tmp_path / "e.ply")
    formats.save_ply(_props(P=0), path)
    assert formats.load_ply(path)["_xyz"].shape == (0, 3)


def test_pt_checkpoint_round_trip_restores_mapper(tmp_path):
    scene, _ = synthetic.make_scene("C1", P=300, W=64, H=48)
    dev = torch.device("cpu")
    m = mapper_step.SurfelMapper(emu_module, scene, dev, fused=False)
    m._global_scores += 1.5
    m._stable_mask[::3] = True
    m._globalkf_id += 7
    path = str(tmp_path / "ckpt.pt")
    formats.save_pt_ckpt(m, path)
    d = formats.load_pt_ckpt(path)
    assert set(d) == set(formats.GROUPS + formats.BOOK)
    m2 = mapper_step.SurfelMapper(emu_module, scene, dev, fused=False)
    m2.set_raw(d, book=d)
    for k in formats.GROUPS + formats.BOOK:
        assert torch.equal(getattr(m2, k).detach(), getattr(m, k).detach()), k
    assert all(len(m2.optimizer.state[p]) == 0 for p in m2.params())  # fresh optimizer, like load_pt_ckpt
    # and through a PLY file (raw parameters only)
    ply = str(tmp_path / "m.ply")
    formats.save_ply({k: getattr(m, k) for k in formats.GROUPS}, ply)
    m3 = mapper_step.SurfelMapper(emu_module, scene, dev, fused=False)
    m3.set_raw(formats.ply_to_scene(ply))
    assert torch.equal(m3._xyz.detach(), m._xyz.detach()) and torch.equal(m3._scaling.detach(), m._scaling.detach())
    assert torch.allclose(m3._rgb.detach(), m._rgb.detach(), atol=1e-6)
    with pytest.raises(KeyError):
        torch.save({"_xyz": m._xyz}, path)
        formats.load_pt_ckpt(path)


def test_pt_ckpt_holds_exactly_P_rows(tmp_path):
    """In fused mode the mapper's attributes are [:P] views into capacity-sized store buffers; the checkpoint must
    hold P rows, not the capacity (torch.save serialises a view's whole storage)."""
    class FakeMapper:
        pass
    m = FakeMapper()
    cap, P = 200_000, 1_000
    cols = dict(_xyz=3, _rgb=3, _opacity=1, _scaling=2, _rotation=4, _local_scores=2, _global_scores=2)
    for k in formats.GROUPS + formats.BOOK:
        if k == "_stable_mask":
            big = torch.zeros(cap, dtype=torch.bool)
        elif k == "_globalkf_id":
            big = torch.arange(cap, dtype=torch.long)
        elif k == "_globalkf_max_scores":
            big = torch.rand(cap)
        else:
            big = torch.randn(cap, cols[k])
        setattr(m, k, big[:P])
    path = str(tmp_path / "views.pt")
    formats.save_pt_ckpt(m, path)
    assert os.path.getsize(path) < 4 * P * (15 + 2 + 2 + 1 + 2 + 1) * 4   # ~P rows of every array, not `cap`
    d = formats.load_pt_ckpt(path)
    for k in formats.GROUPS + formats.BOOK:
        assert torch.equal(d[k], getattr(m, k)), k
        assert d[k].untyped_storage().nbytes() == d[k].numel() * d[k].element_size(), k
