"""On-disk interchange with the reference: 2DGS/3DGS PLY files and ``.pt`` checkpoints.

Mirrors scripts/gaussian/vis_utils.py:186-311 (``construct_list_of_attributes``, ``save_ply``, ``load_ply``) and
scripts/gaussian/gaussian_base.py:477-525 (``load_ply_ckpt``, ``save_pt_ckpt``, ``load_pt_ckpt``).  The reference
writes PLY through the ``plyfile`` package (``PlyData([PlyElement.describe(elements, 'vertex')]).write(path)``): a
binary little-endian PLY whose header lists one ``property float <name>`` per attribute in the order of
``construct_list_of_attributes`` and whose body is the packed float32 records.  ``plyfile`` is not installed in this
image, so the byte layout is written and parsed here directly from that description (parity with the reference's
files is therefore pinned by the PLY format, not by running plyfile: see DESIGN.md).

Host-side code: plain numpy / torch, no GPU needed.
"""
import numpy as np
import torch

C0 = 0.28209479177387814  # SH degree-0 basis constant (vis_utils.py:215)
GROUPS = ("_xyz", "_rgb", "_opacity", "_scaling", "_rotation")
BOOK = ("_global_scores", "_local_scores", "_stable_mask", "_globalkf_id", "_globalkf_max_scores")


def construct_list_of_attributes(save_mode):
    """vis_utils.py:186-205"""
    names = ["x", "y", "z", "nx", "ny", "nz"]
    names += ["f_dc_%d" % i for i in range(3)]
    names += ["f_rest_%d" % i for i in range(45)]
    names.append("opacity")
    if save_mode in ("3dgs", "sky"):
        names += ["scale_%d" % i for i in range(3)]
    elif save_mode == "2dgs":
        names += ["scale_%d" % i for i in range(2)]
    else:
        raise AssertionError("Invalid save_mode.")
    names += ["rot_%d" % i for i in range(4)]
    return names


def _np(t):
    return t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)


def save_ply(props, path, save_mode="2dgs"):
    """props: {'_xyz' [P,3], '_rgb' [P,3] (0..1), '_opacity' [P,1] (logit), '_scaling' [P,2] (log), '_rotation' [P,4]}
    -- the RAW parameters, exactly what the reference stores (vis_utils.py:207-244).  save_mode '3dgs' / 'sky' pad a
    third log-scale of -10 so that 3DGS viewers show flat splats."""
    names = construct_list_of_attributes(save_mode)
    xyz = _np(props["_xyz"]).astype(np.float32)
    P = xyz.shape[0]
    f_dc = ((_np(props["_rgb"]).astype(np.float32) - np.float32(0.5)) / np.float32(C0)).astype(np.float32)
    scale = _np(props["_scaling"]).astype(np.float32)
    if save_mode in ("3dgs", "sky"):
        scale = np.concatenate((scale, -10 * np.ones((P, 1), dtype=np.float32)), axis=1)
    attributes = np.concatenate((xyz, np.zeros_like(xyz), f_dc, np.zeros((P, 45), dtype=np.float32),
                                 _np(props["_opacity"]).astype(np.float32).reshape(P, 1), scale,
                                 _np(props["_rotation"]).astype(np.float32)), axis=1)
    assert attributes.shape[1] == len(names)
    header = ["ply", "format binary_little_endian 1.0", "element vertex %d" % P]
    header += ["property float %s" % n for n in names]
    header.append("end_header")
    with open(path, "wb") as f:
        f.write(("\n".join(header) + "\n").encode("ascii"))
        f.write(np.ascontiguousarray(attributes, dtype="<f4").tobytes())


_PLY_TYPES = {"float": "f4", "float32": "f4", "double": "f8", "float64": "f8", "uchar": "u1", "uint8": "u1",
              "char": "i1", "int8": "i1", "short": "i2", "int16": "i2", "ushort": "u2", "uint16": "u2",
              "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4"}


def read_ply_vertices(path):
    """Structured numpy array of the first element of a PLY file (binary little/big endian or ascii)."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError("%s is not a PLY file" % path)
        fmt, count, props, in_first = None, None, [], False
        while True:
            line = f.readline()
            if not line:
                raise ValueError("PLY header of %s is truncated" % path)
            tok = line.decode("ascii").split()
            if not tok or tok[0] == "comment":
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                if count is None:
                    count, in_first = int(tok[2]), True
                else:
                    in_first = False
            elif tok[0] == "property" and in_first:
                if tok[1] == "list":
                    raise ValueError("list properties are not supported")
                props.append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if fmt == "ascii":
            data = np.loadtxt(f, max_rows=count, ndmin=2)
            out = np.empty(count, dtype=[(n, t) for n, t in props])
            for i, (n, _) in enumerate(props):
                out[n] = data[:, i]
            return out
        order = "<" if fmt == "binary_little_endian" else ">"
        dt = np.dtype([(n, order + t) for n, t in props])
        return np.frombuffer(f.read(count * dt.itemsize), dtype=dt, count=count)


def load_ply(path):
    """vis_utils.py:254-306: raw-parameter dict of numpy arrays ('_rgb' back from the SH DC term, '_scaling' cut to
    its first two columns whatever the file's save_mode was)."""
    v = read_ply_vertices(path)
    names = v.dtype.names
    xyz = np.stack((np.asarray(v["x"]), np.asarray(v["y"]), np.asarray(v["z"])), axis=1)
    opacities = np.asarray(v["opacity"])[..., np.newaxis]
    features_dc = np.zeros((xyz.shape[0], 3, 1))
    for c in range(3):
        features_dc[:, c, 0] = np.asarray(v["f_dc_%d" % c])
    rgb = (features_dc * C0 + 0.5).reshape((-1, 3))
    scale_names = sorted([n for n in names if n.startswith("scale_")], key=lambda x: int(x.split("_")[-1]))
    scales = np.zeros((xyz.shape[0], len(scale_names)))
    for i, n in enumerate(scale_names):
        scales[:, i] = np.asarray(v[n])
    rot_names = sorted([n for n in names if n.startswith("rot")], key=lambda x: int(x.split("_")[-1]))
    rots = np.zeros((xyz.shape[0], len(rot_names)))
    for i, n in enumerate(rot_names):
        rots[:, i] = np.asarray(v[n])
    return {"_xyz": xyz, "_rgb": rgb, "_opacity": opacities, "_scaling": scales[:, :2], "_rotation": rots}


def ply_to_scene(path, device="cpu"):
    """load_ply_ckpt (gaussian_base.py:477-485): float32 tensors of the raw parameters."""
    d = load_ply(path)
    return {k: torch.tensor(d[k], dtype=torch.float32, device=device).contiguous() for k in GROUPS}


def save_pt_ckpt(mapper, path):
    """gaussian_base.py:492-505: the five raw parameters and the score bookkeeping; like the reference, no optimizer
    state (load_pt_ckpt calls setup_optimizer(), i.e. Adam restarts from zero moments)."""
    # exact-size copies: in fused mode the attributes are [:P] views into the surfel store's capacity-sized buffers, and
    # torch.save serialises a view's WHOLE storage (stale rows of pruned surfels included)
    torch.save({k: getattr(mapper, k).detach().clone().contiguous() for k in GROUPS + BOOK}, path)


def load_pt_ckpt(path, device=None):
    """gaussian_base.py:507-525: the dict as saved (tensors moved to ``device`` if given)."""
    d = torch.load(path, map_location=device, weights_only=False)
    missing = [k for k in GROUPS + BOOK if k not in d]
    if missing:
        raise KeyError("checkpoint %s lacks %s" % (path, missing))
    return d
