"""CPU oracle of the trained-model PLY format (numpy).  TEST INFRASTRUCTURE ONLY (tests/ only).

Restates GaussianModel.load_ply / save_ply (/root/reference/scene/gaussian_model.py:225-314) without `plyfile`
(absent here): properties are looked up by name in the parsed header, f_rest is reshaped (P, 3, M-1) and transposed
to (P, M-1, 3), f_dc becomes (P, 1, 3); save_ply writes x y z nx ny nz f_dc f_rest opacity scale rot with
features transposed back to channel-major.  Activations as gaussian_model.py:47-58 (exp, sigmoid, normalize).
Pinned by the reference's own fixtures GS-IRON/base/single_gaussian.ply / two_gaussians.ply (copied to tests/golden/)
and the values their generators hard-code (GS-IRON/base/gaussian.py:4-67, gaussian2.py).
"""
import numpy as np

_SIZES = {"char": 1, "uchar": 1, "int8": 1, "uint8": 1, "short": 2, "ushort": 2, "int16": 2, "uint16": 2, "int": 4,
          "uint": 4, "float": 4, "int32": 4, "uint32": 4, "float32": 4, "double": 8, "float64": 8}
_NP = {"char": "i1", "uchar": "u1", "int8": "i1", "uint8": "u1", "short": "<i2", "ushort": "<u2", "int16": "<i2",
       "uint16": "<u2", "int": "<i4", "uint": "<u4", "float": "<f4", "int32": "<i4", "uint32": "<u4", "float32": "<f4",
       "double": "<f8", "float64": "<f8"}


def read_vertices(raw):
    """bytes of a binary_little_endian PLY -> numpy structured array of the vertex element."""
    end = raw.index(b"end_header\n") + len(b"end_header\n")
    lines = raw[:end].decode("ascii").splitlines()
    assert lines[0] == "ply" and lines[1].split()[:2] == ["format", "binary_little_endian"]
    n, fields, in_vertex = 0, [], False
    for ln in lines[2:]:
        t = ln.split()
        if not t or t[0] in ("comment", "obj_info", "end_header"):
            continue
        if t[0] == "element":
            in_vertex = t[1] == "vertex"
            if in_vertex:
                n = int(t[2])
        elif t[0] == "property" and in_vertex:
            fields.append((t[2], _NP[t[1]]))
    return np.frombuffer(raw, dtype=np.dtype(fields), count=n, offset=end)


def load_ply(raw, max_sh_degree=3):
    """gaussian_model.py:265-314 -> dict of float32 arrays in the reference's parameter shapes."""
    v = read_vertices(raw)
    P = v.shape[0]
    xyz = np.stack((v["x"], v["y"], v["z"]), axis=1)
    opacities = np.asarray(v["opacity"])[..., None]
    fdc = np.zeros((P, 3, 1))
    for c in range(3):
        fdc[:, c, 0] = v["f_dc_%d" % c]
    names = sorted([n for n in v.dtype.names if n.startswith("f_rest_")], key=lambda x: int(x.split("_")[-1]))
    assert len(names) == 3 * (max_sh_degree + 1) ** 2 - 3
    extra = np.zeros((P, len(names)))
    for i, n in enumerate(names):
        extra[:, i] = v[n]
    extra = extra.reshape((P, 3, (max_sh_degree + 1) ** 2 - 1))
    scales = np.stack([v["scale_%d" % i] for i in range(3)], axis=1)
    rots = np.stack([v["rot_%d" % i] for i in range(4)], axis=1)
    f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    return dict(xyz=f32(xyz), features_dc=f32(fdc.transpose(0, 2, 1)), features_rest=f32(extra.transpose(0, 2, 1)),
                opacity=f32(opacities), scaling=f32(scales), rotation=f32(rots))


def activate(p):
    """gaussian_model.py:47-58,107-124: what render() hands to the rasterizer."""
    rot = p["rotation"].astype(np.float64)
    rot = rot / np.maximum(np.linalg.norm(rot, axis=1, keepdims=True), 1e-12)
    return dict(xyz=p["xyz"], shs=np.concatenate([p["features_dc"], p["features_rest"]], axis=1),
                opacity=(1.0 / (1.0 + np.exp(-p["opacity"].astype(np.float64)))).astype(np.float32),
                scales=np.exp(p["scaling"].astype(np.float64)).astype(np.float32), rotations=rot.astype(np.float32))


def save_ply(p):
    """gaussian_model.py:225-258 -> file bytes."""
    P = p["xyz"].shape[0]
    fdc = p["features_dc"].transpose(0, 2, 1).reshape(P, -1)
    frest = p["features_rest"].transpose(0, 2, 1).reshape(P, -1)
    names = ["x", "y", "z", "nx", "ny", "nz"] + ["f_dc_%d" % i for i in range(fdc.shape[1])] + \
            ["f_rest_%d" % i for i in range(frest.shape[1])] + ["opacity"] + ["scale_%d" % i for i in range(3)] + \
            ["rot_%d" % i for i in range(4)]
    attrs = np.concatenate((p["xyz"], np.zeros_like(p["xyz"]), fdc, frest, p["opacity"], p["scaling"], p["rotation"]), axis=1)
    head = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % P + "".join("property float %s\n" % n for n in names) + "end_header\n"
    return head.encode("ascii") + np.ascontiguousarray(attrs, dtype="<f4").tobytes()
