"""Trained-model PLY I/O on the GPU (include/gsr_ply.h, csrc/ply_io.cu).

Host mirror of GaussianModel.load_ply / save_ply (/root/reference/scene/gaussian_model.py:240-314) without the
`plyfile` package: the header is parsed by libgsr_b200.so (attributes are located by name), the raw record bytes are
uploaded as they lie in the file and one kernel transposes them into the tensors the rasterizer takes.
`load_ply` returns the reference's six parameter tensors (raw, optimiser-space values, the reference's shapes);
`load_ply_for_render` returns activated rasterizer inputs (exp / sigmoid / normalize fused into the same kernel).
No CPU fallback: the transposition runs on the device or raises.
"""
import ctypes
import os
from ctypes import POINTER, c_char_p, c_int, c_int32, c_size_t, c_uint32, c_uint64, c_void_p

import numpy as np
import torch

from . import capi

NUM_ATTR = 62


class PlyLayout(ctypes.Structure):
    """gsr_ply_layout"""
    _fields_ = [("num_vertices", c_uint64), ("header_bytes", c_uint64), ("record_bytes", c_uint32), ("n_rest", c_int32),
                ("offset", c_int32 * NUM_ATTR)]


_bound = False


def _lib():
    global _bound
    L = capi.lib()
    if not _bound:
        L.gsr_ply_parse_header.restype = c_int
        L.gsr_ply_parse_header.argtypes = [c_char_p, c_size_t, POINTER(PlyLayout)]
        L.gsr_ply_write_header.restype = c_size_t
        L.gsr_ply_write_header.argtypes = [c_uint64, c_int, c_char_p, c_size_t]
        L.gsr_ply_unpack.restype = c_int
        L.gsr_ply_unpack.argtypes = [POINTER(PlyLayout), c_void_p, c_int, c_int] + [c_void_p] * 6
        L.gsr_ply_pack.restype = c_int
        L.gsr_ply_pack.argtypes = [c_uint64, c_int] + [c_void_p] * 7
        _bound = True
    return L


def parse_header(head_bytes):
    """gsr_ply_parse_header on the first bytes of a file (host only)."""
    lay = PlyLayout()
    capi._check(_lib().gsr_ply_parse_header(head_bytes, len(head_bytes), ctypes.byref(lay)))
    return lay


def write_header(num_vertices, sh_degree=3):
    n = _lib().gsr_ply_write_header(num_vertices, sh_degree, None, 0)
    buf = ctypes.create_string_buffer(n)
    _lib().gsr_ply_write_header(num_vertices, sh_degree, buf, n)
    return buf.raw[:n]


def _read(path):
    with open(path, "rb") as f:
        head = f.read(1 << 16)
    lay = parse_header(head)
    P = lay.num_vertices
    payload = np.fromfile(path, dtype=np.uint8, offset=lay.header_bytes, count=P * lay.record_bytes)
    if payload.size != P * lay.record_bytes:
        raise RuntimeError("ply: file is shorter than its header says (%d of %d payload bytes)" % (payload.size, P * lay.record_bytes))
    return lay, payload


def unpack(lay, records_dev, sh_degree=3, activate=False):
    """gsr_ply_unpack: device payload -> dict of device tensors (xyz, shs [P,M,3], opacity [P,1], scales, rotations)."""
    dev, P, M = records_dev.device, int(lay.num_vertices), (sh_degree + 1) ** 2
    e = lambda *s: torch.empty(*s, dtype=torch.float32, device=dev)
    out = dict(xyz=e(P, 3), shs=e(P, M, 3), opacity=e(P, 1), scales=e(P, 3), rotations=e(P, 4))
    with torch.cuda.device(dev):
        capi._check(_lib().gsr_ply_unpack(ctypes.byref(lay), c_void_p(records_dev.data_ptr()) if P else None, sh_degree,
                                          int(activate), *[c_void_p(out[k].data_ptr()) if P else None
                                                           for k in ("xyz", "shs", "opacity", "scales", "rotations")],
                                          capi._stream()))
    return out


def load_ply(path, device="cuda", max_sh_degree=3):
    """GaussianModel.load_ply (gaussian_model.py:265-314): returns the six parameter tensors in the reference's shapes:
    _xyz [P,3], _features_dc [P,1,3], _features_rest [P,M-1,3], _opacity [P,1], _scaling [P,3], _rotation [P,4]."""
    lay, payload = _read(path)
    if lay.n_rest != 3 * (max_sh_degree + 1) ** 2 - 3:            # the reference's assert, gaussian_model.py:290
        raise AssertionError("f_rest count %d does not match max_sh_degree %d" % (lay.n_rest, max_sh_degree))
    t = unpack(lay, torch.from_numpy(payload).to(device), max_sh_degree, activate=False)
    return dict(xyz=t["xyz"], features_dc=t["shs"][:, :1].contiguous(), features_rest=t["shs"][:, 1:].contiguous(),
                opacity=t["opacity"], scaling=t["scales"], rotation=t["rotations"], active_sh_degree=max_sh_degree)


def load_ply_for_render(path, device="cuda", sh_degree=3):
    """Rasterizer inputs straight from the file: activated scales / opacities / rotations, shs [P,M,3]."""
    lay, payload = _read(path)
    return unpack(lay, torch.from_numpy(payload).to(device), sh_degree, activate=True)


def save_ply(path, xyz, shs, opacity, scaling, rotation):
    """GaussianModel.save_ply (gaussian_model.py:240-258) from raw parameter tensors on the device; shs = cat(features_dc,
    features_rest) [P,M,3]."""
    P, M = xyz.shape[0], shs.shape[1]
    sh_degree = int(round(M ** 0.5)) - 1
    R = 17 + 3 * (M - 1)
    rec = torch.empty(P * R, dtype=torch.float32, device=xyz.device)
    c = lambda t: c_void_p(t.contiguous().data_ptr()) if P else None
    keep = [t.contiguous() for t in (xyz, shs, opacity, scaling, rotation)]
    with torch.cuda.device(xyz.device):
        capi._check(_lib().gsr_ply_pack(P, sh_degree, *[c_void_p(t.data_ptr()) if P else None for t in keep],
                                        c_void_p(rec.data_ptr()) if P else None, capi._stream()))
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "wb") as f:
        f.write(write_header(P, sh_degree))
        f.write(rec.cpu().numpy().tobytes())
