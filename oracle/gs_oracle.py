"""ctypes front-end of the CPU oracle (oracle/gs_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product (gsr_b200, diff_gaussian_rasterization) never
imports this module.

`forward()` mirrors CudaRasterizer::Rasterizer::forward
(/root/reference/diff-gaussian-rasterization-npu/cuda_rasterizer/rasterizer_impl.cu:198-341) and
returns every intermediate the parity tests compare; `backward()` mirrors Rasterizer::backward
(rasterizer_impl.cu:345-450).
"""
import ctypes
import os
import subprocess
from ctypes import POINTER, c_float, c_int, c_size_t, c_ubyte, c_uint32, c_uint64, c_void_p

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "gs_oracle.c")
_LIB = os.path.join(_HERE, "libgs_oracle.so")


def build(force=False):
    """gcc -> oracle/libgs_oracle.so.  -ffp-contract=off: every fused multiply-add is an explicit fmaf()."""
    if not force and os.path.exists(_LIB) and os.path.getmtime(_LIB) >= os.path.getmtime(_SRC):
        return _LIB
    flags = ["-O2", "-std=c11", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off", "-fno-fast-math"]
    try:
        with open("/proc/cpuinfo") as f:
            cpu = f.read()
        if " fma" in cpu and " avx2" in cpu:
            flags += ["-mfma", "-mavx2"]          # fmaf() inlines to vfmadd (still one rounding)
    except OSError:
        pass
    tmp = _LIB + ".tmp.%d" % os.getpid()
    subprocess.check_call(["gcc", *flags, _SRC, "-o", tmp, "-lm"])
    os.replace(tmp, _LIB)
    return _LIB


class _Scene(ctypes.Structure):
    _fields_ = [("P", c_int), ("D", c_int), ("M", c_int), ("W", c_int), ("H", c_int),
                ("background", c_void_p), ("means3D", c_void_p), ("shs", c_void_p),
                ("colors_precomp", c_void_p), ("opacities", c_void_p), ("scales", c_void_p),
                ("scale_modifier", c_float), ("rotations", c_void_p), ("cov3D_precomp", c_void_p),
                ("viewmatrix", c_void_p), ("projmatrix", c_void_p), ("cam_pos", c_void_p),
                ("tan_fovx", c_float), ("tan_fovy", c_float), ("prefiltered", c_int),
                ("antialiasing", c_int)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        path = build()
        try:
            _lib = ctypes.CDLL(path)
        except OSError:
            _lib = ctypes.CDLL(build(force=True))
        _lib.gso_higher_msb.restype = c_uint32
        _lib.gso_higher_msb.argtypes = [c_uint32]
        _lib.gso_num_threads.restype = c_int
        _lib.gso_preprocess.restype = c_int
        for name in ("gso_preprocess", "gso_inclusive_scan", "gso_duplicate", "gso_sort_pairs",
                     "gso_tile_ranges", "gso_render_forward", "gso_render_backward",
                     "gso_preprocess_backward", "gso_mark_visible"):
            getattr(_lib, name).argtypes = None
    return _lib


def _p(a):
    return None if a is None else c_void_p(a.ctypes.data)


def _f32(a):
    if a is None:
        return None
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a if a.size else None


def num_threads():
    return lib().gso_num_threads()


def set_num_threads(n):
    lib().gso_set_num_threads(c_int(n))


def higher_msb(n):
    return int(lib().gso_higher_msb(c_uint32(n)))


class Scene:
    """Holds contiguous fp32 copies of the inputs and the C struct that points at them."""

    def __init__(self, *, W, H, means3D, opacities, viewmatrix, projmatrix, cam_pos, tan_fovx,
                 tan_fovy, background, shs=None, colors_precomp=None, scales=None, rotations=None,
                 cov3D_precomp=None, sh_degree=3, scale_modifier=1.0, prefiltered=False,
                 antialiasing=False):
        self.W, self.H = int(W), int(H)
        self.means3D = _f32(means3D).reshape(-1, 3)
        self.P = self.means3D.shape[0]
        self.opacities = _f32(opacities).reshape(-1)
        self.shs = _f32(shs)
        self.M = 0 if self.shs is None else self.shs.reshape(self.P, -1, 3).shape[1]
        self.colors_precomp = _f32(colors_precomp)
        self.scales, self.rotations = _f32(scales), _f32(rotations)
        self.cov3D_precomp = _f32(cov3D_precomp)
        self.viewmatrix = _f32(viewmatrix).reshape(16)
        self.projmatrix = _f32(projmatrix).reshape(16)
        self.cam_pos = _f32(cam_pos).reshape(3)
        self.background = _f32(background).reshape(3)
        self.D = int(sh_degree)
        self.tan_fovx, self.tan_fovy = float(tan_fovx), float(tan_fovy)
        self.antialiasing, self.prefiltered = bool(antialiasing), bool(prefiltered)
        self.scale_modifier = float(scale_modifier)
        assert (self.shs is None) != (self.colors_precomp is None)
        assert (self.scales is None) == (self.rotations is None)
        assert (self.scales is None) != (self.cov3D_precomp is None)
        self.c = _Scene(self.P, self.D, self.M, self.W, self.H, _p(self.background),
                        _p(self.means3D), _p(self.shs), _p(self.colors_precomp), _p(self.opacities),
                        _p(self.scales), self.scale_modifier, _p(self.rotations),
                        _p(self.cov3D_precomp), _p(self.viewmatrix), _p(self.projmatrix),
                        _p(self.cam_pos), self.tan_fovx, self.tan_fovy, int(self.prefiltered),
                        int(self.antialiasing))

    @property
    def grid(self):
        return (self.W + 15) // 16, (self.H + 15) // 16


def sort_pairs(keys, values, end_bit):
    keys = np.ascontiguousarray(keys, dtype=np.uint64)
    values = np.ascontiguousarray(values, dtype=np.uint32)
    ko, vo = np.empty_like(keys), np.empty_like(values)
    lib().gso_sort_pairs(_p(keys), _p(values), _p(ko), _p(vo), c_size_t(keys.size), c_int(end_bit))
    return ko, vo


def mark_visible(means3D, viewmatrix):
    m = _f32(means3D).reshape(-1, 3)
    out = np.zeros(m.shape[0], dtype=np.uint8)
    lib().gso_mark_visible(c_int(m.shape[0]), _p(m), _p(_f32(viewmatrix).reshape(16)), _p(out))
    return out.astype(bool)


def forward(sc, stages="all", borderline=True):
    """Runs the forward pipeline; returns a dict of every stage's output.

    stages: "preprocess" | "binning" (through tile ranges) | "all".
    """
    L = lib()
    P, W, H = sc.P, sc.W, sc.H
    gx, gy = sc.grid
    o = {}
    o["radii"] = np.zeros(P, np.int32)
    o["tiles_touched"] = np.zeros(P, np.uint32)
    o["depths"] = np.zeros(P, np.float32)
    o["means2D"] = np.zeros((P, 2), np.float32)
    o["cov3D"] = np.zeros((P, 6), np.float32)
    o["conic_opacity"] = np.zeros((P, 4), np.float32)
    o["rgb"] = np.zeros((P, 3), np.float32)
    o["clamped"] = np.zeros((P, 3), np.uint8)
    o["trapped"] = L.gso_preprocess(ctypes.byref(sc.c), _p(o["radii"]), _p(o["tiles_touched"]),
                                    _p(o["depths"]), _p(o["means2D"]), _p(o["cov3D"]),
                                    _p(o["conic_opacity"]), _p(o["rgb"]), _p(o["clamped"]))
    if stages == "preprocess":
        return o
    o["point_offsets"] = np.zeros(P, np.uint32)
    L.gso_inclusive_scan(_p(o["tiles_touched"]), _p(o["point_offsets"]), c_int(P))
    N = int(o["point_offsets"][-1]) if P else 0
    o["num_rendered"] = N
    o["keys_unsorted"] = np.zeros(N, np.uint64)
    o["values_unsorted"] = np.zeros(N, np.uint32)
    L.gso_duplicate(c_int(P), c_int(W), c_int(H), _p(o["means2D"]), _p(o["depths"]),
                    _p(o["point_offsets"]), _p(o["radii"]), _p(o["keys_unsorted"]),
                    _p(o["values_unsorted"]))
    o["bit"] = higher_msb(gx * gy)
    o["keys"], o["point_list"] = sort_pairs(o["keys_unsorted"], o["values_unsorted"], 32 + o["bit"])
    o["ranges"] = np.zeros((gx * gy, 2), np.uint32)
    L.gso_tile_ranges(_p(o["keys"]), c_size_t(N), _p(o["ranges"]), c_int(gx * gy))
    if stages == "binning":
        return o
    o["out_color"] = np.zeros((3, H, W), np.float32)
    o["final_T"] = np.zeros((H, W), np.float32)
    o["n_contrib"] = np.zeros((H, W), np.uint32)
    o["invdepth"] = np.zeros((1, H, W), np.float32)
    o["borderline"] = np.zeros((H, W), np.uint8) if borderline else None
    stats = np.zeros(2, np.uint64)
    feats = sc.colors_precomp if sc.colors_precomp is not None else o["rgb"]
    L.gso_render_forward(c_int(W), c_int(H), _p(o["ranges"]), _p(o["point_list"]), _p(o["means2D"]),
                         _p(feats), _p(o["conic_opacity"]), _p(o["depths"]), _p(sc.background),
                         _p(o["out_color"]), _p(o["final_T"]), _p(o["n_contrib"]), _p(o["invdepth"]),
                         _p(o["borderline"]), _p(stats))
    o["pairs_evaluated"], o["pairs_blended"] = int(stats[0]), int(stats[1])
    return o


def backward(sc, fwd, dL_dcolor, dL_dinvdepth=None):
    """Rasterizer::backward on the oracle's own forward state `fwd`."""
    L = lib()
    P, W, H = sc.P, sc.W, sc.H
    dpix = _f32(dL_dcolor).reshape(3, H, W)
    dinv = None if dL_dinvdepth is None else np.ascontiguousarray(dL_dinvdepth, np.float32).reshape(H, W)
    g = {}
    g["dL_dmean2D"] = np.zeros((P, 3), np.float32)
    g["dL_dconic"] = np.zeros((P, 4), np.float32)
    g["dL_dopacity"] = np.zeros((P, 1), np.float32)
    g["dL_dcolors"] = np.zeros((P, 3), np.float32)
    g["dL_dinvdepths"] = np.zeros((P, 1), np.float32) if dinv is not None else None
    feats = sc.colors_precomp if sc.colors_precomp is not None else fwd["rgb"]
    L.gso_render_backward(c_int(P), c_int(W), c_int(H), c_size_t(fwd["num_rendered"]),
                          _p(fwd["ranges"]), _p(fwd["point_list"]), _p(sc.background),
                          _p(fwd["means2D"]), _p(fwd["conic_opacity"]), _p(feats), _p(fwd["depths"]),
                          _p(fwd["final_T"]), _p(fwd["n_contrib"]), _p(dpix), _p(dinv),
                          _p(g["dL_dmean2D"]), _p(g["dL_dconic"]), _p(g["dL_dopacity"]),
                          _p(g["dL_dcolors"]), _p(g["dL_dinvdepths"]))
    g["dL_dopacity_blend"] = g["dL_dopacity"].copy()   # before the antialiasing rescale
    g["dL_dmeans3D"] = np.zeros((P, 3), np.float32)
    g["dL_dcov3D"] = np.zeros((P, 6), np.float32)
    g["dL_dsh"] = np.zeros((P, sc.M, 3), np.float32)
    g["dL_dscales"] = np.zeros((P, 3), np.float32)
    g["dL_drotations"] = np.zeros((P, 4), np.float32)
    cov_used = sc.cov3D_precomp if sc.cov3D_precomp is not None else fwd["cov3D"]
    L.gso_preprocess_backward(ctypes.byref(sc.c), _p(fwd["radii"]), _p(cov_used), _p(fwd["clamped"]),
                              _p(g["dL_dmean2D"]), _p(g["dL_dconic"]), _p(g["dL_dinvdepths"]),
                              _p(g["dL_dopacity"]), _p(g["dL_dcolors"]), _p(g["dL_dmeans3D"]),
                              _p(g["dL_dcov3D"]), _p(g["dL_dsh"]), _p(g["dL_dscales"]),
                              _p(g["dL_drotations"]))
    return g
