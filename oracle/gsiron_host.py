"""ctypes front-end of oracle/gsiron_host.c: the fork's own forward-only host renderer
(/root/reference/GS-IRON/npu-1/base.cpp:104-447, CPU branch), restated without Eigen / OpenCV.

TEST / BASELINE INFRASTRUCTURE ONLY: imported by tests/ and bench.py's cpu_baseline leg.
"""
import ctypes
import math
import os
import subprocess
from ctypes import c_float, c_int, c_long, c_void_p

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "gsiron_host.c")
_LIB = os.path.join(_HERE, "libgsiron_host.so")
_lib = None

# load_camera (camera_loader.cpp:17-22, :43) and main (base.cpp:453-457)
FX = FY = 1111.1
FOV = 0.69
BASE_MAT = np.array([[-0.9250140190124512, -0.2748899757862091, 0.2622683644294739, -1.0572376251220703],
                     [-0.37993317842483526, 0.6692678928375244, -0.6385383605957031, 2.5740303993225098],
                     [-0.0, -0.6903012990951539, -0.7235219478607177, 2.9166102409362793],
                     [0.0, 0.0, 0.0, 1.0]], np.float32)


def build(force=False):
    if not force and os.path.exists(_LIB) and os.path.getmtime(_LIB) >= os.path.getmtime(_SRC):
        return _LIB
    tmp = _LIB + ".tmp.%d" % os.getpid()
    subprocess.check_call(["gcc", "-O2", "-std=c11", "-fPIC", "-shared", "-fno-fast-math", _SRC, "-o", tmp, "-lm"])
    os.replace(tmp, _LIB)
    return _LIB


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.gsiron_render.restype = c_long
        _lib.gsiron_render.argtypes = [c_int] + [c_void_p] * 7 + [c_int, c_int, c_float, c_float, c_float, c_void_p]
    return _lib


def render(scene, width=800, height=800, base_mat=None, fx=FX, fy=FY, fov=FOV):
    """scene: the rasterizer's activated inputs (means3D, scales, rotations (w,x,y,z), opacities, shs [P,16,3]).
    Returns (image [H,W,3] RGB float32 in [0,1], {"instances": number of (Gaussian, tile) entries})."""
    f = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    xyz, scale, rot = f(scene["means3D"]), f(scene["scales"]), f(scene["rotations"])
    opacity = f(np.asarray(scene["opacities"]).reshape(-1))
    shs = f(scene["shs"])
    P = xyz.shape[0]
    f_dc = f(shs[:, 0, :])                                         # [P,3]
    f_rest = f(shs[:, 1:, :].transpose(0, 2, 1).reshape(P, 45))    # channel-major, the PLY's order (loader.cpp)
    mat = f(BASE_MAT if base_mat is None else base_mat)
    img = np.zeros((height, width, 3), np.float32)
    p = lambda a: c_void_p(a.ctypes.data)
    n = lib().gsiron_render(P, p(xyz), p(scale), p(rot), p(opacity), p(f_dc), p(f_rest), p(mat), width, height,
                            c_float(fx), c_float(fy), c_float(fov), p(img))
    if n < 0:
        raise RuntimeError("gsiron_render failed")
    return img, {"instances": int(n)}
