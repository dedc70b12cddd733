"""ctypes front-end of the CPU oracle for the loss stage (oracle/ssim_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg.
The product (gsr_b200, fused_ssim, diff_gaussian_rasterization) never imports this module.

`fusedssim` / `fusedssim_backward` mirror the two extension functions of the reference
(/root/reference/submodules/fused-ssim/ssim.cu:373-444); `fused_ssim` mirrors the Python entry
(/root/reference/submodules/fused-ssim/fused_ssim/__init__.py:33-41) including the "valid" crop;
`l1_ssim_loss` is the loss of /root/reference/train.py:118-126.
"""
import ctypes
import os
import subprocess
from ctypes import c_double, c_float, c_int, c_void_p

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "ssim_oracle.c")
_LIB = os.path.join(_HERE, "libssim_oracle.so")

C1 = 0.01 ** 2
C2 = 0.03 ** 2


def build(force=False):
    if not force and os.path.exists(_LIB) and os.path.getmtime(_LIB) >= os.path.getmtime(_SRC):
        return _LIB
    flags = ["-O2", "-std=c11", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off", "-fno-fast-math"]
    try:
        with open("/proc/cpuinfo") as f:
            cpu = f.read()
        if " fma" in cpu and " avx2" in cpu:
            flags += ["-mfma", "-mavx2"]
    except OSError:
        pass
    tmp = _LIB + ".tmp.%d" % os.getpid()
    subprocess.check_call(["gcc", *flags, _SRC, "-o", tmp, "-lm"])
    os.replace(tmp, _LIB)
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        try:
            _lib = ctypes.CDLL(build())
        except OSError:
            _lib = ctypes.CDLL(build(force=True))
        _lib.sso_num_threads.restype = c_int
        _lib.sso_ssim_forward.argtypes = [c_int, c_int, c_int, c_float, c_float] + [c_void_p] * 6
        _lib.sso_ssim_backward.argtypes = [c_int, c_int, c_int] + [c_void_p] * 7
        _lib.sso_l1_ssim_loss.argtypes = [c_int, c_int, c_int, c_float, c_float, c_float, c_float,
                                          c_void_p, c_void_p, c_void_p, c_void_p]
    return _lib


def num_threads():
    return lib().sso_num_threads()


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a):
    return None if a is None else a.ctypes.data_as(c_void_p)


def fusedssim(C1_, C2_, img1, img2, train=True):
    """[B,CH,H,W] x2 -> (ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12); the last three are None if not train."""
    img1, img2 = _f32(img1), _f32(img2)
    B, CH, H, W = img1.shape
    m = np.zeros_like(img1)
    d = [np.zeros_like(img1) for _ in range(3)] if train else [None] * 3
    lib().sso_ssim_forward(B * CH, H, W, C1_, C2_, _ptr(img1), _ptr(img2), _ptr(m), *[_ptr(x) for x in d])
    return m, d[0], d[1], d[2]


def fusedssim_backward(C1_, C2_, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12):
    img1, img2, dL_dmap = _f32(img1), _f32(img2), _f32(dL_dmap)
    B, CH, H, W = img1.shape
    g = np.zeros_like(img1)
    lib().sso_ssim_backward(B * CH, H, W, _ptr(img1), _ptr(img2), _ptr(dL_dmap), _ptr(_f32(dm_dmu1)),
                            _ptr(_f32(dm_dsigma1_sq)), _ptr(_f32(dm_dsigma12)), _ptr(g))
    return g


def fused_ssim(img1, img2, padding="same", upstream=1.0):
    """Returns (mean ssim as float64, d(mean ssim * upstream)/dimg1)."""
    assert padding in ("same", "valid")
    m, d0, d1, d2 = fusedssim(C1, C2, img1, img2, True)
    dl = np.zeros_like(m)
    if padding == "valid":
        crop = m[:, :, 5:-5, 5:-5]
        dl[:, :, 5:-5, 5:-5] = np.float32(upstream / crop.size)
    else:
        crop = m
        dl[...] = np.float32(upstream / crop.size)
    g = fusedssim_backward(C1, C2, img1, img2, dl, d0, d1, d2)
    return float(crop.astype(np.float64).mean()), g


def l1_ssim_loss(img, gt, lambda_dssim=0.2, upstream=1.0, want_grad=True):
    """img, gt: [CH,H,W] or [B,CH,H,W].  Returns dict(loss, l1, ssim, dL_dimg)."""
    img, gt = _f32(img), _f32(gt)
    planes = int(np.prod(img.shape[:-2]))
    H, W = img.shape[-2:]
    out = (c_double * 3)()
    g = np.zeros_like(img) if want_grad else None
    lib().sso_l1_ssim_loss(planes, H, W, C1, C2, lambda_dssim, upstream, _ptr(img), _ptr(gt),
                           ctypes.cast(out, c_void_p), _ptr(g))
    return {"loss": out[0], "l1": out[1], "ssim": out[2], "dL_dimg": g}
