"""ctypes driver of oracle/_ref/libref_rasterizer.so -- the UNMODIFIED reference CUDA rasterizer
(/root/reference/diff-gaussian-rasterization-npu/cuda_rasterizer/*.cu) built for sm_100a by
oracle/build_ref.py.  TEST INFRASTRUCTURE ONLY.  Needs a GPU; used by the `-m gpu` parity tests
(GPU-vs-GPU bit-exactness), by tests/golden/make_golden.py and by bench.py as the same-box bar.
"""
import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_longlong, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libref_rasterizer.so")
_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = ctypes.CDLL(LIB_PATH)
        L.ref_state_create.restype = c_void_p
        L.ref_state_destroy.argtypes = [c_void_p]
        L.ref_forward.restype = c_int
        L.ref_forward.argtypes = [c_void_p, c_int, c_int, c_int, c_void_p, c_int, c_int] + [c_void_p] * 5 + \
            [c_float] + [c_void_p] * 5 + [c_float, c_float, c_int, c_void_p, c_void_p, c_int, c_void_p, c_int]
        L.ref_backward.restype = c_int
        L.ref_backward.argtypes = [c_void_p, c_int, c_int] + [c_void_p] * 6 + [c_float] + [c_void_p] * 5 + \
            [c_float, c_float] + [c_void_p] * 13 + [c_int, c_int]
        L.ref_copy.restype = c_int
        L.ref_copy.argtypes = [c_void_p, c_void_p, ctypes.c_size_t]
        L.ref_get.restype = c_longlong
        L.ref_get.argtypes = [c_void_p, c_char_p, ctypes.POINTER(c_void_p)]
        _lib = L
    return _lib


def _ptr(t):
    if t is None or t.numel() == 0:
        return None
    assert t.is_cuda and t.is_contiguous()
    return c_void_p(t.data_ptr())


_DTYPES = dict(depths=torch.float32, clamped=torch.uint8, means2D=torch.float32, cov3D=torch.float32,
               conic_opacity=torch.float32, rgb=torch.float32, tiles_touched=torch.int32,
               point_offsets=torch.int32, keys=torch.int64, point_list=torch.int32, final_T=torch.float32,
               n_contrib=torch.int32, ranges=torch.int32, keys_unsorted=torch.int64,
               point_list_unsorted=torch.int32)


class RefRasterizer:
    def __init__(self):
        self.st = c_void_p(lib().ref_state_create())

    def close(self):
        if self.st:
            lib().ref_state_destroy(self.st)
            self.st = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def forward(self, *, means3D, opacities, viewmatrix, projmatrix, cam_pos, background, width, height,
                tan_fovx, tan_fovy, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, sh_degree=3, scale_modifier=1.0, prefiltered=False, antialiasing=False,
                debug=False):
        dev = means3D.device
        P = means3D.shape[0]
        M = 0 if shs is None else shs.shape[1]
        # the reference's binding zero-fills these (rasterize_points.cu:69-76)
        out_color = torch.zeros(3, height, width, dtype=torch.float32, device=dev)
        out_invdepth = torch.zeros(1, height, width, dtype=torch.float32, device=dev)
        radii = torch.zeros(P, dtype=torch.int32, device=dev)
        torch.cuda.synchronize()
        r = lib().ref_forward(self.st, P, sh_degree, M, _ptr(background), width, height, _ptr(means3D), _ptr(shs),
                              _ptr(colors_precomp), _ptr(opacities), _ptr(scales), scale_modifier, _ptr(rotations),
                              _ptr(cov3D_precomp), _ptr(viewmatrix), _ptr(projmatrix), _ptr(cam_pos), tan_fovx,
                              tan_fovy, int(prefiltered), _ptr(out_color), _ptr(out_invdepth), int(antialiasing),
                              _ptr(radii), int(debug))
        if r < 0:
            raise RuntimeError("reference forward failed (%d)" % r)
        self.P, self.W, self.H, self.M, self.R = P, width, height, M, r
        return dict(num_rendered=r, out_color=out_color, out_invdepth=out_invdepth, radii=radii)

    def get(self, name):
        """Copy of a named intermediate of the last forward as a torch tensor."""
        p = c_void_p()
        n = lib().ref_get(self.st, name.encode(), ctypes.byref(p))
        if n < 0:
            raise KeyError(name)
        dt = _DTYPES[name]
        out = torch.empty(int(n), dtype=dt, device="cuda")
        if n:
            nbytes = int(n) * out.element_size()
            if lib().ref_copy(c_void_p(out.data_ptr()), p, ctypes.c_size_t(nbytes)) != 0:
                raise RuntimeError("ref_copy failed")
        return out

    def intermediates(self):
        o = {k: self.get(k) for k in ("depths", "means2D", "cov3D", "conic_opacity", "rgb", "clamped",
                                      "tiles_touched", "point_offsets", "keys", "point_list", "final_T",
                                      "n_contrib", "ranges")}
        P, W, H = self.P, self.W, self.H
        o["means2D"] = o["means2D"].view(P, 2); o["cov3D"] = o["cov3D"].view(P, 6)
        o["conic_opacity"] = o["conic_opacity"].view(P, 4); o["rgb"] = o["rgb"].view(P, 3)
        o["clamped"] = o["clamped"].view(P, 3)
        o["final_T"] = o["final_T"].view(H, W); o["n_contrib"] = o["n_contrib"].view(H, W)
        o["ranges"] = o["ranges"].view(-1, 2)
        return o

    def backward(self, radii, *, dL_dcolor, dL_dinvdepth=None, means3D, opacities, viewmatrix, projmatrix,
                 cam_pos, background, tan_fovx, tan_fovy, shs=None, colors_precomp=None, scales=None,
                 rotations=None, cov3D_precomp=None, sh_degree=3, scale_modifier=1.0, antialiasing=False,
                 debug=False):
        dev = means3D.device
        P, M = self.P, self.M
        z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=dev)   # rasterize_points.cu:163-172
        g = dict(dL_dmeans2D=z(P, 3), dL_dconic=z(P, 2, 2), dL_dopacity=z(P, 1), dL_dcolors=z(P, 3),
                 dL_dinvdepths=z(P, 1) if dL_dinvdepth is not None else None, dL_dmeans3D=z(P, 3),
                 dL_dcov3D=z(P, 6), dL_dsh=z(P, M, 3), dL_dscales=z(P, 3), dL_drotations=z(P, 4))
        torch.cuda.synchronize()
        r = lib().ref_backward(self.st, sh_degree, M, _ptr(background), _ptr(means3D), _ptr(shs), _ptr(colors_precomp),
                               _ptr(opacities), _ptr(scales), scale_modifier, _ptr(rotations), _ptr(cov3D_precomp),
                               _ptr(viewmatrix), _ptr(projmatrix), _ptr(cam_pos), tan_fovx, tan_fovy, _ptr(radii),
                               _ptr(dL_dcolor), _ptr(dL_dinvdepth), _ptr(g["dL_dmeans2D"]), _ptr(g["dL_dconic"]),
                               _ptr(g["dL_dopacity"]), _ptr(g["dL_dcolors"]), _ptr(g["dL_dinvdepths"]),
                               _ptr(g["dL_dmeans3D"]), _ptr(g["dL_dcov3D"]), _ptr(g["dL_dsh"]), _ptr(g["dL_dscales"]),
                               _ptr(g["dL_drotations"]), int(antialiasing), int(debug))
        if r < 0:
            raise RuntimeError("reference backward failed (%d)" % r)
        return g
