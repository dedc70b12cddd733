"""ctypes binding of the C ABI (include/gsr.h, libgsr_b200.so) on torch CUDA tensors.

This is the path the `-m gpu` parity tests and bench.py's device-timed leg call: Python only owns
device memory (torch tensors) and streams; every kernel is in libgsr_b200.so.  It also decodes the
opaque scratch buffers with gsr_get_*_layout so that tests can compare every intermediate
(depths, records, tiles_touched, offsets, sorted keys, point list, ranges, final_T, n_contrib).
No fallback: a missing library raises.
"""
import ctypes
import math
import os
from ctypes import (CFUNCTYPE, POINTER, Structure, c_char_p, c_double, c_float, c_int, c_size_t, c_ubyte,
                    c_uint32, c_uint64, c_void_p)

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GSR_B200_LIB") or os.path.join(_HERE, "libgsr_b200.so")   # override: kernel A/B experiments only

ALLOC_FN = CFUNCTYPE(c_void_p, c_size_t, c_void_p)


class ForwardArgs(Structure):
    _fields_ = [("P", c_int), ("D", c_int), ("M", c_int), ("width", c_int), ("height", c_int),
                ("background", c_void_p), ("means3D", c_void_p), ("shs", c_void_p),
                ("colors_precomp", c_void_p), ("opacities", c_void_p), ("scales", c_void_p),
                ("scale_modifier", c_float), ("rotations", c_void_p), ("cov3D_precomp", c_void_p),
                ("viewmatrix", c_void_p), ("projmatrix", c_void_p), ("cam_pos", c_void_p),
                ("tan_fovx", c_float), ("tan_fovy", c_float), ("prefiltered", c_int),
                ("antialiasing", c_int), ("debug", c_int), ("out_color", c_void_p),
                ("out_invdepth", c_void_p), ("radii", c_void_p)]


class BackwardArgs(Structure):
    _fields_ = [("P", c_int), ("D", c_int), ("M", c_int), ("R", c_int), ("width", c_int), ("height", c_int),
                ("background", c_void_p), ("means3D", c_void_p), ("shs", c_void_p),
                ("colors_precomp", c_void_p), ("opacities", c_void_p), ("scales", c_void_p),
                ("scale_modifier", c_float), ("rotations", c_void_p), ("cov3D_precomp", c_void_p),
                ("viewmatrix", c_void_p), ("projmatrix", c_void_p), ("cam_pos", c_void_p),
                ("tan_fovx", c_float), ("tan_fovy", c_float), ("radii", c_void_p),
                ("geom_buffer", c_void_p), ("binning_buffer", c_void_p), ("image_buffer", c_void_p),
                ("dL_dpix", c_void_p), ("dL_dinvdepth_pix", c_void_p), ("dL_dmean2D", c_void_p),
                ("dL_dconic", c_void_p), ("dL_dopacity", c_void_p), ("dL_dcolor", c_void_p),
                ("dL_dinvdepth", c_void_p), ("dL_dmean3D", c_void_p), ("dL_dcov3D", c_void_p),
                ("dL_dsh", c_void_p), ("dL_dscale", c_void_p), ("dL_drot", c_void_p),
                ("antialiasing", c_int), ("debug", c_int), ("accumulate", c_int)]


class GeometryLayout(Structure):
    _fields_ = [(n, c_size_t) for n in ("depths", "records", "cov3D", "clamped", "tiles_touched",
                                        "point_offsets", "depth_order", "sorted_offsets", "total")]


class BinningLayout(Structure):
    _fields_ = [(n, c_size_t) for n in ("tile_keys", "point_list", "tile_keys_alt", "point_list_alt",
                                        "sort_temp", "total")]


class ImageLayout(Structure):
    _fields_ = [(n, c_size_t) for n in ("final_T", "n_contrib", "ranges", "total")]


class AdamSegment(ctypes.Structure):
    """gsr_adam_segment (include/gsr_optim.h)."""
    _fields_ = [("param", c_void_p), ("grad", c_void_p), ("exp_avg", c_void_p), ("exp_avg_sq", c_void_p),
                ("rows", c_size_t), ("cols", c_int), ("split_col", c_int), ("lr_a", c_float), ("lr_b", c_float)]


EXPORTS = ["gsr_forward", "gsr_backward", "gsr_forward_batch", "gsr_backward_batch", "gsr_backward_blend", "gsr_backward_params_batch", "gsr_mark_visible", "gsr_geometry_bytes", "gsr_image_bytes",
           "gsr_binning_bytes", "gsr_sort_temp_bytes", "gsr_sort_pairs", "gsr_sort_static_max_pairs", "gsr_sort_pairs_static",
           "gsr_get_geometry_layout",
           "gsr_get_binning_layout", "gsr_get_image_layout", "gsr_materialize_keys", "gsr_set_profiling",
           "gsr_last_stage_ms",
           "gsr_last_error", "gsr_version",
           # include/gsr_loss.h
           "gsr_ssim_forward", "gsr_ssim_backward", "gsr_l1_ssim_partials_bytes", "gsr_l1_ssim_loss_forward",
           "gsr_l1_ssim_loss_backward", "gsr_exposure_clamp_forward", "gsr_exposure_partials_bytes",
           "gsr_exposure_clamp_backward",
           # include/gsr_optim.h
           "gsr_adam_step", "gsr_densification_stats",
           # include/gsr_dist.h (bound in gsr_b200/exchange.py)
           "gsr_exchange_signal_words", "gsr_exchange_shard", "gsr_exchange_step",
           # include/gsr_ply.h (bound in gsr_b200/ply.py)
           "gsr_ply_parse_header", "gsr_ply_write_header", "gsr_ply_unpack", "gsr_ply_pack",
           # include/gsr_frame.h (bound in gsr_b200/frames.py)
           "gsr_frame_to_rgb8"]

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libgsr_b200.so is not built (python gaussian-splatting-npu_b200/build.py); "
                               "there is no fallback path")
        L = ctypes.CDLL(LIB_PATH)
        L.gsr_forward.restype = c_int
        L.gsr_forward.argtypes = [POINTER(ForwardArgs), ALLOC_FN, c_void_p, ALLOC_FN, c_void_p, ALLOC_FN,
                                  c_void_p, c_void_p, POINTER(c_int)]
        L.gsr_backward.restype = c_int
        L.gsr_backward.argtypes = [POINTER(BackwardArgs), c_void_p]
        L.gsr_mark_visible.restype = c_int
        L.gsr_mark_visible.argtypes = [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]
        for n in ("gsr_geometry_bytes",):
            getattr(L, n).restype = c_size_t
            getattr(L, n).argtypes = [c_int]
        L.gsr_image_bytes.restype = c_size_t
        L.gsr_image_bytes.argtypes = [c_int, c_int]
        L.gsr_binning_bytes.restype = c_size_t
        L.gsr_binning_bytes.argtypes = [c_size_t, c_int, c_int]
        L.gsr_sort_temp_bytes.restype = c_size_t
        L.gsr_sort_temp_bytes.argtypes = [c_size_t]
        L.gsr_sort_pairs.restype = c_int
        L.gsr_sort_pairs.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_void_p, c_void_p]
        L.gsr_sort_static_max_pairs.restype = c_size_t
        L.gsr_sort_static_max_pairs.argtypes = []
        L.gsr_sort_pairs_static.restype = c_int
        L.gsr_sort_pairs_static.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_void_p]
        L.gsr_get_geometry_layout.argtypes = [c_int, POINTER(GeometryLayout)]
        L.gsr_get_binning_layout.argtypes = [c_size_t, c_int, c_int, POINTER(BinningLayout)]
        L.gsr_get_image_layout.argtypes = [c_int, c_int, POINTER(ImageLayout)]
        L.gsr_materialize_keys.restype = c_int
        L.gsr_materialize_keys.argtypes = [c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]
        L.gsr_forward_batch.restype = c_int
        L.gsr_forward_batch.argtypes = [c_int, POINTER(ForwardArgs), ALLOC_FN, POINTER(c_void_p), ALLOC_FN, POINTER(c_void_p), ALLOC_FN,
                                        POINTER(c_void_p), c_void_p, POINTER(c_void_p), POINTER(c_int)]
        L.gsr_backward_batch.restype = c_int
        L.gsr_backward_batch.argtypes = [c_int, POINTER(BackwardArgs), c_void_p, POINTER(c_void_p)]
        L.gsr_set_profiling.argtypes = [c_int]
        L.gsr_last_stage_ms.restype = c_float
        L.gsr_last_stage_ms.argtypes = [c_int]
        L.gsr_last_error.restype = c_char_p
        L.gsr_version.restype = c_char_p
        # loss stage (include/gsr_loss.h)
        L.gsr_ssim_forward.restype = c_int
        L.gsr_ssim_forward.argtypes = [c_int, c_int, c_int, c_float, c_float] + [c_void_p] * 7
        L.gsr_ssim_backward.restype = c_int
        L.gsr_ssim_backward.argtypes = [c_int, c_int, c_int] + [c_void_p] * 8
        L.gsr_l1_ssim_partials_bytes.restype = c_size_t
        L.gsr_l1_ssim_partials_bytes.argtypes = [c_int, c_int, c_int]
        L.gsr_l1_ssim_loss_forward.restype = c_int
        L.gsr_l1_ssim_loss_forward.argtypes = [c_int, c_int, c_int, c_float, c_float, c_float] + [c_void_p] * 8
        L.gsr_l1_ssim_loss_backward.restype = c_int
        L.gsr_l1_ssim_loss_backward.argtypes = [c_int, c_int, c_int, c_float] + [c_void_p] * 8
        L.gsr_exposure_clamp_forward.restype = c_int
        L.gsr_exposure_clamp_forward.argtypes = [c_int, c_int] + [c_void_p] * 4
        L.gsr_exposure_partials_bytes.restype = c_size_t
        L.gsr_exposure_partials_bytes.argtypes = [c_int, c_int]
        L.gsr_exposure_clamp_backward.restype = c_int
        L.gsr_exposure_clamp_backward.argtypes = [c_int, c_int] + [c_void_p] * 7
        # optimizer stage (include/gsr_optim.h)
        L.gsr_adam_step.restype = c_int
        L.gsr_adam_step.argtypes = [POINTER(AdamSegment), c_int, c_int, c_double, c_double, c_double, c_void_p, c_int, c_void_p]
        L.gsr_densification_stats.restype = c_int
        L.gsr_densification_stats.argtypes = [c_int] + [c_void_p] * 6
        _lib = L
    return _lib


def last_error():
    return lib().gsr_last_error().decode()


def _check(status):
    if status != 0:
        raise RuntimeError("gsr error %d: %s" % (status, last_error()))


def _ptr(t):
    if t is None or t.numel() == 0:
        return None
    assert t.is_cuda and t.is_contiguous()
    return c_void_p(t.data_ptr())


def _stream():
    return c_void_p(torch.cuda.current_stream().cuda_stream)


class _Buf:
    """One growable scratch tensor + the C callback that (re)allocates it."""

    def __init__(self, device):
        self.t = torch.empty(0, dtype=torch.uint8, device=device)
        self.device = device
        self.cb = ALLOC_FN(lambda nbytes, _user: self.alloc(nbytes))

    def alloc(self, nbytes):
        self.t = torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)
        return self.t.data_ptr()


class _GrowBuf(_Buf):
    """A scratch tensor that only grows (25 % headroom): a request that fits is served without touching the allocator."""

    def alloc(self, nbytes):
        if self.t.numel() < nbytes:
            self.t = torch.empty(0, dtype=torch.uint8, device=self.device)     # release before growing
            self.t = torch.empty(int(nbytes) + int(nbytes) // 4, dtype=torch.uint8, device=self.device)
        return self.t.data_ptr()


class Workspace:
    """Reusable scratch of gsr_forward (geometry / binning / image buffers) for callers that keep ONE forward state alive
    at a time per workspace (a forward-only render loop, or forward immediately followed by its backward).  The binning
    buffer's size follows the number of (Gaussian, tile) pairs of the view; sizing it per call makes the caching
    allocator split, merge and occasionally cudaMalloc -- a growth-only buffer takes that out of the loop.  The state
    returned by forward(workspace=ws) aliases these tensors: it is valid until the next forward with the same ws."""

    def __init__(self, device):
        self.geom, self.binning, self.img = _GrowBuf(device), _GrowBuf(device), _GrowBuf(device)


STAGES = ["preprocess", "scan", "duplicate", "sort", "ranges", "render_fwd", "render_bwd",
          "preprocess_bwd", "host_sync", "depth_sort"]


def stage_times_ms():
    L = lib()
    return {n: float(L.gsr_last_stage_ms(i)) for i, n in enumerate(STAGES)}


def set_profiling(on):
    lib().gsr_set_profiling(1 if on else 0)


def forward(*, means3D, opacities, viewmatrix, projmatrix, cam_pos, background, width, height,
            tan_fovx, tan_fovy, shs=None, colors_precomp=None, scales=None, rotations=None,
            cov3D_precomp=None, sh_degree=3, scale_modifier=1.0, prefiltered=False,
            antialiasing=False, debug=False, want_invdepth=True, workspace=None):
    """gsr_forward on torch CUDA tensors.  Returns a state dict holding outputs + scratch buffers (fresh per call, or
    the caller's `workspace`)."""
    L = lib()
    dev = means3D.device
    P = means3D.shape[0]
    out_color = torch.empty(3, height, width, dtype=torch.float32, device=dev)
    out_invdepth = torch.empty(1, height, width, dtype=torch.float32, device=dev) if want_invdepth else None
    radii = torch.empty(P, dtype=torch.int32, device=dev)
    M = 0 if shs is None else shs.shape[1]
    a = ForwardArgs(P, sh_degree, M, width, height, _ptr(background), _ptr(means3D), _ptr(shs),
                    _ptr(colors_precomp), _ptr(opacities), _ptr(scales), scale_modifier, _ptr(rotations),
                    _ptr(cov3D_precomp), _ptr(viewmatrix), _ptr(projmatrix), _ptr(cam_pos), tan_fovx, tan_fovy,
                    int(prefiltered), int(antialiasing), int(debug), _ptr(out_color), _ptr(out_invdepth), _ptr(radii))
    if workspace is not None:
        geom, binning, img = workspace.geom, workspace.binning, workspace.img
    else:
        geom, binning, img = _Buf(dev), _Buf(dev), _Buf(dev)
    n = c_int(0)
    with torch.cuda.device(dev):
        _check(L.gsr_forward(ctypes.byref(a), geom.cb, None, binning.cb, None, img.cb, None, _stream(), ctypes.byref(n)))
    return dict(num_rendered=n.value, out_color=out_color, out_invdepth=out_invdepth, radii=radii,
                geom=geom.t, binning=binning.t, img=img.t, P=P, W=width, H=height, M=M)


def backward(state, *, dL_dcolor, dL_dinvdepth=None, means3D, opacities, viewmatrix, projmatrix, cam_pos,
             background, tan_fovx, tan_fovy, shs=None, colors_precomp=None, scales=None, rotations=None,
             cov3D_precomp=None, sh_degree=3, scale_modifier=1.0, antialiasing=False, debug=False, out=None,
             accumulate=False):
    """gsr_backward.  `out` may map output names to preallocated contiguous tensors (e.g. views of a
    gradient bucket) that the kernels then write directly; with `accumulate` the five parameter
    gradients are added into them (visible rows only) instead of overwriting."""
    L = lib()
    dev = means3D.device
    P, W, H, M = state["P"], state["W"], state["H"], state["M"]
    out = out or {}
    shapes = dict(dL_dmeans2D=(P, 3), dL_dconic=(P, 2, 2), dL_dopacity=(P, 1), dL_dcolors=(P, 3),
                  dL_dinvdepths=(P, 1) if dL_dinvdepth is not None else None, dL_dmeans3D=(P, 3), dL_dcov3D=(P, 6),
                  dL_dsh=(P, M, 3) if M else None, dL_dscales=(P, 3) if scales is not None else None,
                  dL_drotations=(P, 4) if scales is not None else None)
    g = {}
    for k, shp in shapes.items():
        v = out.get(k) if shp is not None else None
        if v is not None:       # the caller's tensor is written in place; nothing is allocated for this output
            assert v.is_contiguous() and v.numel() == math.prod(shp) and v.data_ptr() % 16 == 0 and v.dtype == torch.float32
            g[k] = v
        else:
            g[k] = torch.empty(*shp, dtype=torch.float32, device=dev) if shp is not None else None
    a = BackwardArgs(P, sh_degree, M, state["num_rendered"], W, H, _ptr(background), _ptr(means3D), _ptr(shs),
                     _ptr(colors_precomp), _ptr(opacities), _ptr(scales), scale_modifier, _ptr(rotations),
                     _ptr(cov3D_precomp), _ptr(viewmatrix), _ptr(projmatrix), _ptr(cam_pos), tan_fovx, tan_fovy,
                     _ptr(state["radii"]), _ptr(state["geom"]), _ptr(state["binning"]), _ptr(state["img"]),
                     _ptr(dL_dcolor), _ptr(dL_dinvdepth), _ptr(g["dL_dmeans2D"]), _ptr(g["dL_dconic"]),
                     _ptr(g["dL_dopacity"]), _ptr(g["dL_dcolors"]), _ptr(g["dL_dinvdepths"]), _ptr(g["dL_dmeans3D"]),
                     _ptr(g["dL_dcov3D"]), _ptr(g["dL_dsh"]), _ptr(g["dL_dscales"]), _ptr(g["dL_drotations"]),
                     int(antialiasing), int(debug), int(accumulate))
    with torch.cuda.device(dev):
        _check(L.gsr_backward(ctypes.byref(a), _stream()))
    return g


_CAMERA_KEYS = ("viewmatrix", "projmatrix", "cam_pos", "tan_fovx", "tan_fovy")


def _view_streams(streams, V):
    if not streams:
        return None
    return (c_void_p * V)(*[c_void_p(streams[v % len(streams)].cuda_stream) for v in range(V)])


def forward_batch(cameras, *, means3D, opacities, background, width, height, shs, scales, rotations, sh_degree=3,
                  scale_modifier=1.0, prefiltered=False, antialiasing=False, debug=False, want_invdepth=True, workspaces=None,
                  streams=None):
    """gsr_forward_batch: the views of a step in one call -- the per-Gaussian stage runs once for all cameras (parameters
    and SH rows read once), the rest per view.  `cameras`: list of dicts(viewmatrix, projmatrix, cam_pos, tan_fovx,
    tan_fovy[, width, height]).  `streams`: optional torch.cuda.Stream list; view v's chain (scan, sorts, emission, blend)
    runs on streams[v % len(streams)], the one-launch stage on the current stream, which continues after all of them.
    Returns one state per view, identical to forward()'s."""
    L = lib()
    dev = means3D.device
    P, M, V = means3D.shape[0], shs.shape[1], len(cameras)
    arr = (ForwardArgs * V)()
    bufs, outs = [], []
    for v, c in enumerate(cameras):
        w, h = c.get("width", width), c.get("height", height)
        out_color = torch.empty(3, h, w, dtype=torch.float32, device=dev)
        out_invdepth = torch.empty(1, h, w, dtype=torch.float32, device=dev) if want_invdepth else None
        radii = torch.empty(P, dtype=torch.int32, device=dev)
        arr[v] = ForwardArgs(P, sh_degree, M, w, h, _ptr(background), _ptr(means3D), _ptr(shs), None, _ptr(opacities),
                             _ptr(scales), scale_modifier, _ptr(rotations), None, _ptr(c["viewmatrix"]), _ptr(c["projmatrix"]),
                             _ptr(c["cam_pos"]), c["tan_fovx"], c["tan_fovy"], int(prefiltered), int(antialiasing), int(debug),
                             _ptr(out_color), _ptr(out_invdepth), _ptr(radii))
        ws = workspaces[v] if workspaces is not None else None
        bufs.append((ws.geom, ws.binning, ws.img) if ws is not None else (_Buf(dev), _Buf(dev), _Buf(dev)))
        outs.append((out_color, out_invdepth, radii, w, h))
    # one callback per buffer kind; the user pointer carries the view index (+1: NULL arrives as None)
    cbs = [ALLOC_FN(lambda nbytes, user, k=k: bufs[int(user) - 1][k].alloc(nbytes)) for k in range(3)]
    users = (c_void_p * V)(*[c_void_p(v + 1) for v in range(V)])
    n = (c_int * V)()
    with torch.cuda.device(dev):
        _check(L.gsr_forward_batch(V, arr, cbs[0], users, cbs[1], users, cbs[2], users, _stream(), _view_streams(streams, V), n))
    return [dict(num_rendered=n[v], out_color=outs[v][0], out_invdepth=outs[v][1], radii=outs[v][2], geom=bufs[v][0].t,
                 binning=bufs[v][1].t, img=bufs[v][2].t, P=P, W=outs[v][3], H=outs[v][4], M=M) for v in range(V)]


def backward_batch(states, cameras, *, dL_dcolor, dL_dinvdepth=None, means3D, opacities, background, shs, scales, rotations,
                   sh_degree=3, scale_modifier=1.0, antialiasing=False, debug=False, out=None, accumulate=False, scratch=None,
                   streams=None):
    """gsr_backward_batch: per view the backward blend, then ONE per-Gaussian launch that sums the views' gradients in
    registers and writes (or, with `accumulate`, adds) the five parameter gradients once.  dL_dcolor / dL_dinvdepth: one
    tensor per view (or one shared by all).  `out`: the five parameter-gradient tensors (e.g. views of the exchange
    bucket); `scratch`: optional list of per-view dicts with dL_dmeans2D / dL_dconic / dL_dcolors / dL_dinvdepths.
    Returns (parameter gradients, list of per-view dicts)."""
    L = lib()
    dev = means3D.device
    V = len(states)
    P, M = states[0]["P"], states[0]["M"]
    e = lambda *s: torch.empty(*s, dtype=torch.float32, device=dev)
    out = out or {}
    g = {}
    for k, shp in dict(dL_dmeans3D=(P, 3), dL_dsh=(P, M, 3), dL_dopacity=(P, 1), dL_dscales=(P, 3), dL_drotations=(P, 4)).items():
        g[k] = out[k] if out.get(k) is not None else e(*shp)
        assert g[k].is_contiguous() and g[k].numel() == math.prod(shp) and g[k].data_ptr() % 16 == 0
    per_view = []
    arr = (BackwardArgs * V)()
    for v, (st, c) in enumerate(zip(states, cameras)):
        s = scratch[v] if scratch is not None else {}
        pv = dict(dL_dmeans2D=s.get("dL_dmeans2D", None), dL_dconic=s.get("dL_dconic", None), dL_dcolors=s.get("dL_dcolors", None),
                  dL_dinvdepths=s.get("dL_dinvdepths", None))
        pv["dL_dmeans2D"] = pv["dL_dmeans2D"] if pv["dL_dmeans2D"] is not None else e(P, 3)
        pv["dL_dconic"] = pv["dL_dconic"] if pv["dL_dconic"] is not None else e(P, 2, 2)
        pv["dL_dcolors"] = pv["dL_dcolors"] if pv["dL_dcolors"] is not None else e(P, 3)
        dcol = dL_dcolor[v] if isinstance(dL_dcolor, (list, tuple)) else dL_dcolor
        ddep = dL_dinvdepth[v] if isinstance(dL_dinvdepth, (list, tuple)) else dL_dinvdepth
        if ddep is not None and pv["dL_dinvdepths"] is None:
            pv["dL_dinvdepths"] = e(P, 1)
        first = v == 0
        arr[v] = BackwardArgs(P, sh_degree, M, st["num_rendered"], st["W"], st["H"], _ptr(background), _ptr(means3D), _ptr(shs),
                              None, _ptr(opacities), _ptr(scales), scale_modifier, _ptr(rotations), None, _ptr(c["viewmatrix"]),
                              _ptr(c["projmatrix"]), _ptr(c["cam_pos"]), c["tan_fovx"], c["tan_fovy"], _ptr(st["radii"]),
                              _ptr(st["geom"]), _ptr(st["binning"]), _ptr(st["img"]), _ptr(dcol), _ptr(ddep),
                              _ptr(pv["dL_dmeans2D"]), _ptr(pv["dL_dconic"]), _ptr(g["dL_dopacity"]) if first else None,
                              _ptr(pv["dL_dcolors"]), _ptr(pv["dL_dinvdepths"]) if ddep is not None else None,
                              _ptr(g["dL_dmeans3D"]) if first else None, None, _ptr(g["dL_dsh"]) if first else None,
                              _ptr(g["dL_dscales"]) if first else None, _ptr(g["dL_drotations"]) if first else None,
                              int(antialiasing), int(debug), int(accumulate))
        per_view.append(pv)
    with torch.cuda.device(dev):
        _check(L.gsr_backward_batch(V, arr, _stream(), _view_streams(streams, V)))
    return g, per_view


def mark_visible(means3D, viewmatrix, projmatrix):
    P = means3D.shape[0]
    out = torch.zeros(P, dtype=torch.bool, device=means3D.device)
    with torch.cuda.device(means3D.device):
        _check(lib().gsr_mark_visible(P, _ptr(means3D), _ptr(viewmatrix), _ptr(projmatrix), c_void_p(out.data_ptr()), _stream()))
    return out


def _aligned(t):
    off = (-t.data_ptr()) % 256
    return t[off:]


def _view(buf, offset, count, dtype):
    esize = torch.empty(0, dtype=dtype).element_size()
    return _aligned(buf)[offset:offset + count * esize].view(dtype)


def introspect(state):
    """Decode the scratch buffers of a forward `state` into named torch views."""
    L = lib()
    P, W, H, N = state["P"], state["W"], state["H"], state["num_rendered"]
    gl, bl, il = GeometryLayout(), BinningLayout(), ImageLayout()
    L.gsr_get_geometry_layout(P, ctypes.byref(gl))
    L.gsr_get_binning_layout(N, W, H, ctypes.byref(bl))
    L.gsr_get_image_layout(W, H, ctypes.byref(il))
    T = ((W + 15) // 16) * ((H + 15) // 16)
    rec = _view(state["geom"], gl.records, P * 12, torch.float32).view(P, 12)
    clamp = _view(state["geom"], gl.clamped, P, torch.uint8)
    o = dict(
        depths=_view(state["geom"], gl.depths, P, torch.float32),
        means2D=rec[:, 0:2], conic_opacity=rec[:, [2, 4, 3, 5]], rgb=rec[:, 6:9], invdepth_g=rec[:, 9],
        cov3D=_view(state["geom"], gl.cov3D, P * 6, torch.float32).view(P, 6),
        clamped=torch.stack([(clamp >> c) & 1 for c in range(3)], dim=1),
        tiles_touched=_view(state["geom"], gl.tiles_touched, P, torch.int32),
        point_offsets=_view(state["geom"], gl.point_offsets, P, torch.int32),
        final_T=_view(state["img"], il.final_T, H * W, torch.float32).view(H, W),
        n_contrib=_view(state["img"], il.n_contrib, H * W, torch.int32).view(H, W),
        ranges=_view(state["img"], il.ranges, T * 2, torch.int32).view(T, 2),
    )
    if N:
        o["tile_keys"] = _view(state["binning"], bl.tile_keys, N, torch.int32)
        o["point_list"] = _view(state["binning"], bl.point_list, N, torch.int32)
        # the reference's 64-bit sorted keys, rebuilt on demand (the hot path never materialises them)
        keys = torch.empty(N, dtype=torch.int64, device=state["geom"].device)
        with torch.cuda.device(keys.device):
            _check(L.gsr_materialize_keys(P, N, W, H, c_void_p(state["geom"].data_ptr()), c_void_p(state["binning"].data_ptr()),
                                          c_void_p(keys.data_ptr()), _stream()))
        o["keys"] = keys
    else:
        o["tile_keys"] = torch.empty(0, dtype=torch.int32, device=state["geom"].device)
        o["keys"] = torch.empty(0, dtype=torch.int64, device=state["geom"].device)
        o["point_list"] = torch.empty(0, dtype=torch.int32, device=state["geom"].device)
    P_vis = int((state["radii"] > 0).sum().item())
    o["depth_order"] = _view(state["geom"], gl.depth_order, P, torch.int32)[:P_vis]
    o["sorted_offsets"] = _view(state["geom"], gl.sorted_offsets, P, torch.int32)[:P_vis]
    return o


def sort_pairs(keys, values, end_bit):
    """gsr_sort_pairs on int64 keys / int32 values (bit patterns of u64 / u32). Inputs are clobbered."""
    L = lib()
    n = keys.numel()
    ko, vo = torch.empty_like(keys), torch.empty_like(values)
    temp = torch.empty(int(L.gsr_sort_temp_bytes(n)), dtype=torch.uint8, device=keys.device)
    with torch.cuda.device(keys.device):
        _check(L.gsr_sort_pairs(_ptr(keys) if n else None, _ptr(values) if n else None, c_void_p(ko.data_ptr()) if n else None,
                                c_void_p(vo.data_ptr()) if n else None, n, end_bit, c_void_p(temp.data_ptr()), _stream()))
    return ko, vo


def sort_pairs_static(keys, values, ko, vo, end_bit):
    """gsr_sort_pairs_static: the 64-bit sort on the library's built-in scratch array, for callers that
    cannot allocate device memory.  All four tensors come from the caller (this wrapper allocates
    nothing, so that it can stand in for such a caller); keys/values are clobbered, the result lands in
    ko/vo.  At most gsr_sort_static_max_pairs() pairs; calls that overlap on the GPU race on the array."""
    L = lib()
    n = keys.numel()
    with torch.cuda.device(keys.device):
        _check(L.gsr_sort_pairs_static(_ptr(keys) if n else None, _ptr(values) if n else None,
                                       c_void_p(ko.data_ptr()) if n else None, c_void_p(vo.data_ptr()) if n else None,
                                       n, end_bit, _stream()))
    return ko, vo


# ---- loss stage (include/gsr_loss.h) ------------------------------------------------------------------
SSIM_C1 = 0.01 ** 2
SSIM_C2 = 0.03 ** 2


def _planes(img):
    H, W = img.shape[-2:]
    return (img.numel() // (H * W) if H * W else 0), H, W


def ssim_forward(img1, img2, train=True, C1=SSIM_C1, C2=SSIM_C2):
    """gsr_ssim_forward: (ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12); the last three are None if not train."""
    planes, H, W = _planes(img1)
    m = torch.empty_like(img1)
    d = [torch.empty_like(img1) for _ in range(3)] if train else [None] * 3
    with torch.cuda.device(img1.device):
        _check(lib().gsr_ssim_forward(planes, H, W, C1, C2, _ptr(img1), _ptr(img2), _ptr(m), *[_ptr(x) for x in d], _stream()))
    return m, d[0], d[1], d[2]


def ssim_backward(img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12):
    planes, H, W = _planes(img1)
    g = torch.empty_like(img1)
    with torch.cuda.device(img1.device):
        _check(lib().gsr_ssim_backward(planes, H, W, _ptr(img1), _ptr(img2), _ptr(dL_dmap), _ptr(dm_dmu1), _ptr(dm_dsigma1_sq),
                                       _ptr(dm_dsigma12), _ptr(g), _stream()))
    return g


class LossScratch:
    """Preallocated buffers of the fused L1 + D-SSIM loss for one image shape (no allocation per step)."""

    def __init__(self, shape, device):
        e = lambda: torch.empty(shape, dtype=torch.float32, device=device)
        self.dm = [e(), e(), e()]
        self.dL_dimage = e()
        self.loss = torch.empty(3, dtype=torch.float32, device=device)
        H, W = shape[-2:]
        planes = int(torch.Size(shape).numel() // (H * W))
        self.partials = torch.empty(int(lib().gsr_l1_ssim_partials_bytes(planes, H, W)) // 8, dtype=torch.float64, device=device)


def l1_ssim_loss_forward(image, gt, lambda_dssim, scratch, train=True, C1=SSIM_C1, C2=SSIM_C2):
    """gsr_l1_ssim_loss_forward: fills scratch.loss = (loss, mean L1, mean ssim) and the derivative maps."""
    planes, H, W = _planes(image)
    d = scratch.dm if train else [None] * 3
    with torch.cuda.device(image.device):
        _check(lib().gsr_l1_ssim_loss_forward(planes, H, W, C1, C2, lambda_dssim, _ptr(image), _ptr(gt), _ptr(scratch.loss),
                                              *[_ptr(x) for x in d], _ptr(scratch.partials), _stream()))
    return scratch.loss


def l1_ssim_loss_backward(image, gt, lambda_dssim, scratch, upstream=None):
    """gsr_l1_ssim_loss_backward into scratch.dL_dimage (upstream: device scalar tensor or None = 1)."""
    planes, H, W = _planes(image)
    with torch.cuda.device(image.device):
        _check(lib().gsr_l1_ssim_loss_backward(planes, H, W, lambda_dssim, _ptr(image), _ptr(gt), *[_ptr(x) for x in scratch.dm],
                                               _ptr(upstream), _ptr(scratch.dL_dimage), _stream()))
    return scratch.dL_dimage


def exposure_clamp_forward(image, exposure=None):
    """gsr_exposure_clamp_forward: clamp(exposure[:3,:3]^T . image + exposure[:3,3], 0, 1) of one [3,H,W] image
    (gaussian_renderer/__init__.py:112-119); exposure None = clamp only."""
    assert image.dim() == 3 and image.shape[0] == 3 and image.dtype == torch.float32
    if exposure is not None:
        assert exposure.dtype == torch.float32 and tuple(exposure.shape) == (3, 4)
    out = torch.empty_like(image)
    with torch.cuda.device(image.device):
        _check(lib().gsr_exposure_clamp_forward(image.shape[1], image.shape[2], _ptr(image), _ptr(exposure), _ptr(out), _stream()))
    return out


def exposure_clamp_backward(image, exposure, dL_dout, need_exposure_grad=True):
    """gsr_exposure_clamp_backward: (dL_dimage [3,H,W], dL_dexposure [3,4] or None)."""
    H, W = image.shape[1:]
    dimg = torch.empty_like(image)
    dexp = part = None
    if exposure is not None and need_exposure_grad:
        dexp = torch.zeros(3, 4, dtype=torch.float32, device=image.device)
        part = torch.empty(int(lib().gsr_exposure_partials_bytes(H, W)) // 8, dtype=torch.float64, device=image.device)
    with torch.cuda.device(image.device):
        _check(lib().gsr_exposure_clamp_backward(H, W, _ptr(image), _ptr(exposure), _ptr(dL_dout), _ptr(dimg), _ptr(dexp),
                                                 _ptr(part), _stream()))
    return dimg, dexp


# ---- optimizer stage (include/gsr_optim.h) --------------------------------------------------------------
def adam_step(segments, step, beta1=0.9, beta2=0.999, eps=1e-15, visible=None, zero_grad=False):
    """gsr_adam_step.  segments: list of dicts(param, grad, exp_avg, exp_avg_sq [rows, cols] contiguous CUDA fp32,
    lr or (lr_a, lr_b, split_col)).  One launch for all of them."""
    arr = (AdamSegment * len(segments))()
    for i, sg in enumerate(segments):
        p = sg["param"]
        rows = p.shape[0]
        cols = p.numel() // rows if rows else 1
        if "lr" in sg:
            lr_a = lr_b = float(sg["lr"])
            split = cols
        else:
            lr_a, lr_b, split = float(sg["lr_a"]), float(sg["lr_b"]), int(sg["split_col"])
        for k in ("param", "grad", "exp_avg", "exp_avg_sq"):
            assert sg[k].is_cuda and sg[k].is_contiguous() and sg[k].dtype == torch.float32 and sg[k].numel() == p.numel()
        arr[i] = AdamSegment(p.data_ptr(), sg["grad"].data_ptr(), sg["exp_avg"].data_ptr(), sg["exp_avg_sq"].data_ptr(),
                             rows, cols, split, lr_a, lr_b)
    vis = None
    if visible is not None:
        assert visible.is_cuda and visible.is_contiguous() and visible.dtype in (torch.bool, torch.uint8)
        vis = c_void_p(visible.data_ptr())
    with torch.cuda.device(segments[0]["param"].device):
        _check(lib().gsr_adam_step(arr, len(segments), int(step), beta1, beta2, eps, vis, int(zero_grad), _stream()))


def densification_stats(grad_means2D, radii, xyz_gradient_accum, denom, max_radii2D):
    """gsr_densification_stats (train.py:166-167, gaussian_model.py:471-473), in place."""
    P = radii.shape[0]
    assert grad_means2D.shape == (P, 3) and grad_means2D.is_contiguous() and radii.dtype == torch.int32
    with torch.cuda.device(radii.device):
        _check(lib().gsr_densification_stats(P, _ptr(grad_means2D), _ptr(radii), _ptr(xyz_gradient_accum), _ptr(denom),
                                             _ptr(max_radii2D), _stream()))
