"""The C-ABI library loads without a GPU, exports every symbol include/*.h declares, and its
argument validation / layout functions work (no compute calls)."""
import ctypes
import os
import re

import pytest

import cases  # noqa: F401  (sys.path)
from gsr_b200 import capi

ROOT = cases.ROOT


def _declared_functions():
    src = ""
    for h in sorted(os.listdir(os.path.join(ROOT, "include"))):
        if h.endswith(".h"):
            src += open(os.path.join(ROOT, "include", h)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gsr_[a-z_0-9]+)\s*\(", src)) - {"gsr_alloc_fn"})


def test_library_exports_every_declared_symbol():
    L = ctypes.CDLL(capi.LIB_PATH)
    names = _declared_functions()
    assert len(names) >= 14
    for n in names:
        assert hasattr(L, n), "libgsr_b200.so does not export " + n
    assert sorted(capi.EXPORTS) == names


def test_version_and_layouts():
    L = capi.lib()
    assert b"sm_100a" in L.gsr_version()
    g = capi.GeometryLayout()
    L.gsr_get_geometry_layout(1000, ctypes.byref(g))
    offs = [g.depths, g.records, g.cov3D, g.clamped, g.tiles_touched, g.point_offsets, g.depth_order,
            g.sorted_offsets, g.total]
    assert offs == sorted(offs) and all(o % 256 == 0 for o in offs[:-1])
    assert g.records - g.depths >= 4000 and g.cov3D - g.records >= 48000
    assert L.gsr_geometry_bytes(1000) == g.total
    b = capi.BinningLayout()
    L.gsr_get_binning_layout(12345, 640, 360, ctypes.byref(b))
    assert b.point_list - b.tile_keys >= 12345 * 4 and b.sort_temp > b.point_list_alt > b.tile_keys_alt
    assert b.total == L.gsr_binning_bytes(12345, 640, 360)
    i = capi.ImageLayout()
    L.gsr_get_image_layout(167, 101, ctypes.byref(i))
    assert i.n_contrib - i.final_T >= 167 * 101 * 4 and i.total == L.gsr_image_bytes(167, 101)
    assert L.gsr_sort_temp_bytes(0) > 0 and L.gsr_sort_temp_bytes(1 << 20) > L.gsr_sort_temp_bytes(1 << 10)


def test_argument_validation_without_gpu():
    L = capi.lib()
    n = ctypes.c_int(-7)
    cb = capi.ALLOC_FN(lambda nbytes, user: None)
    assert L.gsr_forward(None, cb, None, cb, None, cb, None, None, ctypes.byref(n)) == -1
    assert b"null" in L.gsr_last_error()
    a = capi.ForwardArgs()
    a.P, a.width, a.height = 10, 0, 16
    assert L.gsr_forward(ctypes.byref(a), cb, None, cb, None, cb, None, None, ctypes.byref(n)) == -1
    a.width = 16
    a.out_color = 1   # non-null sentinel; P=10 but inputs missing -> invalid before anything is touched
    assert L.gsr_forward(ctypes.byref(a), cb, None, cb, None, cb, None, None, ctypes.byref(n)) == -1
    assert b"missing" in L.gsr_last_error()
    assert L.gsr_backward(None, None) == -1
    assert L.gsr_mark_visible(-1, None, None, None, None, None) == -1
    assert L.gsr_mark_visible(0, None, None, None, None, None) == 0
    assert L.gsr_sort_pairs(None, None, None, None, 0, 80, None, None) == -1
    assert L.gsr_last_stage_ms(0) == -1.0
    # the allocation-free sort refuses what its built-in array cannot hold, before touching the device
    assert L.gsr_sort_static_max_pairs() == 1 << 20
    assert L.gsr_sort_pairs_static(1, 1, 1, 1, (1 << 20) + 1, 45, None) == -1
    assert b"GSR_SORT_STATIC_MAX_PAIRS" in L.gsr_last_error()
    assert L.gsr_sort_pairs_static(None, None, None, None, 0, 45, None) == 0


def test_backward_batch_validates_before_any_cuda_call():
    """gsr_backward_batch / _blend / _params_batch reject bad arguments with GSR_ERR_INVALID before they touch the device
    (on this GPU-less box a CUDA call would come back as GSR_ERR_CUDA = -2 instead): SH degree outside 0..3, missing
    camera pointers, views that do not share their Gaussians."""
    L = capi.lib()
    assert L.gsr_backward_batch(1, None, None, None) == -1
    arr = (capi.BackwardArgs * 2)()
    for a in arr:
        a.P, a.D, a.M, a.R, a.width, a.height = 10, 3, 16, 5, 32, 32
        a.scale_modifier = 1.0
        for name in ("background", "means3D", "shs", "opacities", "scales", "rotations", "viewmatrix", "projmatrix", "cam_pos",
                     "radii", "geom_buffer", "binning_buffer", "image_buffer", "dL_dpix", "dL_dmean2D", "dL_dconic",
                     "dL_dopacity", "dL_dcolor", "dL_dmean3D", "dL_dsh", "dL_dscale", "dL_drot"):
            setattr(a, name, 256)                         # aligned non-null sentinel, never dereferenced on the host
    assert L.gsr_backward_batch(9, arr, None, None) == -1 and b"n_views" in L.gsr_last_error()
    arr[1].D = 5
    assert L.gsr_backward_batch(2, arr, None, None) == -1 and b"SH degree" in L.gsr_last_error()
    arr[1].D = 3
    arr[1].cam_pos = None
    assert L.gsr_backward_params_batch(2, arr, None) == -1 and b"camera" in L.gsr_last_error()
    arr[1].cam_pos = 256
    arr[1].opacities = 512
    assert L.gsr_backward_batch(2, arr, None, None) == -1 and b"share" in L.gsr_last_error()
    arr[1].opacities = 256
    arr[0].dL_dconic = 260
    assert L.gsr_backward_blend(arr, None) == -1 and b"aligned" in L.gsr_last_error()


def test_python_api_has_no_cpu_fallback():
    import torch
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    rs = GaussianRasterizationSettings(8, 8, 1.0, 1.0, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 3,
                                       torch.zeros(3), False, False, False)
    r = GaussianRasterizer(rs)
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(torch.zeros(4, 3), torch.zeros(4, 3), torch.zeros(4, 1), scales=torch.zeros(4, 3), rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair or precomputed 3D covariance"):
        r(torch.zeros(4, 3), torch.zeros(4, 3), torch.zeros(4, 1), shs=torch.zeros(4, 16, 3), scales=torch.zeros(4, 3))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        r(torch.zeros(4, 3), torch.zeros(4, 3), torch.zeros(4, 1), shs=torch.zeros(4, 16, 3), scales=torch.zeros(4, 3),
          rotations=torch.zeros(4, 4))
    with pytest.raises(RuntimeError, match="means3D must have dimensions"):
        r(torch.zeros(4, 2), torch.zeros(4, 3), torch.zeros(4, 1), shs=torch.zeros(4, 16, 3), scales=torch.zeros(4, 3),
          rotations=torch.zeros(4, 4))
    assert list(GaussianRasterizationSettings._fields) == [
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug", "antialiasing"]


def test_loss_stage_argument_validation_without_gpu():
    L = capi.lib()
    assert L.gsr_l1_ssim_partials_bytes(3, 1080, 1920) == 60 * 34 * 3 * 16 + 256
    # empty problems are no-ops, NULL images are rejected before any launch
    assert L.gsr_ssim_forward(0, 16, 16, 1e-4, 9e-4, None, None, None, None, None, None, None) == 0
    assert L.gsr_ssim_forward(1, 16, 16, 1e-4, 9e-4, None, None, None, None, None, None, None) == -1
    assert "NULL" in capi.last_error()
    assert L.gsr_ssim_backward(1, 16, 16, None, None, None, None, None, None, None, None) == -1
    assert L.gsr_ssim_forward(70000, 16, 16, 1e-4, 9e-4, None, None, None, None, None, None, None) == -1
    assert "65535" in capi.last_error()


def test_fused_ssim_package_has_no_cpu_fallback():
    import torch
    from fused_ssim import fused_ssim, l1_ssim_loss, allowed_padding
    assert allowed_padding == ["same", "valid"]
    a = torch.rand(1, 3, 16, 16)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        fused_ssim(a, a)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        l1_ssim_loss(a[0], a[0])
    with pytest.raises(AssertionError):
        fused_ssim(a, a, padding="reflect")


def test_headers_are_plain_c():
    """The boundary is a C ABI: every header under include/ compiles as C99 on its own (no C++, no torch, no CUDA types)."""
    import glob
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for h in sorted(glob.glob(os.path.join(root, "include", "*.h"))):
        r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-x", "c", "-fsyntax-only", "-"],
                           input='#include "%s"\n' % h, capture_output=True, text=True)
        assert r.returncode == 0, (h, r.stderr[:500])
