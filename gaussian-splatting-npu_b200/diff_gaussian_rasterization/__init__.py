"""Drop-in `diff_gaussian_rasterization` package backed by the B200-native rasterizer.

Public surface (names, argument order, return tuples, error text) is the reference's
/root/reference/diff-gaussian-rasterization-npu/diff_gaussian_rasterization/__init__.py:
  GaussianRasterizationSettings (:143-156), GaussianRasterizer (:158-207),
  rasterize_gaussians (:21-42), _RasterizeGaussians (:44-141)
so `gaussian_renderer/__init__.py` and `train.py` of the reference run on it unchanged.
`SparseGaussianAdam` is deliberately NOT exported (train.py:37-41 then keeps separate_sh off).

The compute path is the in-tree torch extension `_C` (csrc/torch_binding.cpp) over the C ABI of
libgsr_b200.so (include/gsr.h).  There is no CPU or PyTorch fallback: if the extension is missing
the import fails loudly.
"""
from typing import NamedTuple

import torch
import torch.nn as nn

try:
    from . import _C
except ImportError as exc:  # pragma: no cover - exercised only on a broken build
    raise ImportError(
        "diff_gaussian_rasterization._C (the sm_100a CUDA extension) is not built or cannot be "
        "loaded; run `python gaussian-splatting-npu_b200/build.py`. There is no fallback path."
    ) from exc

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians", "grad_sink"]

import contextlib


class _SinkState:      # process-wide on purpose: autograd runs CUDA backward nodes on its own device threads
    sink = None


_tls = _SinkState


@contextlib.contextmanager
def grad_sink(means3D, shs, opacities, scales, rotations, accumulate=False, defer=False):
    """Extension over the reference API (opt-in; nothing changes unless it is used).

    Inside this context every `backward()` through the rasterizer writes the five PARAMETER gradients
    straight into the given preallocated tensors -- typically views of one flat all-reduce bucket
    (`gsr_b200.dist.FlatGaussians.grad_views()`) -- and hands autograd `None` for those inputs, so no
    [P,59] gradient is allocated, copied or added by AccumulateGrad.  With `accumulate=True` the kernels
    ADD into the visible rows (second and later views of one training step); with False they overwrite
    every row (first view).  means2D / colors_precomp / cov3D_precomp gradients flow through autograd
    as usual (densification reads means2D.grad per view).

    `defer=True` is for a context that spans ALL views of a step: every backward runs its blend stage at once and the
    per-Gaussian stage is postponed; when the context closes, ONE launch per (up to) 8 views sums their parameter
    gradients in registers and writes the sink once (gsr_backward_params_batch; `accumulate` then says whether that
    first write adds or overwrites).  The sink tensors hold the step's gradients only after the context has closed.
    Configurations the batched kernel does not cover (M != 16, colors_precomp, cov3D_precomp) run undeferred: the first
    such backward follows `accumulate`, so give every view of such a step its own context as before.
    Views share a launch only when they pass the SAME five parameter tensors (same storage) to the rasterizer: a caller
    that builds them afresh per render -- the reference's `pc.get_features` is a new `torch.cat` every time, `get_opacity`
    a new sigmoid (scene/gaussian_model.py) -- gets a correct result, but each view is then written out on its own when
    the next one arrives, i.e. nothing is batched.  Compute the activated tensors once per step to get the batching.
    The write-out happens on the stream that is current where the context closes (or where a later view forces it); it
    waits for the blends of the parked views, whatever stream those ran on.
    """
    prev = getattr(_tls, "sink", None)
    none = torch.empty(0)
    _tls.sink = tuple(none if t is None else t for t in (means3D, shs, opacities, scales, rotations)) + (bool(accumulate), bool(defer))
    try:
        yield
    finally:
        _tls.sink = prev
        if defer:
            _C.flush_grad_sink()


class GaussianRasterizationSettings(NamedTuple):
    """Per-call camera / image configuration.  Field order is part of the API (reference :143-156)."""
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    antialiasing: bool


def _camera_args(rs):
    return (rs.scale_modifier,), (rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy)


class _RasterizeGaussians(torch.autograd.Function):
    """autograd bridge: 9 inputs -> (color[3,H,W], radii[P] int32, invdepth[1,H,W])  (reference :44-141)."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                cov3Ds_precomp, raster_settings):
        rs = raster_settings
        (mod,), cam = _camera_args(rs)
        # argument order of RasterizeGaussiansCUDA (rasterize_points.h:18-39)
        out = _C.rasterize_gaussians(
            rs.bg, means3D, colors_precomp, opacities, scales, rotations, mod, cov3Ds_precomp,
            *cam, rs.image_height, rs.image_width, sh, rs.sh_degree, rs.campos,
            rs.prefiltered, rs.antialiasing, rs.debug)
        num_rendered, color, radii, geom_buf, binning_buf, img_buf, invdepths = out
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
                              opacities, geom_buf, binning_buf, img_buf)
        return color, radii, invdepths

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii, grad_out_depth):
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, opacities,
         geom_buf, binning_buf, img_buf) = ctx.saved_tensors
        (mod,), cam = _camera_args(rs)
        sink = getattr(_tls, "sink", None)
        if sink is not None:
            (grad_means2D, grad_colors_precomp, _o, _m, grad_cov3Ds_precomp, _s, _sc, _r) = \
                _C.rasterize_gaussians_backward_into(
                    rs.bg, means3D, radii, colors_precomp, opacities, scales, rotations, mod, cov3Ds_precomp,
                    *cam, grad_out_color, grad_out_depth, sh, rs.sh_degree, rs.campos, geom_buf,
                    ctx.num_rendered, binning_buf, img_buf, rs.antialiasing, rs.debug, *sink)
            return (None, grad_means2D, None, grad_colors_precomp, None, None, None, grad_cov3Ds_precomp, None)
        # argument order of RasterizeGaussiansBackwardCUDA (rasterize_points.cu:126-151)
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp,
         grad_sh, grad_scales, grad_rotations) = _C.rasterize_gaussians_backward(
            rs.bg, means3D, radii, colors_precomp, opacities, scales, rotations, mod, cov3Ds_precomp,
            *cam, grad_out_color, grad_out_depth, sh, rs.sh_degree, rs.campos, geom_buf,
            ctx.num_rendered, binning_buf, img_buf, rs.antialiasing, rs.debug)
        # one gradient per forward input, in input order; raster_settings gets None
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities,
                grad_scales, grad_rotations, grad_cov3Ds_precomp, None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                        cov3Ds_precomp, raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales,
                                     rotations, cov3Ds_precomp, raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """bool[P]: which points pass the near-plane frustum test of this camera (reference :163-172)."""
        rs = self.raster_settings
        with torch.no_grad():
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None, dc=None):
        # `dc` is what gaussian_renderer/__init__.py:90-100 passes on its separate_sh path (only taken when
        # SparseGaussianAdam is importable from this package, i.e. under GSR_EXPORT_SPARSE_ADAM=1): the degree-0
        # coefficients [P,1,3] apart from the rest [P,M-1,3]; the kernels read one [P,M,3] tensor
        if dc is not None and shs is not None:
            shs = torch.cat([dc, shs], dim=1)
        elif dc is not None:
            shs = dc
        have_sh, have_col = shs is not None, colors_precomp is not None
        if have_sh == have_col:
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        have_sr = scales is not None or rotations is not None
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                (have_sr and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')

        def or_empty(t):
            return torch.Tensor([]) if t is None else t

        return rasterize_gaussians(means3D, means2D, or_empty(shs), or_empty(colors_precomp),
                                   opacities, or_empty(scales), or_empty(rotations),
                                   or_empty(cov3D_precomp), self.raster_settings)


# The reference's own rasterizer package does not export SparseGaussianAdam, which keeps train.py:37-41 on its
# torch.optim.Adam / single-SH-tensor path.  Opt in to the sparse optimizer + separate_sh path:
import os as _os

if _os.environ.get("GSR_EXPORT_SPARSE_ADAM") == "1":
    from gsr_b200.optim import SparseGaussianAdam  # noqa: E402,F401
    __all__.append("SparseGaussianAdam")
