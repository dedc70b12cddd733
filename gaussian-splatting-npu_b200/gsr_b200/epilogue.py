"""Render epilogue between the rasterizer and the loss: exposure affine + clamp, one launch each way.

Host mirror of /root/reference/gaussian_renderer/__init__.py:112-119
    if use_trained_exp:
        exposure = pc.get_exposure_from_name(...)                                   # [3,4]
        rendered_image = matmul(rendered_image.permute(1,2,0), exposure[:3,:3]).permute(2,0,1) + exposure[:3,3,None,None]
    rendered_image = rendered_image.clamp(0, 1)
over include/gsr_loss.h (gsr_exposure_clamp_forward / _backward in libgsr_b200.so).  No CPU path."""
import torch

from . import capi


class _ExposureClamp(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, exposure):
        image = image.contiguous()
        expo = exposure.contiguous() if exposure is not None else None
        ctx.save_for_backward(image, expo)
        return capi.exposure_clamp_forward(image, expo)

    @staticmethod
    def backward(ctx, dL_dout):
        image, expo = ctx.saved_tensors
        need_e = expo is not None and ctx.needs_input_grad[1]
        dimg, dexp = capi.exposure_clamp_backward(image, expo, dL_dout.contiguous(), need_e)
        return dimg, dexp


def exposure_clamp(rendered_image, exposure=None):
    """clamp(apply_exposure(rendered_image), 0, 1); exposure: [3,4] (pc.get_exposure_from_name) or None when
    use_trained_exp is off.  Differentiable w.r.t. both."""
    if not rendered_image.is_cuda:
        raise RuntimeError("exposure_clamp needs CUDA tensors: there is no CPU path")
    return _ExposureClamp.apply(rendered_image, exposure)
