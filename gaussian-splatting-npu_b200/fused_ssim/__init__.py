"""fused_ssim -- drop-in for the reference's fused-ssim package, on the sm_100a kernels of libgsr_b200.so.

Mirrors /root/reference/submodules/fused-ssim/fused_ssim/__init__.py:1-41 (same names, arguments,
padding modes and autograd structure): `fused_ssim(img1, img2, padding="same", train=True)`,
`FusedSSIMMap`, `allowed_padding`; the extension module keeps the reference's name `fused_ssim_cuda`
(`fusedssim`, `fusedssim_backward`).  train.py:24-28,121-122 uses it unchanged.

Addition (not in the reference): `l1_ssim_loss(image, gt, lambda_dssim)` -- the whole loss expression of
train.py:118-126, (1-l)*L1 + l*(1-SSIM), as two kernel launches (csrc/ssim.cu) instead of the L1 ops,
the SSIM map, its mean and their autograd graph.

There is no CPU fallback: a missing extension raises at import, CPU tensors raise in the binding.
"""
import torch

try:
    from fused_ssim_cuda import fusedssim, fusedssim_backward, l1_ssim_loss_forward, l1_ssim_loss_backward
except ImportError as e:  # pragma: no cover
    raise ImportError("fused_ssim_cuda (sm_100a extension) is not built: run "
                      "`python gaussian-splatting-npu_b200/build.py`; there is no fallback path") from e

allowed_padding = ["same", "valid"]


class FusedSSIMMap(torch.autograd.Function):
    @staticmethod
    def forward(ctx, C1, C2, img1, img2, padding="same", train=True):
        ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12 = fusedssim(C1, C2, img1, img2, train)

        if padding == "valid":
            ssim_map = ssim_map[:, :, 5:-5, 5:-5]

        ctx.save_for_backward(img1.detach(), img2, dm_dmu1, dm_dsigma1_sq, dm_dsigma12)
        ctx.C1 = C1
        ctx.C2 = C2
        ctx.padding = padding

        return ssim_map

    @staticmethod
    def backward(ctx, opt_grad):
        img1, img2, dm_dmu1, dm_dsigma1_sq, dm_dsigma12 = ctx.saved_tensors
        C1, C2, padding = ctx.C1, ctx.C2, ctx.padding
        dL_dmap = opt_grad
        if padding == "valid":
            dL_dmap = torch.zeros_like(img1)
            dL_dmap[:, :, 5:-5, 5:-5] = opt_grad
        grad = fusedssim_backward(C1, C2, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12)
        return None, None, grad, None, None, None


def fused_ssim(img1, img2, padding="same", train=True):
    C1 = 0.01 ** 2
    C2 = 0.03 ** 2

    assert padding in allowed_padding

    map = FusedSSIMMap.apply(C1, C2, img1, img2, padding, train)
    return map.mean()


class _L1SSIMLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, gt, lambda_dssim):
        train = image.requires_grad
        out, d0, d1, d2 = l1_ssim_loss_forward(image, gt, lambda_dssim, 0.01 ** 2, 0.03 ** 2, train)
        ctx.save_for_backward(image.detach(), gt, d0, d1, d2)
        ctx.lambda_dssim = lambda_dssim
        ctx.terms = out          # (loss, mean L1, mean ssim), device
        return out[0]

    @staticmethod
    def backward(ctx, grad_loss):
        image, gt, d0, d1, d2 = ctx.saved_tensors
        # grad_loss stays on the device: the kernel reads the scalar, no host sync
        grad = l1_ssim_loss_backward(image, gt, ctx.lambda_dssim, d0, d1, d2, grad_loss.reshape(1))
        return grad, None, None


def l1_ssim_loss(image, gt, lambda_dssim=0.2):
    """(1 - lambda) * mean|image - gt| + lambda * (1 - mean SSIM(image, gt)), padding "same"; image/gt [CH,H,W] or [B,CH,H,W]."""
    return _L1SSIMLoss.apply(image, gt, float(lambda_dssim))
