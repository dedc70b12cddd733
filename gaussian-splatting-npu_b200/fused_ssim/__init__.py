"""fused_ssim -- drop-in for the reference's fused-ssim package, on the sm_100a kernels of libgsr_b200.so.

Same public surface as /root/reference/submodules/fused-ssim/fused_ssim/__init__.py:1-41 (names, arguments,
padding modes, autograd slots -- the API a caller binds to), bodies restated: `fused_ssim(img1, img2, padding="same", train=True)`,
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
_BORDER = 5          # half of the 11-tap window: what padding="valid" crops on every side
_SSIM_C1, _SSIM_C2 = 0.01 ** 2, 0.03 ** 2


def _crop(t):
    return t[:, :, _BORDER:-_BORDER, _BORDER:-_BORDER]


class FusedSSIMMap(torch.autograd.Function):
    """Autograd node over the two extension entry points.  Its signature -- apply(C1, C2, img1, img2, padding, train)
    returning the SSIM map, six backward slots with the gradient in the third -- is the reference's
    (fused_ssim/__init__.py:8-31) because callers may use the class directly; the body is restated here."""

    @staticmethod
    def forward(ctx, C1, C2, img1, img2, padding="same", train=True):
        full_map, d_mu1, d_sigma1_sq, d_sigma12 = fusedssim(C1, C2, img1, img2, train)
        ctx.consts = (C1, C2, padding)
        ctx.save_for_backward(img1.detach(), img2, d_mu1, d_sigma1_sq, d_sigma12)
        return _crop(full_map) if padding == "valid" else full_map

    @staticmethod
    def backward(ctx, grad_map):
        C1, C2, padding = ctx.consts
        img1, img2, d_mu1, d_sigma1_sq, d_sigma12 = ctx.saved_tensors
        if padding == "valid":          # the kernel wants a full-size upstream map: zero outside the cropped window
            full = torch.zeros_like(img1)
            _crop(full).copy_(grad_map)
            grad_map = full
        d_img1 = fusedssim_backward(C1, C2, img1, img2, grad_map, d_mu1, d_sigma1_sq, d_sigma12)
        return (None, None, d_img1, None, None, None)


def fused_ssim(img1, img2, padding="same", train=True):
    """Mean SSIM of two [B,CH,H,W] images (reference signature, fused_ssim/__init__.py:33-41)."""
    assert padding in allowed_padding
    return FusedSSIMMap.apply(_SSIM_C1, _SSIM_C2, img1, img2, padding, train).mean()


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
