"""CPU tests of the loss-stage oracle (oracle/ssim_oracle.c).

Pin 1: the reference's own test recipe for fused-ssim
(/root/reference/submodules/fused-ssim/tests/test.py:14-55,66-91): a torch conv2d SSIM
(`ssim()`, identical to /root/reference/utils/loss_utils.py:43-78) compared with torch.isclose on the
value and on the gradient, padding "same"; "valid" is compared with the same conv2d map cropped by 5
(the reference compares it with pytorch_msssim, which is not installed here: same definition).
Pin 2 (tests/test_ssim_golden.py): outputs of the unmodified reference CUDA kernel run on a B200.
"""
from math import exp

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import ssim_oracle as so


def _window(channel):
    g = torch.Tensor([exp(-(x - 11 // 2) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)])
    g = (g / g.sum()).unsqueeze(1)
    w2 = g.mm(g.t()).float().unsqueeze(0).unsqueeze(0)
    return w2.expand(channel, 1, 11, 11).contiguous()


def torch_ssim_map(img1, img2):
    """The map of /root/reference/utils/loss_utils.py:54-73 (before the mean)."""
    ch = img1.size(-3)
    w = _window(ch).to(img1.dtype).to(img1.device)
    mu1 = F.conv2d(img1, w, padding=5, groups=ch)
    mu2 = F.conv2d(img2, w, padding=5, groups=ch)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1 = F.conv2d(img1 * img1, w, padding=5, groups=ch) - mu1_sq
    s2 = F.conv2d(img2 * img2, w, padding=5, groups=ch) - mu2_sq
    s12 = F.conv2d(img1 * img2, w, padding=5, groups=ch) - mu1_mu2
    return ((2 * mu1_mu2 + so.C1) * (2 * s12 + so.C2)) / ((mu1_sq + mu2_sq + so.C1) * (s1 + s2 + so.C2))


def test_window_matches_reference_literals():
    # FS/ssim.cu:10-20 holds the fp32 window of FS/tests/test.py:14-16
    g = torch.Tensor([exp(-(x - 5) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)])
    g = (g / g.sum()).numpy()
    lit = np.array([0.001028380123898387, 0.0075987582094967365, 0.036000773310661316, 0.10936068743467331,
                    0.21300552785396576, 0.26601171493530273, 0.21300552785396576, 0.10936068743467331,
                    0.036000773310661316, 0.0075987582094967365, 0.001028380123898387], np.float32)
    assert np.allclose(g, lit, rtol=0, atol=1e-8)


@pytest.mark.parametrize("shape", [(1, 3, 64, 64), (2, 5, 37, 75), (1, 1, 11, 11), (1, 3, 7, 5), (1, 3, 96, 160)])
def test_oracle_vs_torch_conv2d_same(shape):
    torch.manual_seed(sum(shape))
    img1 = torch.rand(shape, dtype=torch.float64, requires_grad=True)
    img2 = torch.rand(shape, dtype=torch.float64)
    ref_map = torch_ssim_map(img1, img2)
    ref_val = ref_map.mean()
    ref_val.backward()
    m, d0, d1, d2 = so.fusedssim(so.C1, so.C2, img1.detach().numpy(), img2.numpy(), True)
    assert np.abs(m - ref_map.detach().numpy()).max() <= 2e-5
    val, g = so.fused_ssim(img1.detach().numpy(), img2.numpy(), "same")
    # the reference's own criterion (torch.isclose defaults: rtol 1e-5, atol 1e-8) on the value ...
    assert torch.isclose(torch.tensor(val, dtype=torch.float32), ref_val.float())
    # ... and on the gradient, normalised (gradients of a mean are ~1e-6: compare relative to the largest)
    gr = img1.grad.numpy()
    assert np.abs(g - gr).max() <= 1e-4 * np.abs(gr).max()


def test_oracle_valid_padding():
    torch.manual_seed(7)
    img1 = torch.rand((1, 3, 48, 40), dtype=torch.float64, requires_grad=True)
    img2 = torch.rand((1, 3, 48, 40), dtype=torch.float64)
    ref_val = torch_ssim_map(img1, img2)[:, :, 5:-5, 5:-5].mean()
    ref_val.backward()
    val, g = so.fused_ssim(img1.detach().numpy(), img2.numpy(), "valid")
    assert abs(val - ref_val.item()) <= 1e-6
    gr = img1.grad.numpy()
    assert np.abs(g - gr).max() <= 1e-4 * np.abs(gr).max()


def test_identical_images_give_one():
    rng = np.random.default_rng(3)
    a = rng.random((1, 3, 40, 56), dtype=np.float32)
    m, *_ = so.fusedssim(so.C1, so.C2, a, a.copy(), False)
    assert np.abs(m - 1.0).max() <= 1e-5


def test_l1_ssim_loss_matches_train_py_formula():
    # train.py:118-126: loss = (1-l)*L1 + l*(1-ssim); gradient against torch autograd in float64
    torch.manual_seed(11)
    img = torch.rand((3, 50, 70), dtype=torch.float64, requires_grad=True)
    gt = torch.rand((3, 50, 70), dtype=torch.float64)
    lam = 0.2
    l1 = torch.abs(img - gt).mean()
    ss = torch_ssim_map(img.unsqueeze(0), gt.unsqueeze(0)).mean()
    loss = (1.0 - lam) * l1 + lam * (1.0 - ss)
    (loss * 3.0).backward()
    r = so.l1_ssim_loss(img.detach().numpy(), gt.numpy(), lam, upstream=3.0)
    assert abs(r["loss"] - loss.item()) <= 1e-6
    assert abs(r["l1"] - l1.item()) <= 1e-6 and abs(r["ssim"] - ss.item()) <= 1e-6
    gr = img.grad.numpy()
    assert np.abs(r["dL_dimg"] - gr).max() <= 1e-4 * np.abs(gr).max()


def test_backward_is_linear_in_dL_dmap():
    rng = np.random.default_rng(5)
    a = rng.random((1, 2, 33, 47), dtype=np.float32)
    b = rng.random((1, 2, 33, 47), dtype=np.float32)
    m, d0, d1, d2 = so.fusedssim(so.C1, so.C2, a, b, True)
    u = rng.standard_normal(a.shape).astype(np.float32)
    v = rng.standard_normal(a.shape).astype(np.float32)
    gu = so.fusedssim_backward(so.C1, so.C2, a, b, u, d0, d1, d2)
    gv = so.fusedssim_backward(so.C1, so.C2, a, b, v, d0, d1, d2)
    guv = so.fusedssim_backward(so.C1, so.C2, a, b, u + v, d0, d1, d2)
    assert np.abs(guv - (gu + gv)).max() <= 1e-4 * np.abs(guv).max()
