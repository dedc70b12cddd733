"""The epilogue oracle (oracle/epilogue_oracle.py) pinned against the reference's own torch lines
(/root/reference/gaussian_renderer/__init__.py:112-119, restated verbatim below) and torch autograd, on CPU."""
import numpy as np
import pytest
import torch

from oracle import epilogue_oracle as eo


def reference_lines(rendered_image, exposure, use_trained_exp):
    if use_trained_exp:
        rendered_image = torch.matmul(rendered_image.permute(1, 2, 0), exposure[:3, :3]).permute(2, 0, 1) + exposure[:3, 3, None, None]
    return rendered_image.clamp(0, 1)


def make(H, W, seed):
    rng = np.random.default_rng(seed)
    img = (rng.random((3, H, W)) * 1.6 - 0.3).astype(np.float32)          # values on both sides of both clamp bounds
    img[:, 0, 0] = (0.0, 1.0, 0.5)                                        # exactly on the bounds
    E = (np.eye(3, 4) + rng.standard_normal((3, 4)) * 0.1).astype(np.float32)
    dL = rng.standard_normal((3, H, W)).astype(np.float32)
    return img, E, dL


@pytest.mark.parametrize("H,W", [(1, 1), (7, 13), (64, 96)])
@pytest.mark.parametrize("trained", [False, True])
def test_oracle_vs_reference_torch_expression(H, W, trained):
    img, E, dL = make(H, W, H * W)
    ti = torch.from_numpy(img).requires_grad_(True)
    te = torch.from_numpy(E).requires_grad_(True)
    out = reference_lines(ti, te, trained)
    out.backward(torch.from_numpy(dL))
    y, _ = eo.forward(img, E if trained else None)
    assert np.abs(y - out.detach().numpy()).max() <= 1e-6
    dx, dE = eo.backward(img, E if trained else None, dL)
    # a pixel whose unclamped value is within rounding of a bound may fall on either side of the mask: compare off them
    with torch.no_grad():
        pre = (torch.matmul(ti.permute(1, 2, 0), te[:3, :3]).permute(2, 0, 1) + te[:3, 3, None, None]) if trained else ti.detach()
    safe = ((pre.abs() > 1e-5) & ((pre - 1).abs() > 1e-5)).numpy()
    if trained:
        safe = np.broadcast_to(safe.all(axis=0), safe.shape)
    assert np.abs(dx - ti.grad.numpy())[safe].max(initial=0.0) <= 1e-5
    if trained:
        if safe.all():
            assert np.abs(dE - te.grad.numpy()).max() <= 1e-4 * max(1.0, np.abs(te.grad.numpy()).max())
    else:
        assert dE is None
        assert np.array_equal(dx[:, 0, 0], dL[:, 0, 0])                      # torch.clamp passes gradient ON the bounds
