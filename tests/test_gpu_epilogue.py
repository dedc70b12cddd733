"""GPU parity of the render epilogue (csrc/epilogue.cu) through the C ABI against oracle/epilogue_oracle.py and the
reference's torch lines (/root/reference/gaussian_renderer/__init__.py:112-119) run on this GPU.
Tolerance: image 1e-6 max-abs (three fp32 products + bias; the kernel may contract them into FMAs), dL_dimage 1e-5,
dL_dexposure 1e-4 relative (north_star's gradient bound); the clamp mask is compared away from values within 1e-5 of
a bound, where fp32 contraction may legitimately flip it."""
import numpy as np
import pytest

from oracle import epilogue_oracle as eo
from test_epilogue_cpu import make, reference_lines

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("H,W", [(1, 1), (7, 13), (545, 979), (1080, 1920)])
@pytest.mark.parametrize("trained", [False, True])
def test_exposure_clamp_vs_oracle_and_torch(cuda, H, W, trained):
    import torch
    from gsr_b200.epilogue import exposure_clamp
    img, E, dL = make(H, W, H + W)
    ti = torch.from_numpy(img).to(cuda).requires_grad_(True)
    te = torch.from_numpy(E).to(cuda).requires_grad_(True)
    out = exposure_clamp(ti, te if trained else None)
    out.backward(torch.from_numpy(dL).to(cuda))
    y, pre = eo.forward(img, E if trained else None)
    assert np.abs(out.detach().cpu().numpy() - y).max() <= 1e-6
    safe = (np.abs(pre) > 1e-5) & (np.abs(pre - 1) > 1e-5)
    if trained:
        safe = np.broadcast_to(safe.all(axis=0), safe.shape)
    dx, dE = eo.backward(img, E if trained else None, dL)
    assert np.abs(ti.grad.cpu().numpy() - dx)[safe].max(initial=0.0) <= 1e-5
    # the reference's own lines on this GPU
    ri = torch.from_numpy(img).to(cuda).requires_grad_(True)
    re_ = torch.from_numpy(E).to(cuda).requires_grad_(True)
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        rout = reference_lines(ri, re_, trained)
        rout.backward(torch.from_numpy(dL).to(cuda))
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    assert (out.detach() - rout.detach()).abs().max().item() <= 1e-6
    assert (ti.grad - ri.grad).abs().cpu().numpy()[safe].max(initial=0.0) <= 1e-5
    if trained:
        n_unsafe = int((~safe).sum())
        scale = max(1.0, float(np.abs(dE).max()))
        slack = 1e-4 * scale + n_unsafe * 4.0 * float(np.abs(dL).max())      # a flipped mask moves one pixel's term
        assert np.abs(te.grad.cpu().numpy() - dE).max() <= slack
        assert (te.grad - re_.grad).abs().max().item() <= slack
    else:
        assert te.grad is None
        assert torch.equal(ti.grad[:, 0, 0], torch.from_numpy(dL[:, 0, 0]).to(cuda))   # gradient passes ON the bounds


def test_deterministic_and_image_only_grad(cuda):
    import torch
    from gsr_b200 import capi
    img, E, dL = make(300, 500, 9)
    t = [torch.from_numpy(a).to(cuda) for a in (img, E, dL)]
    a = capi.exposure_clamp_backward(*t)
    b = capi.exposure_clamp_backward(*t)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    c = capi.exposure_clamp_backward(*t, need_exposure_grad=False)
    assert c[1] is None and torch.equal(c[0], a[0])


def test_error_paths(cuda):
    import torch
    from gsr_b200 import capi
    L = capi.lib()
    z = torch.zeros(3, 4, 4, device=cuda)
    assert L.gsr_exposure_clamp_forward(-1, 4, None, None, None, None) != 0 and "negative" in capi.last_error()
    assert L.gsr_exposure_clamp_forward(4, 4, None, None, None, None) != 0
    assert L.gsr_exposure_clamp_forward(0, 4, None, None, None, None) == 0
    e = torch.zeros(3, 4, device=cuda)
    rc = L.gsr_exposure_clamp_backward(4, 4, z.data_ptr(), None, z.data_ptr(), z.data_ptr(), e.data_ptr(), None, None)
    assert rc != 0 and "dL_dexposure needs" in capi.last_error()
