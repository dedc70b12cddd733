"""GPU parity of the frame output (csrc/frame_out.cu) through the C ABI: bit-exact against oracle/frame_oracle.py and
against the reference's torch expressions run on this GPU (train.py:81; torchvision save_image as in render.py:44)."""
import io

import numpy as np
import pytest

from oracle import frame_oracle as fo
from test_frames_cpu import make

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("H,W", [(1, 1), (5, 7), (33, 130), (546, 979), (1080, 1920)])
def test_to_rgb8_bit_exact(cuda, H, W):
    import torch
    from gsr_b200 import frames
    img = make(H, W, H + W)
    if H * W > 2000:
        img[1, 3, 5] = np.nan
    t = torch.from_numpy(img).to(cuda)
    for mode, name in ((frames.TRUNCATE, "truncate"), (frames.ROUND, "round")):
        got = frames.to_rgb8(t, mode)
        assert got.shape == (H, W, 3) and got.dtype == torch.uint8
        assert np.array_equal(got.cpu().numpy(), fo.to_rgb8(img, name)), name
    clean = torch.nan_to_num(t, nan=0.0)
    assert torch.equal(frames.to_rgb8(clean, frames.TRUNCATE), (torch.clamp(clean, min=0, max=1.0) * 255).byte().permute(1, 2, 0).contiguous())
    assert torch.equal(frames.to_rgb8(clean, frames.ROUND), clean.clone().mul(255).add_(0.5).clamp_(0, 255).permute(1, 2, 0).to(torch.uint8))
    for x0, x1 in ((W // 2, W), (1, W), (0, 0), (W // 3, W - W // 5)):                  # crops, incl. unaligned and empty
        assert np.array_equal(frames.to_rgb8(t, frames.ROUND, x0, x1).cpu().numpy(), fo.to_rgb8(img, "round", x0, x1))


def test_unaligned_base_pointer(cuda):
    import torch
    from gsr_b200 import frames
    img = make(16, 64, 3)
    big = torch.zeros(3 * 16 * 64 + 1, device=cuda)
    big[1:] = torch.from_numpy(img).reshape(-1).to(cuda)
    view = big[1:].view(3, 16, 64)                                                    # 4-byte aligned only
    assert np.array_equal(frames.to_rgb8(view, frames.TRUNCATE).cpu().numpy(), fo.to_rgb8(img, "truncate"))


def test_frame_bytes_and_png_of_a_render(cuda, tmp_path):
    """A render through the product path -> viewer bytes and PNG file; the PNG equals torchvision's for the same render."""
    import torch
    import cases
    from gsr_b200 import frames
    PIL = pytest.importorskip("PIL.Image")
    tv = pytest.importorskip("torchvision.utils")
    scene, cam, opts = cases.get_case("odd167x101")
    color = torch.from_numpy(cases.run_cuda(scene, cam, opts)["out_color"]).to(cuda)
    mv = frames.frame_bytes(color)
    assert mv.nbytes == cam.image_height * cam.image_width * 3
    assert np.array_equal(np.asarray(mv), fo.to_rgb8(color.cpu().numpy(), "truncate"))
    for half in (False, True):
        p, q = str(tmp_path / "ours.png"), str(tmp_path / "ref.png")
        frames.save_png(color, p, right_half=half)
        tv.save_image(color[..., color.shape[-1] // 2:] if half else color, q)       # render.py:40-45
        assert np.array_equal(np.array(PIL.open(p).convert("RGB")), np.array(PIL.open(q).convert("RGB")))


def test_error_paths(cuda):
    import torch
    from gsr_b200 import capi, frames
    t = torch.zeros(3, 8, 8, device=cuda)
    with pytest.raises(RuntimeError, match="column range"):
        frames.to_rgb8(t, frames.ROUND, 4, 9)
    with pytest.raises(RuntimeError, match="unknown mode"):
        frames.to_rgb8(t, 7)
