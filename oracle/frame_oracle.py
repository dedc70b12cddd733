"""CPU oracle of the frame output (TEST INFRASTRUCTURE ONLY: tests/, smoke(), bench cpu legs).

numpy restatement of
  /root/reference/train.py:81       (torch.clamp(img, 0, 1) * 255).byte().permute(1, 2, 0)             mode "truncate"
  /root/reference/render.py:44-45   torchvision save_image: img.mul(255).add_(0.5).clamp_(0, 255).permute(1, 2, 0).to(uint8)
                                    (and :40-42, the right half of the image under train_test_exp)     mode "round"
fp32 arithmetic with one rounding per operation, as torch's elementwise kernels do.  Pinned against those torch
expressions (and torchvision + PIL for the PNG) in tests/test_frames_cpu.py.  NaN -> 0 (torch leaves it undefined)."""
import numpy as np

F = np.float32


def to_rgb8(image, mode="truncate", x_begin=0, x_end=None):
    x = np.asarray(image, F)[:, :, x_begin:x_end]
    nan = np.isnan(x)
    x = np.where(nan, F(0), x)
    if mode == "truncate":
        y = np.clip(x, F(0), F(1)) * F(255)
    elif mode == "round":
        y = np.clip((x * F(255)).astype(F) + F(0.5), F(0), F(255))
    else:
        raise ValueError(mode)
    return np.ascontiguousarray(y.astype(np.uint8).transpose(1, 2, 0))
