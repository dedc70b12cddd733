"""Runs the loss stage a few times on a 1080p view (for ncu captures): API fwd/bwd and the fused loss pair."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-npu_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402
from gsr_b200 import capi  # noqa: E402

H, W = 1080, 1920
dev = torch.device("cuda:0")
torch.manual_seed(3)
img = torch.rand(3, H, W, device=dev)
gt = (img + 0.05 * torch.randn_like(img)).clamp(0, 1)
sc = capi.LossScratch((3, H, W), dev)
dl = torch.randn_like(img)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    capi.l1_ssim_loss_forward(img, gt, 0.2, sc)
    capi.l1_ssim_loss_backward(img, gt, 0.2, sc)
    m, d0, d1, d2 = capi.ssim_forward(img, gt, True)
    capi.ssim_backward(img, gt, dl, d0, d1, d2)
torch.cuda.synchronize()
print("ok", sc.loss.tolist())
