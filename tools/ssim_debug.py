import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-npu_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np, torch
from gsr_b200 import capi
from oracle import ssim_oracle as so
H, W = int(sys.argv[1]), int(sys.argv[2])
torch.manual_seed(0)
a = torch.rand(1, 3, H, W, device="cuda"); b = torch.rand(1, 3, H, W, device="cuda")
m, d0, d1, d2 = capi.ssim_forward(a, b, True)
torch.cuda.synchronize()
om, *_ = so.fusedssim(so.C1, so.C2, a.cpu().numpy(), b.cpu().numpy(), True)
print("NO_TMA=%s %dx%d equal=%s" % (os.environ.get("GSR_SSIM_NO_TMA"), H, W, np.array_equal(m.cpu().numpy(), om)))
