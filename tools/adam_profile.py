"""Runs the Adam step on a C2-sized flat bucket a few times (for ncu captures and A/B timing)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-npu_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402
from gsr_b200.dist import FlatGaussians  # noqa: E402
from gsr_b200.optim import GaussianAdam  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 3_000_000
dev = torch.device("cuda:0")
G = FlatGaussians(P, dev)
G.params.normal_()
G.bucket.normal_()
opt = GaussianAdam(G)
STATE = sys.argv[2] if len(sys.argv) > 2 else "random"   # random | zero (untouched rows: g = m = v = 0) | denormal (decayed m)
if STATE == "zero":
    G.bucket.zero_()
elif STATE == "denormal":
    G.bucket.zero_()
    opt.exp_avg.fill_(1e-41)
    opt.exp_avg_sq.fill_(1e-24)
elif STATE == "half":                                      # 57 % of the rows never seen, the rest with gradients
    opt.step(zero_grad=False)
    rows = (torch.rand(P, device=dev) < 0.57).nonzero().squeeze(1)
    for buf in (G.bucket, opt.exp_avg, opt.exp_avg_sq):
        for name in ("xyz", "shs", "opacity", "scales", "rotations"):
            G._view(buf, name)[rows] = 0
vis = torch.rand(P, device=dev) < 0.43
for mode in ("dense", "sparse"):
    ts = []
    for _ in range(6):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        opt.step(visible=vis if mode == "sparse" else None, zero_grad=(STATE == "random"))
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    t = sorted(ts)[len(ts) // 2]
    rows = int(vis.sum()) if mode == "sparse" else P
    print("%s: %.3f ms  %.0f GB/s (32 B x 59 x %d rows)" % (mode, t, 32 * 59 * rows / t / 1e6, rows))
