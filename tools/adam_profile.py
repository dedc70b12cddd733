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
vis = torch.rand(P, device=dev) < 0.43
for mode in ("dense", "sparse"):
    ts = []
    for _ in range(6):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        opt.step(visible=vis if mode == "sparse" else None, zero_grad=True)
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    t = sorted(ts)[len(ts) // 2]
    rows = int(vis.sum()) if mode == "sparse" else P
    print("%s: %.3f ms  %.0f GB/s (32 B x 59 x %d rows)" % (mode, t, 32 * 59 * rows / t / 1e6, rows))
