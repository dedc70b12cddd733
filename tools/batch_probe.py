"""One batched step (gsr_forward_batch + gsr_backward_batch, V views of a workload) a few times: driver for ncu / timing.
    python tools/batch_probe.py [workload] [V]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-npu_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import torch
import bench

wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
V = int(sys.argv[2]) if len(sys.argv) > 2 else 8
dev = torch.device("cuda:0")
torch.cuda.set_device(0)
job = bench.Job(wl, V, V, 0, 1, dev, "nccl", max_streams=3)
for name, fn in (("single", lambda: job.step_render(1)), ("streams3", lambda: job.step_render(3)), ("batched", lambda: job.step_batched(1)), ("batched_streams3", lambda: job.step_batched(3))):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        fn()
    b.record()
    torch.cuda.synchronize()
    print(json.dumps({"workload": wl, "V": V, "mode": name, "ms_per_step": a.elapsed_time(b) / 10}))
