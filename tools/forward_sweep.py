"""BASELINE.json configs[4]: forward-only render sweep, 0.5M..10M Gaussians at 1920x1080 over 1000 views
(sort + blend throughput).  Device-timed through the C ABI, inputs resident; one JSON line per scene size.

    python tools/forward_sweep.py [--views 1000] [--sizes 0.5,1,2,5,10] > profiles/rNN_c5_sweep.jsonl
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-npu_b200")):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402
from gsr_b200 import capi, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--views", type=int, default=1000)
ap.add_argument("--sizes", default="0.5,1,2,5,10")
ap.add_argument("--streams", type=int, default=3, help="view v runs on CUDA stream v %% streams (one host thread)")
args = ap.parse_args()
dev = torch.device("cuda:0")
W, H = 1920, 1080
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
cams = synth.orbit_cameras(args.views, W, H, 5005)
cam_t = [(t(c.world_view_transform), t(c.full_proj_transform), t(c.camera_center), c.tanfovx, c.tanfovy) for c in cams]
bg = torch.zeros(3, device=dev)
for k, m in enumerate(float(x) for x in args.sizes.split(",")):
    P = int(m * 1e6)
    scene = synth.trained_like_scene(P, 5005 + k, s0=0.004)
    g = {n: t(scene[n]) for n in ("means3D", "opacities", "shs", "scales", "rotations")}

    def fwd(c, ws=None):
        return capi.forward(workspace=ws, width=W, height=H, viewmatrix=c[0], projmatrix=c[1], cam_pos=c[2], tan_fovx=c[3], tan_fovy=c[4],
                            background=bg, sh_degree=3, want_invdepth=False, **g)
    for c in cam_t[:400]:                  # >= 0.3 s of work: SM clocks ramp over the first few hundred ms
        fwd(c)                             # (tools/forward_gap.py: 1.50 ms/view cold vs 0.67 warm at 1M)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n_sum = 0
    main = torch.cuda.current_stream()
    streams = [main] + [torch.cuda.Stream(device=dev) for _ in range(max(1, args.streams) - 1)]
    spaces = [capi.Workspace(dev) for _ in streams]          # one live forward state per stream
    for v, c in enumerate(cam_t[:4 * len(streams)]):         # grow the workspaces outside the timed region
        with torch.cuda.stream(streams[v % len(streams)]):
            fwd(c, spaces[v % len(streams)])
    torch.cuda.synchronize()
    e0.record()
    for s_ in streams[1:]:
        s_.wait_stream(main)
    for v, c in enumerate(cam_t):
        with torch.cuda.stream(streams[v % len(streams)]):
            n_sum += fwd(c, spaces[v % len(streams)])["num_rendered"]
    for s_ in streams[1:]:
        main.wait_stream(s_)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    capi.set_profiling(True)
    stages = {}
    for c in cam_t[:8]:
        fwd(c)
        for s, v in capi.stage_times_ms().items():
            if v >= 0:
                stages.setdefault(s, []).append(v)
    capi.set_profiling(False)
    print(json.dumps({"workload": "c5: forward-only, %d views 1920x1080" % len(cam_t), "P": P, "views": len(cam_t), "streams": len(streams),
                      "views_per_s": len(cam_t) / (ms / 1e3), "ms_per_view": ms / len(cam_t),
                      "mean_num_rendered": n_sum / len(cam_t),
                      "stage_ms_median_of_8": {s: float(np.median(v)) for s, v in stages.items()}}), flush=True)
    del g, scene
    torch.cuda.empty_cache()
