"""Where does forward-only time go beyond the stage kernels?  (C5 sweep: ms/view vs sum of stage times.)"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-npu_b200")):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402
from gsr_b200 import capi, synth  # noqa: E402

P = int(float(sys.argv[1]) * 1e6) if len(sys.argv) > 1 else 1_000_000
NV = 200
dev = torch.device("cuda:0")
W, H = 1920, 1080
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
cams = synth.orbit_cameras(1000, W, H, 5005)[:NV]
cam_t = [(t(c.world_view_transform), t(c.full_proj_transform), t(c.camera_center), c.tanfovx, c.tanfovy) for c in cams]
bg = torch.zeros(3, device=dev)
scene = synth.trained_like_scene(P, 5005, s0=0.004)
g = {n: t(scene[n]) for n in ("means3D", "opacities", "shs", "scales", "rotations")}


def fwd(c):
    return capi.forward(width=W, height=H, viewmatrix=c[0], projmatrix=c[1], cam_pos=c[2], tan_fovx=c[3], tan_fovy=c[4],
                        background=bg, sh_degree=3, want_invdepth=False, **g)


for _ in range(3):                       # long warm-up: clocks ramp for the first few hundred ms
    for c in cam_t:
        fwd(c)
torch.cuda.synchronize()
# host time of one call with an idle GPU (no queueing): python + ctypes + launches + the mid-forward sync
th = []
for c in cam_t[:50]:
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    fwd(c)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    th.append((t1 - t0) * 1e3)
print("host time of one call, idle GPU: median %.3f ms (min %.3f)" % (float(np.median(th)), min(th)))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
e0.record()
for c in cam_t:
    fwd(c)
e1.record()
t_host = time.perf_counter() - t0
torch.cuda.synchronize()
print("P=%d  gpu %.3f ms/view   host loop %.3f ms/view" % (P, e0.elapsed_time(e1) / NV, t_host * 1e3 / NV))
capi.set_profiling(True)
tot = {}
for c in cam_t:
    fwd(c)
    for s, v in capi.stage_times_ms().items():
        if v >= 0:
            tot[s] = tot.get(s, 0.0) + v
capi.set_profiling(False)
print("stage sum %.3f ms/view:" % (sum(tot.values()) / NV), {s: round(v / NV, 3) for s, v in tot.items()})
