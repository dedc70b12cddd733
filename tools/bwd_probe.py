"""Where does a single-stream forward+backward spend its time?  Per-call CUDA-event and host times in several calling
modes (fresh scratch vs capi.Workspace, fresh gradient tensors vs out=), for the library named by GSR_B200_LIB.

    GSR_B200_LIB=path/to/lib.so python tools/bwd_probe.py [workload]
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-npu_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import torch
import bench
from gsr_b200 import capi, synth

wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
w, scene, cam = bench.make_workload(wl)
dev = torch.device("cuda:0")
t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)
dcol, ddep = (t(x) for x in synth.upstream_grads(w["H"], w["W"], 5))
kw = dict(means3D=t(scene["means3D"]), opacities=t(scene["opacities"]), viewmatrix=t(cam.world_view_transform),
          projmatrix=t(cam.full_proj_transform), cam_pos=t(cam.camera_center), background=torch.zeros(3, device=dev),
          tan_fovx=cam.tanfovx, tan_fovy=cam.tanfovy, shs=t(scene["shs"]), scales=t(scene["scales"]), rotations=t(scene["rotations"]),
          sh_degree=3, antialiasing=w["aa"])
W, H = w["W"], w["H"]


def ev():
    return torch.cuda.Event(enable_timing=True)


def run(mode, iters=20, sync_each=False):
    ws = capi.Workspace(dev) if "ws" in mode else None
    out = None
    if "out" in mode:
        st = capi.forward(width=W, height=H, **kw)
        out = capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, **kw)
    for _ in range(10):
        st = capi.forward(width=W, height=H, workspace=ws, **kw)
        capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, out=out, **kw)
    torch.cuda.synchronize()
    marks, host = [], []
    t_all0 = time.perf_counter()
    e_all0, e_all1 = ev(), ev()
    e_all0.record()
    for _ in range(iters):
        a, b, c = ev(), ev(), ev()
        h0 = time.perf_counter()
        a.record()
        st = capi.forward(width=W, height=H, workspace=ws, **kw)
        h1 = time.perf_counter()
        b.record()
        capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, out=out, **kw)
        c.record()
        h2 = time.perf_counter()
        if sync_each:
            torch.cuda.synchronize()
        marks.append((a, b, c))
        host.append((h1 - h0, h2 - h1))
    e_all1.record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t_all0) / iters * 1e3
    f = [a.elapsed_time(b) for a, b, c in marks]
    bw = [b.elapsed_time(c) for a, b, c in marks]
    return {"mode": mode, "sync_each": sync_each, "per_iter_ms": e_all0.elapsed_time(e_all1) / iters, "wall_ms": wall,
            "fwd_ms_med": float(np.median(f)), "bwd_ms_med": float(np.median(bw)), "bwd_ms_max": float(np.max(bw)),
            "host_fwd_ms": float(np.median([h[0] for h in host])) * 1e3, "host_bwd_ms": float(np.median([h[1] for h in host])) * 1e3}


def bwd_only(iters=20):
    st = capi.forward(width=W, height=H, **kw)
    out = capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, **kw)
    torch.cuda.synchronize()
    a, b = ev(), ev()
    a.record()
    for _ in range(iters):
        capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, out=out, **kw)
    b.record()
    torch.cuda.synchronize()
    return {"mode": "bwd only, same state, out=", "per_iter_ms": a.elapsed_time(b) / iters}


print(json.dumps({"lib": os.path.basename(capi.LIB_PATH), "workload": wl}))
for mode, se in (("fresh", False), ("fresh", True), ("ws", False), ("ws+out", False), ("ws+out", True), ("out", False)):
    print(json.dumps(run(mode, sync_each=se)))
    sys.stdout.flush()
print(json.dumps(bwd_only()))
