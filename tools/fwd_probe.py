"""Forward + backward of one view back to back, device-timed, for the library named by GSR_B200_LIB and the
environment as given (GSR_COLOR_INLINE=1 keeps the fused preprocess kernel):  python tools/fwd_probe.py [workload]"""
import json
import os
import sys

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
ws = capi.Workspace(dev)
st = capi.forward(width=W, height=H, workspace=ws, **kw)
out = capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, **kw)
res = {"workload": wl, "lib": os.path.basename(capi.LIB_PATH), "GSR_COLOR_INLINE": os.environ.get("GSR_COLOR_INLINE")}
for mode in ("fwd", "fwd+bwd"):
    for _ in range(30):
        st = capi.forward(width=W, height=H, workspace=ws, **kw)
        if mode != "fwd":
            capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, out=out, **kw)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(100):
        st = capi.forward(width=W, height=H, workspace=ws, **kw)
        if mode != "fwd":
            capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, out=out, **kw)
    b.record()
    torch.cuda.synchronize()
    res[mode + "_ms"] = a.elapsed_time(b) / 100
capi.set_profiling(True)
acc = {}
for _ in range(9):
    st = capi.forward(width=W, height=H, workspace=ws, **kw)
    for k, v in capi.stage_times_ms().items():
        acc.setdefault(k, []).append(v)
capi.set_profiling(False)
res["stage_ms"] = {k: round(float(np.median(v)), 4) for k, v in acc.items() if k in ("preprocess", "sh_color", "scan", "depth_sort", "duplicate", "sort", "ranges", "render_fwd")}
res["color_sum"] = float(st["out_color"].double().sum().item())
print(json.dumps(res))
