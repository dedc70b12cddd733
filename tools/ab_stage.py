"""A/B of kernel variants on one workload: per-stage device times (library profiling mode) and gradient agreement.

    python tools/ab_stage.py [--workload c2] libA.so libB.so ...      (first library = the baseline of the comparison)

Each library runs in its own process (GSR_B200_LIB); gradients of the first are kept in /tmp and compared with the others.
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import json, os, sys
ROOT = sys.argv[1]
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-npu_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np, torch
import bench
from gsr_b200 import capi, synth
wl, tag = sys.argv[2], sys.argv[3]
w, scene, cam = bench.make_workload(wl)
dev = torch.device("cuda:0")
t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)
dcol, ddep = (t(x) for x in synth.upstream_grads(w["H"], w["W"], 5))
kw = dict(means3D=t(scene["means3D"]), opacities=t(scene["opacities"]), viewmatrix=t(cam.world_view_transform),
          projmatrix=t(cam.full_proj_transform), cam_pos=t(cam.camera_center), background=torch.zeros(3, device=dev),
          tan_fovx=cam.tanfovx, tan_fovy=cam.tanfovy, shs=t(scene["shs"]), scales=t(scene["scales"]), rotations=t(scene["rotations"]),
          sh_degree=3, antialiasing=w["aa"])
for _ in range(30):
    st = capi.forward(width=w["W"], height=w["H"], **kw); g = capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, **kw)
torch.cuda.synchronize()
capi.set_profiling(True)
acc = {}
for _ in range(15):
    st = capi.forward(width=w["W"], height=w["H"], **kw)
    f = capi.stage_times_ms()
    g = capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, **kw)
    for k, v in capi.stage_times_ms().items():
        acc.setdefault(k, []).append(v)
capi.set_profiling(False)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
torch.cuda.synchronize(); ev[0].record()
for _ in range(20):
    st = capi.forward(width=w["W"], height=w["H"], **kw); g = capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, **kw)
ev[1].record(); torch.cuda.synchronize()
out = {k: float(np.median(v)) for k, v in acc.items()}
out["fwd_bwd_ms_back_to_back"] = ev[0].elapsed_time(ev[1]) / 20
keys = ["dL_dmeans2D", "dL_dconic", "dL_dopacity", "dL_dcolors", "dL_dinvdepths", "dL_dmeans3D", "dL_dscales", "dL_drotations"]
np.savez("/tmp/ab_%s.npz" % tag, **{k: g[k].cpu().numpy() for k in keys}, out_color=st["out_color"].cpu().numpy())
print(json.dumps(out))
'''


def main():
    args = sys.argv[1:]
    wl = "c2"
    if args and args[0] == "--workload":
        wl, args = args[1], args[2:]
    import numpy as np
    base = None
    for i, lib in enumerate(args):
        env = dict(os.environ, GSR_B200_LIB=os.path.abspath(lib))
        r = subprocess.run([sys.executable, "-c", CHILD, ROOT, wl, str(i)], env=env, capture_output=True, text=True)
        if r.returncode != 0:
            print(json.dumps({"lib": lib, "error": r.stderr[-1500:]}))
            continue
        res = {"lib": os.path.basename(lib), "workload": wl, "stage_ms": json.loads(r.stdout.strip().splitlines()[-1])}
        cur = np.load("/tmp/ab_%d.npz" % i)
        if base is None:
            base = cur
        else:
            res["vs_first"] = {k: float(np.abs(cur[k].astype(np.float64) - base[k]).max() / (np.abs(base[k]).max() + 1e-30)) for k in cur.files}
        print(json.dumps(res))
        sys.stdout.flush()


if __name__ == "__main__":
    main()
