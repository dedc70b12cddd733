"""Minimal driver for ncu: the C2 workload's forward+backward through the C ABI, nothing else.
    python tools/profile_step.py [--steps 2] [--workload c2]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-npu_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import torch

import bench
from gsr_b200 import capi, synth
from gsr_b200.dist import FlatGaussians

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=2)
ap.add_argument("--workload", default="c2")
a = ap.parse_args()
w, scene, cam = bench.make_workload(a.workload)
dev = torch.device("cuda:0")
G = FlatGaussians.from_numpy(scene, dev)
t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)
dcol, ddep = (t(x) for x in synth.upstream_grads(w["H"], w["W"], 5))
common = dict(means3D=G.xyz, opacities=G.opacity, viewmatrix=t(cam.world_view_transform), projmatrix=t(cam.full_proj_transform),
              cam_pos=t(cam.camera_center), background=torch.zeros(3, device=dev), tan_fovx=cam.tanfovx, tan_fovy=cam.tanfovy,
              shs=G.shs, scales=G.scales, rotations=G.rotations, sh_degree=3, antialiasing=w["aa"])
for _ in range(a.steps):
    st = capi.forward(width=w["W"], height=w["H"], **common)
    g = capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, out=G.grad_views(), **common)
torch.cuda.synchronize()
print("ok N=%d" % st["num_rendered"])
