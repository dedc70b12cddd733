"""Measurement probe: what the warps of the blend kernels meet on a workload (visits, contributing lanes, tails).

    python tools/blend_stats.py [c2|c3|c4] > gpurun_out/blend_stats_c2.json
"""
import ctypes
import json
import os
import sys
from ctypes import c_int, c_void_p

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-npu_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import bench  # noqa: E402
import cases  # noqa: E402
from gsr_b200 import capi  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c2"
w, scene, cam = bench.make_workload(name)
kw = cases._to_dev(cases.inputs_of(scene, cam, dict(bg=(0, 0, 0), antialiasing=w["aa"], mode="sh")))
st = capi.forward(width=w["W"], height=w["H"], **kw)
torch.cuda.synchronize()
it = capi.introspect(st)
L = capi.lib()
bl = capi.BinningLayout()
L.gsr_get_binning_layout(st["num_rendered"], w["W"], w["H"], ctypes.byref(bl))
gl = capi.GeometryLayout()
L.gsr_get_geometry_layout(st["P"], ctypes.byref(gl))
masks = capi._view(st["binning"], bl.tile_keys_alt, st["num_rendered"], torch.uint8)
records = capi._view(st["geom"], gl.records, st["P"] * 12, torch.float32)
probe = ctypes.CDLL(os.path.join(ROOT, "tools", "probe", "blend_stats.so"))
probe.blend_stats.argtypes = [c_void_p] * 4 + [c_int, c_int] + [c_void_p] * 4
cnt = torch.zeros(128, dtype=torch.int64, device="cuda")
rc = probe.blend_stats(it["ranges"].data_ptr(), it["point_list"].data_ptr(), masks.data_ptr(), records.data_ptr(), w["W"], w["H"],
                       it["final_T"].data_ptr(), it["n_contrib"].data_ptr(), cnt.data_ptr(), None)
torch.cuda.synchronize()
assert rc == 0, rc
c = cnt.cpu().numpy()
r = it["ranges"].to(torch.int64)
length = (r[:, 1] - r[:, 0]).cpu().numpy()
gx, gy = (w["W"] + 15) // 16, (w["H"] + 15) // 16
nc = it["n_contrib"].to(torch.int64)
Hp, Wp = gy * 16, gx * 16
ncp = torch.zeros(Hp, Wp, dtype=torch.int64, device="cuda")
ncp[:w["H"], :w["W"]] = nc
tmax = ncp.view(gy, 16, gx, 16).permute(0, 2, 1, 3).reshape(gy * gx, 256).max(dim=1).values.cpu().numpy()
pop = np.array([bin(m).count("1") for m in range(256)])
mask_np = masks.cpu().numpy()
out = {
    "workload": name, "P": st["P"], "P_vis": int((st["radii"] > 0).sum()), "N": st["num_rendered"], "tiles": int(gx * gy),
    "tile_list_len": {"mean": float(length.mean()), "p50": float(np.percentile(length, 50)), "p90": float(np.percentile(length, 90)),
                      "p99": float(np.percentile(length, 99)), "max": int(length.max())},
    "tile_max_n_contrib": {"mean": float(tmax.mean()), "p50": float(np.percentile(tmax, 50)), "p90": float(np.percentile(tmax, 90)),
                           "p99": float(np.percentile(tmax, 99)), "max": int(tmax.max()), "sum": int(tmax.sum())},
    "mask_bits_per_instance": float(pop[mask_np].mean()), "mask_zero_frac": float((mask_np == 0).mean()),
    "bwd": {"visits": int(c[0]), "visits_beyond_warp_max": int(c[1]), "visits_with_contrib": int(c[2]), "pairs": int(c[3]),
            "visits_alpha_ok_ignoring_last": int(c[12]),
            "instances_staged": int(c[4]), "warp_batches_skipped": int(c[5]), "warp_batches_processed": int(c[6]),
            "lanes_per_contributing_visit": float(c[3]) / max(1, int(c[2])),
            "lane_hist": [int(x) for x in c[16:49]],
            "lane_private_walk_passes(window 4,8,16,32)": [int(x) for x in c[80:84]],
            "halves_touched_hist(0,1,2)": [int(x) for x in c[64:67]], "quarters_touched_hist(0..4)": [int(x) for x in c[72:77]]},
    "fwd": {"visits": int(c[7]), "visits_with_blend": int(c[8]), "pairs": int(c[9]), "visits_alpha_ok_ignoring_done": int(c[10]),
            "instances_staged": int(c[11])},
}
print(json.dumps(out, indent=1))
