"""Stage-by-stage device times of the UNMODIFIED reference CUDA rasterizer (oracle/_ref, built for sm_100a) next to this
library's kernels, same box, same workload, same view -- from one ncu launch list.

    # on the GPU box (profiler range = ITERS single-view fwd+bwd of each implementation, after warm-up):
    ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
        --log-file gpurun_out/ref_stages.csv python tools/ref_stage_table.py run c2
    # anywhere:
    python tools/ref_stage_table.py table gpurun_out/ref_stages.csv > profiles/rNN_ref_stage_table.md

ncu serialises launches and runs them cold-cache, so these are per-kernel times, not a step time (bench.py's
`reference_cuda_same_box` and `value` are the step times); what the table is for is the kernel-against-kernel comparison.
Test / measurement infrastructure only: the product never loads oracle/_ref.
"""
import collections
import csv
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ITERS = 3

# stage -> (regex of reference kernels, regex of this library's kernels); reference file:line of each stage in
# /root/reference/diff-gaussian-rasterization-npu/cuda_rasterizer/
STAGES = [
    ("preprocess fwd (forward.cu:154-272)", r"^preprocessCUDA#fwd", r"^preprocess_fwd_kernel"),
    ("scan of tiles_touched (rasterizer_impl.cu:280)", r"DeviceScan", r"^scan_compact_kernel"),
    ("duplicateWithKeys (rasterizer_impl.cu:70-111)", r"^duplicateWithKeys", r"^(sorted_scan_kernel|emit_tiles_kernel)"),
    ("radix sort (rasterizer_impl.cu:306-311)", r"DeviceRadixSort", r"^(sort_hist_kernel|onesweep_kernel)"),
    ("identifyTileRanges (rasterizer_impl.cu:116-138)", r"^identifyTileRanges", r"^tile_ranges_kernel"),
    ("blend fwd (forward.cu:277-400)", r"^renderCUDA#fwd", r"^render_fwd_kernel"),
    ("blend bwd (backward.cu:452-638)", r"^renderCUDA#bwd", r"^render_bwd_kernel"),
    ("per-Gaussian bwd (backward.cu:147-326, 398-449)", r"^(computeCov2DCUDA|preprocessCUDA#bwd)", r"^preprocess_bwd_kernel"),
]


def run(workload):
    for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-npu_b200"), os.path.join(ROOT, "tests")):
        sys.path.insert(0, p)
    import numpy as np
    import torch
    import bench
    from gsr_b200 import capi, synth
    from oracle import ref_cuda
    w, scene, cam = bench.make_workload(workload)
    dev = torch.device("cuda:0")
    t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)
    dcol, ddep = (t(x) for x in synth.upstream_grads(w["H"], w["W"], 5))
    kw = dict(means3D=t(scene["means3D"]), opacities=t(scene["opacities"]), viewmatrix=t(cam.world_view_transform),
              projmatrix=t(cam.full_proj_transform), cam_pos=t(cam.camera_center), background=torch.zeros(3, device=dev),
              tan_fovx=cam.tanfovx, tan_fovy=cam.tanfovy, shs=t(scene["shs"]), scales=t(scene["scales"]),
              rotations=t(scene["rotations"]), sh_degree=3, antialiasing=w["aa"])
    R = ref_cuda.RefRasterizer()

    def ref_iter():
        f = R.forward(width=w["W"], height=w["H"], **kw)
        R.backward(f["radii"], dL_dcolor=dcol, dL_dinvdepth=ddep, **kw)

    def our_iter():
        st = capi.forward(width=w["W"], height=w["H"], **kw)
        capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, **kw)

    for _ in range(3):
        ref_iter(); our_iter()
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    for _ in range(ITERS):
        ref_iter()
    torch.cuda.synchronize()
    for _ in range(ITERS):
        our_iter()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    R.close()


def table(path):
    rows = list(csv.DictReader(l for l in open(path) if l.startswith('"')))
    seen = collections.Counter()
    times = collections.defaultdict(float)       # tagged kernel name -> summed us over ITERS iterations
    counts = collections.Counter()
    for r in rows:
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        name = r["Kernel Name"].split("(")[0].replace("void ", "").strip()
        name = re.sub(r"<.*", "", name) if not name.startswith(("cub::", "at::")) else name
        us = float(r["Metric Value"].replace(",", "")) / (1e3 if r["Metric Unit"] in ("ns", "nsecond") else 1.0)
        if name in ("preprocessCUDA", "renderCUDA"):           # the reference uses each name in forward AND backward
            k = seen[name]
            seen[name] += 1
            name += "#fwd" if k % 2 == 0 else "#bwd"
        times[name] += us
        counts[name] += 1
    used = set()
    print("| stage (reference file:line) | reference kernels | reference us | this library's kernels | ours us | reference / ours |")
    print("|---|---|---|---|---|---|")
    tot_r = tot_o = 0.0
    for stage, rx_ref, rx_our in STAGES:
        def pick(rx):
            ks = [k for k in times if re.search(rx, k)]
            used.update(ks)
            return ks, sum(times[k] for k in ks) / ITERS
        kr, ur = pick(rx_ref)
        ko, uo = pick(rx_our)
        tot_r += ur
        tot_o += uo
        fmt = lambda ks: ", ".join("%s x%d" % (re.sub(r"^cub::(detail::)?", "cub::", k)[:48], round(counts[k] / ITERS)) for k in sorted(ks)) or "-"
        print("| %s | %s | %.1f | %s | %.1f | %s |" % (stage, fmt(kr), ur, fmt(ko), uo, ("%.2fx" % (ur / uo)) if uo else "-"))
    rest = [k for k in times if k not in used]
    print("| **sum of the stages** | | **%.1f** | | **%.1f** | **%.2fx** |" % (tot_r, tot_o, tot_r / tot_o if tot_o else 0))
    print()
    print("Other launches inside the profiled range (fills, copies of the harnesses): " +
          ", ".join("%s x%d = %.1f us/iter" % (k[:60], round(counts[k] / ITERS), times[k] / ITERS) for k in sorted(rest)))


if __name__ == "__main__":
    if sys.argv[1] == "run":
        run(sys.argv[2] if len(sys.argv) > 2 else "c2")
    else:
        table(sys.argv[2])
