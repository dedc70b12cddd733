"""Per-kernel DRAM traffic (dram__bytes_read.sum + dram__bytes_write.sum) of the LAST step in an
`ncu --set full` report of tools/profile_step.py, summed per stage -> JSON for bench.py's
`roofline.traffic`.

    python tools/ncu_traffic.py gpurun_out/full.ncu-rep c2 > profiles/rNN_xx_traffic.json
"""
import csv
import io
import json
import subprocess
import sys

STAGE_OF = {"preprocess_fwd_kernel": "preprocess", "scan_compact_kernel": "scan", "sorted_scan_kernel": "duplicate",
            "emit_tiles_kernel": "duplicate", "tile_ranges_kernel": "ranges", "render_fwd_kernel": "render_fwd",
            "render_bwd_kernel": "render_bwd", "preprocess_bwd_kernel": "preprocess_bwd"}
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    rep, workload = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    ki, ri, wi = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    launches = []
    for r in rows[2:]:
        name = r[ki].split("(")[0].replace("void ", "").split("<")[0]
        b = float(r[ri].replace(",", "")) * SCALE[units[ri]] + float(r[wi].replace(",", "")) * SCALE[units[wi]]
        launches.append((name, b))
    # last step = everything from the last preprocess_fwd_kernel on
    start = max(i for i, (n, _) in enumerate(launches) if n == "preprocess_fwd_kernel")
    out, sort_seen = {}, 0
    for name, b in launches[start:]:
        if name in ("sort_hist_kernel", "onesweep_kernel"):
            # the first hist + 4 onesweep launches are the depth sort, the rest the tile sort
            stage = "depth_sort" if sort_seen < 5 else "sort"
            sort_seen += 1
        else:
            stage = STAGE_OF.get(name)
        if stage:
            out[stage] = out.get(stage, 0.0) + b
    print(json.dumps({"workload": workload, "source": "ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, one step",
                      "traffic_bytes": {k: int(v) for k, v in out.items()}}, indent=1))


if __name__ == "__main__":
    main()
