"""GPU-box utility: run every parity case through (this repo's CUDA path, the CPU oracle, the
reference CUDA build oracle/_ref) and print/save the mismatch report.

    python tests/parity_report.py [case ...] [--json gpurun_out/parity.json]
"""
import json
import sys
import time
import traceback

import numpy as np

import cases


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    out_json = None
    if "--json" in sys.argv:
        out_json = sys.argv[sys.argv.index("--json") + 1]
        args = [a for a in args if a != out_json]
    names = args or (cases.SMALL_CASES + cases.MEDIUM_CASES)
    from oracle import ref_cuda
    have_ref = ref_cuda.available()
    report = {}
    for name in names:
        try:
            scene, cam, opts = cases.get_case(name)
            grads = cases.grads_for(cam)
            t = time.time(); o = cases.run_oracle(scene, cam, opts, grads); t_or = time.time() - t
            c = cases.run_cuda(scene, cam, opts, grads)
            rep = {"P": int(scene["means3D"].shape[0]), "N": int(o["num_rendered"]), "oracle_s": round(t_or, 3),
                   "borderline_pixels": int(o["borderline"].sum()),
                   "cuda_vs_oracle": cases.compare(c, o, o["borderline"])}
            if have_ref:
                r = cases.run_ref(scene, cam, opts, grads)
                rep["ref_vs_oracle"] = cases.compare(r, o, o["borderline"])
                rep["cuda_vs_ref"] = cases.compare(c, r, None)
            report[name] = rep
        except Exception:
            report[name] = {"error": traceback.format_exc()}
        print("=== %s" % name)
        print(json.dumps(report[name], indent=1, default=str))
        sys.stdout.flush()
    if out_json:
        with open(out_json, "w") as f:
            json.dump(report, f, indent=1, default=str)


if __name__ == "__main__":
    main()
