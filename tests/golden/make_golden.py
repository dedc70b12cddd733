"""Generates the golden vectors in tests/golden/*.npz by running the UNMODIFIED reference CUDA
rasterizer (oracle/_ref, built from /root/reference by oracle/build_ref.py) on a B200.

    gpurun -- 'python tests/golden/make_golden.py gpurun_out/golden [case ...]'     # then copy *.npz here

Each file holds, for one seeded case of tests/cases.py, the reference's outputs: every integer
artefact in full (radii, tiles_touched, offsets, sorted keys, point list, tile ranges, n_contrib),
the images (colour, final_T, inverse depth) and the per-Gaussian float arrays / gradients, the big
ones sub-sampled every `stride` rows to keep the fixtures small.  The CPU test-suite checks the
oracle against these files (tests/test_oracle_golden.py): that is what pins the oracle.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import cases  # noqa: E402


def main(out_dir, names=None):
    os.makedirs(out_dir, exist_ok=True)
    import torch
    for name in (names or cases.GOLDEN_CASES):
        scene, cam, opts = cases.get_case(name)
        grads = cases.grads_for(cam)
        r = cases.run_ref(scene, cam, opts, grads)
        P = scene["means3D"].shape[0]
        stride = 1 if P <= 2000 else 16
        rows = np.arange(0, P, stride)
        out = dict(stride=np.int64(stride), num_rendered=np.int64(r["num_rendered"]),
                   gpu=np.bytes_(torch.cuda.get_device_name(0)))
        for k in ("radii", "tiles_touched", "point_offsets", "keys", "point_list", "ranges", "n_contrib"):
            out[k] = r[k]
        for k in ("out_color", "final_T", "invdepth"):
            out[k] = r[k]
        for k in ("depths", "means2D", "conic_opacity", "rgb", "cov3D", "clamped"):
            out[k] = r[k][rows]
        for k in cases.BWD_KEYS:
            if r.get(k) is not None and r[k].size:
                out[k] = r[k][rows]
                out[k + "_colsum"] = r[k].astype(np.float64).reshape(P, -1).sum(axis=0)
        path = os.path.join(out_dir, name + ".npz")
        np.savez_compressed(path, **out)
        print(name, "->", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(HERE), sys.argv[2:])
