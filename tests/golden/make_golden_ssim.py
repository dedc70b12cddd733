"""Generates tests/golden/ssim_*.npz by running the UNMODIFIED reference fused-ssim CUDA kernels
(oracle/_ref/fused_ssim_cuda_ref.so, built from /root/reference/submodules/fused-ssim by
oracle/build_ref.py:build_fused_ssim) on a B200.

    gpurun -- 'python tests/golden/make_golden_ssim.py gpurun_out/golden'     # then copy ssim_*.npz here

Each file holds the reference's ssim_map, the three derivative maps and dL_dimg1 for the seeded
dL_dmap of tests/ssim_cases.py.  tests/test_ssim_golden.py checks the CPU oracle against them.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import ssim_cases  # noqa: E402


def main(out_dir):
    import torch
    from oracle import build_ref
    ref = build_ref.load_fused_ssim()
    assert ref is not None, "oracle/_ref/fused_ssim_cuda_ref.so is missing"
    os.makedirs(out_dir, exist_ok=True)
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    for name in ssim_cases.GOLDEN:
        a, b, dl = ssim_cases.make(name)
        ta, tb, tdl = (torch.from_numpy(x).cuda() for x in (a, b, dl))
        m, d0, d1, d2 = ref.fusedssim(C1, C2, ta, tb, True)
        g = ref.fusedssim_backward(C1, C2, ta, tb, tdl, d0, d1, d2)
        torch.cuda.synchronize()
        path = os.path.join(out_dir, "ssim_" + name + ".npz")
        np.savez_compressed(path, ssim_map=m.cpu().numpy(), dm_dmu1=d0.cpu().numpy(), dm_dsigma1_sq=d1.cpu().numpy(),
                            dm_dsigma12=d2.cpu().numpy(), dL_dimg1=g.cpu().numpy(),
                            gpu=np.bytes_(torch.cuda.get_device_name(0)))
        print(name, "->", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else HERE)
