"""CPU: the loss-stage oracle against golden vectors of the UNMODIFIED reference fused-ssim CUDA kernel
run on a B200 (tests/golden/ssim_*.npz, made by tests/golden/make_golden_ssim.py).  This is pin 2 of
oracle/ssim_oracle.c (pin 1: tests/test_ssim_oracle_cpu.py)."""
import os

import numpy as np
import pytest

import ssim_cases
from oracle import ssim_oracle as so

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
# The SSIM API kernels are compared BIT FOR BIT (np.array_equal).  Tolerances below are for the fused-loss variant
# (one reciprocal instead of seven divisions) and for comparisons with float64 torch recipes:
MAP_ATOL = 2e-5          # ssim_map (values in [-1, 1]); ill-conditioned where sigma^2 << C2 (smooth images)
DERIV_RTOL = 2e-4        # derivative maps / gradients, relative to the largest magnitude of the array


def _load(name):
    path = os.path.join(GOLD, "ssim_" + name + ".npz")
    if not os.path.exists(path):
        pytest.skip("golden vector %s not generated yet" % os.path.basename(path))
    return np.load(path)


@pytest.mark.parametrize("name", ssim_cases.GOLDEN)
def test_oracle_matches_reference_cuda(name):
    g = _load(name)
    a, b, dl = ssim_cases.make(name)
    m, d0, d1, d2 = so.fusedssim(so.C1, so.C2, a, b, True)
    # the oracle spells the reference's FFMA placement (read from its sm_100a SASS): bit-identical
    assert np.array_equal(m, g["ssim_map"])
    for got, key in ((d0, "dm_dmu1"), (d1, "dm_dsigma1_sq"), (d2, "dm_dsigma12")):
        assert np.array_equal(got, g[key]), key
    gi = so.fusedssim_backward(so.C1, so.C2, a, b, dl, d0, d1, d2)
    assert np.array_equal(gi, g["dL_dimg1"])
