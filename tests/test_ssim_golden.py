"""CPU: the loss-stage oracle against golden vectors of the UNMODIFIED reference fused-ssim CUDA kernel
run on a B200 (tests/golden/ssim_*.npz, made by tests/golden/make_golden_ssim.py).  This is pin 2 of
oracle/ssim_oracle.c (pin 1: tests/test_ssim_oracle_cpu.py)."""
import os

import numpy as np
import pytest

import ssim_cases
from oracle import ssim_oracle as so

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
# fp32 tolerances of the loss stage: maps are O(1) rational expressions of window sums
MAP_ATOL = 2e-5          # ssim_map (values in [-1, 1])
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
    assert np.abs(m - g["ssim_map"]).max() <= MAP_ATOL
    for got, key in ((d0, "dm_dmu1"), (d1, "dm_dsigma1_sq"), (d2, "dm_dsigma12")):
        want = g[key]
        assert np.abs(got - want).max() <= DERIV_RTOL * np.abs(want).max(), key
    # backward from the REFERENCE's derivative maps (isolates the backward kernel)
    gi = so.fusedssim_backward(so.C1, so.C2, a, b, dl, g["dm_dmu1"], g["dm_dsigma1_sq"], g["dm_dsigma12"])
    want = g["dL_dimg1"]
    assert np.abs(gi - want).max() <= DERIV_RTOL * np.abs(want).max()
