"""CPU half of the acceptance run (tests/test_gpu_reference_callers.py is the GPU half): the reference's own caller
modules -- train.py, gaussian_renderer/, scene/ -- import unchanged against this repo's `diff_gaussian_rasterization`
and `fused_ssim`, and pick the code paths the integration notes promise (fused SSIM on, sparse Adam / separate_sh
off unless GSR_EXPORT_SPARSE_ADAM=1).  No compute.  Runs in a subprocess: the reference uses top-level package
names (`utils`, `scene`, `arguments`) that must not leak into the test session."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "gaussian-splatting-npu_b200")
STUBS = os.path.join(ROOT, "tests", "stubs")

PROBE = r"""
import os, sys, inspect
import train, gaussian_renderer, diff_gaussian_rasterization as dgr, fused_ssim
from scene import Scene, GaussianModel
assert os.path.realpath(dgr.__file__).startswith(os.path.realpath(sys.argv[1])), dgr.__file__
assert os.path.realpath(fused_ssim.__file__).startswith(os.path.realpath(sys.argv[1])), fused_ssim.__file__
assert os.path.realpath(train.__file__).startswith(os.path.realpath(sys.argv[2])), train.__file__
assert train.FUSED_SSIM_AVAILABLE is True
assert train.SPARSE_ADAM_AVAILABLE is (os.environ.get("GSR_EXPORT_SPARSE_ADAM") == "1")
assert gaussian_renderer.GaussianRasterizer is dgr.GaussianRasterizer
# the keyword set render() passes (gaussian_renderer/__init__.py:88-110) is accepted by this package's forward()
params = inspect.signature(dgr.GaussianRasterizer.forward).parameters
for k in ("means3D", "means2D", "dc", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp"):
    assert k in params, k
print("callers-import-ok")
"""


def _callers_dir():
    sys.path.insert(0, ROOT)
    from oracle import build_ref
    d = build_ref.callers_dir()
    if d is None:
        pytest.skip("reference callers not available (neither /root/reference nor oracle/_ref/callers)")
    return d


@pytest.mark.parametrize("sparse", ["0", "1"])
def test_reference_callers_import_on_this_package(sparse):
    d = _callers_dir()
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([PKG, STUBS, d])
    env["GSR_EXPORT_SPARSE_ADAM"] = sparse
    r = subprocess.run([sys.executable, "-c", PROBE, PKG, d], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "callers-import-ok" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]
