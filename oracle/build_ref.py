"""Recipe for oracle/_ref/libref_rasterizer.so: the UNMODIFIED reference CUDA rasterizer compiled
for sm_100a from where its sources lie under /root/reference (nothing is copied into the repo).

    python oracle/build_ref.py

Runs only where /root/reference exists (the build container).  The .so is git-ignored but travels
to the GPU box with the snapshot.  Two switches work around the fork's half-finished rename
(SURVEY.md section 0): `-include cstdint` (rasterizer_impl.h uses uintptr_t without the header) and
`-DNPURasterizer=CudaRasterizer` (rasterizer.h declares namespace NPURasterizer, the .cu files
define CudaRasterizer::...).  Default nvcc flags otherwise (-fmad=true, IEEE div/sqrt), exactly
what the reference's own CMakeLists.txt would use.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/diff-gaussian-rasterization-npu"
OUT_DIR = os.path.join(HERE, "_ref")
OUT = os.path.join(OUT_DIR, "libref_rasterizer.so")


def available():
    return os.path.isdir(os.path.join(REF, "cuda_rasterizer"))


def build(force=False):
    if not available():
        return OUT if os.path.exists(OUT) else None
    srcs = [os.path.join(REF, "cuda_rasterizer", f) for f in ("forward.cu", "backward.cu", "rasterizer_impl.cu")]
    srcs.append(os.path.join(HERE, "ref_harness.cu"))
    if not force and os.path.exists(OUT) and all(os.path.getmtime(s) <= os.path.getmtime(OUT) for s in srcs):
        return OUT
    os.makedirs(OUT_DIR, exist_ok=True)
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-shared",
           "-Xcompiler", "-fPIC", "-include", "cstdint", "-DNPURasterizer=CudaRasterizer",
           "-I", os.path.join(REF, "third_party", "glm"), "-I", os.path.join(REF, "cuda_rasterizer"),
           "-diag-suppress", "177,550", "-w",
           *srcs, "-o", OUT + ".tmp", "-cudart", "shared"]
    subprocess.check_call(cmd)
    os.replace(OUT + ".tmp", OUT)
    return OUT


# ---- the reference's fused-ssim extension (loss stage) --------------------------
FS = "/root/reference/submodules/fused-ssim"
FS_OUT = os.path.join(OUT_DIR, "fused_ssim_cuda_ref.so")


def fused_ssim_available():
    return os.path.exists(os.path.join(FS, "ssim.cu"))


def build_fused_ssim(force=False):
    """UNMODIFIED FS/ssim.cu + FS/ext.cpp compiled for sm_100a as the torch extension module
    `fused_ssim_cuda_ref` (the module name is the only change: -DTORCH_EXTENSION_NAME, so that it can be
    imported next to the product's fused_ssim_cuda).  FS/setup.py is not run.  nvcc needs ~6 minutes
    for the torch headers."""
    if not fused_ssim_available():
        return FS_OUT if os.path.exists(FS_OUT) else None
    srcs = [os.path.join(FS, "ssim.cu"), os.path.join(FS, "ext.cpp")]
    if not force and os.path.exists(FS_OUT) and all(os.path.getmtime(s) <= os.path.getmtime(FS_OUT) for s in srcs):
        return FS_OUT
    import sysconfig
    import torch
    from torch.utils import cpp_extension as ce
    os.makedirs(OUT_DIR, exist_ok=True)
    inc = []
    for p in ce.include_paths("cuda"):
        inc += ["-I", p]
    inc += ["-I", sysconfig.get_paths()["include"]]
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-shared", "-Xcompiler", "-fPIC",
           "-std=c++17", "-DTORCH_EXTENSION_NAME=fused_ssim_cuda_ref", "-DTORCH_API_INCLUDE_EXTENSION_H",
           "-D_GLIBCXX_USE_CXX11_ABI=%d" % abi, "-w", *inc, *srcs, "-o", FS_OUT + ".tmp",
           "-L", tlib, "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python",
           "-Xlinker", "-rpath," + tlib]
    subprocess.check_call(cmd)
    os.replace(FS_OUT + ".tmp", FS_OUT)
    return FS_OUT


def load_fused_ssim():
    """Imports oracle/_ref/fused_ssim_cuda_ref.so (GPU box: the prebuilt file).  None if it does not exist."""
    if not os.path.exists(FS_OUT):
        return None
    import importlib.util
    import torch  # noqa: F401  (libtorch must be loaded first)
    spec = importlib.util.spec_from_file_location("fused_ssim_cuda_ref", FS_OUT)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
    print(build_fused_ssim(force="--force" in sys.argv))
