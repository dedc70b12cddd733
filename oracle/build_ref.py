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




# ---- the reference's Python callers of the path (acceptance run) -----------------------
# north_star: "gaussian_renderer/__init__.py and train.py use it unchanged".  The `-m gpu` acceptance test
# (tests/test_gpu_reference_callers.py) runs those files AS THEY ARE on this package.  /root/reference does not
# exist on the GPU box, so -- exactly like the compiled reference above -- the files are staged, byte for byte,
# into oracle/_ref/callers/ (git-ignored: never part of the repo's history; travels with the gpurun snapshot).
CALLERS_SRC = "/root/reference"
CALLERS_OUT = os.path.join(OUT_DIR, "callers")
CALLER_FILES = ["train.py", "render.py", "gaussian_renderer/__init__.py", "gaussian_renderer/network_gui.py",
                "scene/__init__.py", "scene/cameras.py", "scene/colmap_loader.py", "scene/dataset_readers.py",
                "scene/gaussian_model.py", "arguments/__init__.py", "utils/camera_utils.py", "utils/general_utils.py",
                "utils/graphics_utils.py", "utils/image_utils.py", "utils/loss_utils.py", "utils/sh_utils.py",
                "utils/system_utils.py"]


def callers_available():
    return os.path.exists(os.path.join(CALLERS_SRC, "train.py"))


def callers_dir():
    """Where the unmodified caller files can be imported from: the reference tree itself when present, else the staged
    copy; None when neither exists."""
    if callers_available():
        return CALLERS_SRC
    return CALLERS_OUT if os.path.exists(os.path.join(CALLERS_OUT, "train.py")) else None


def stage_callers():
    if not callers_available():
        return CALLERS_OUT if os.path.isdir(CALLERS_OUT) else None
    import shutil
    for rel in CALLER_FILES:
        dst = os.path.join(CALLERS_OUT, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(CALLERS_SRC, rel), dst)
    return CALLERS_OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
    print(build_fused_ssim(force="--force" in sys.argv))
    print(stage_callers())
