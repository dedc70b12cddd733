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


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
