"""In-tree build of the sm_100a rasterizer: libgsr_b200.so (CUDA kernels + C ABI, no torch) and
diff_gaussian_rasterization/_C*.so (the thin torch binding that keeps the reference's pybind surface).

    python gaussian-splatting-npu_b200/build.py [--force]

nvcc cross-compiles for sm_100a without a GPU.  Outputs stay in the tree (git-ignored) so that they
travel with the snapshot to the GPU box.
"""
import os
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "gsr_b200", "libgsr_b200.so")
PKG = os.path.join(HERE, "diff_gaussian_rasterization")

CU_SOURCES = ["gsr_api.cu", "preprocess_fwd.cu", "binning.cu", "radix_sort.cu", "render_fwd.cu",
              "render_bwd.cu", "preprocess_bwd.cu", "ssim.cu", "adam.cu", "ply_io.cu", "epilogue.cu", "frame_out.cu", "exchange.cu", "preprocess_batch.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xptxas", "-v", "--expt-relaxed-constexpr"]


def _newer(target, deps):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(d) <= t for d in deps)


def ext_path():
    return os.path.join(PKG, "_C" + sysconfig.get_config_var("EXT_SUFFIX"))


def build_lib(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, "gsr_common.cuh"), os.path.join(CSRC, "adam_math.cuh"), os.path.join(CSRC, "preprocess_math.cuh"), os.path.join(ROOT, "include", "gsr.h"),
               os.path.join(ROOT, "include", "gsr_dist.h"),
               os.path.join(ROOT, "include", "gsr_loss.h"), os.path.join(ROOT, "include", "gsr_optim.h"),
               os.path.join(ROOT, "include", "gsr_ply.h"), os.path.join(ROOT, "include", "gsr_frame.h")]
    objs, jobs = [], []
    for src in CU_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src.replace(".cu", ".o"))
        objs.append(o)
        if force or not _newer(o, [s] + headers):
            jobs.append((s, o))

    def compile_one(job):
        s, o = job
        r = subprocess.run(["nvcc", *NVCC_FLAGS, "-c", s, "-o", o], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed on %s:\n%s\n%s" % (s, r.stdout, r.stderr))
        with open(o + ".ptxas.log", "w") as f:
            f.write(r.stderr)
        if verbose:
            sys.stderr.write(r.stderr)

    with ThreadPoolExecutor(max_workers=8) as ex:
        list(ex.map(compile_one, jobs))
    if force or jobs or not _newer(LIB, objs):
        subprocess.check_call(["nvcc", "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB + ".tmp",
                               *objs, "-cudart", "shared"])
        os.replace(LIB + ".tmp", LIB)
    return LIB


def ssim_ext_path():
    return os.path.join(HERE, "fused_ssim_cuda" + sysconfig.get_config_var("EXT_SUFFIX"))


def _compile_ext(name, sources, out, defines=()):
    import torch
    from torch.utils import cpp_extension as ce
    inc = []
    for p in ce.include_paths("cuda") if "device_type" in ce.include_paths.__code__.co_varnames else ce.include_paths(cuda=True):
        inc += ["-I", p]
    inc += ["-I", sysconfig.get_paths()["include"], "-I", os.path.join(ROOT, "include"), "-I", "/usr/local/cuda/include"]
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    rel = os.path.relpath(os.path.dirname(LIB), os.path.dirname(out))
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-DTORCH_EXTENSION_NAME=" + name, "-DTORCH_API_INCLUDE_EXTENSION_H",
           "-D_GLIBCXX_USE_CXX11_ABI=%d" % abi, *defines, *inc, *sources, "-o", out + ".tmp",
           "-L", tlib, "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python",
           "-L", os.path.dirname(LIB), "-lgsr_b200", "-L", "/usr/local/cuda/lib64", "-lcudart",
           "-Wl,-rpath,$ORIGIN/" + rel, "-Wl,-rpath," + tlib]
    subprocess.check_call(cmd)
    os.replace(out + ".tmp", out)
    return out


def build_ext(force=False):
    """_C (rasterizer + loss stage) and the top-level fused_ssim_cuda module (the reference's name)."""
    # the extensions link libgsr_b200.so dynamically (rpath): only an ABI change -- the headers -- needs a rebuild
    hdrs = [os.path.join(ROOT, "include", "gsr.h"), os.path.join(ROOT, "include", "gsr_loss.h")]
    main_src = os.path.join(CSRC, "torch_binding.cpp")
    ssim_src = os.path.join(CSRC, "ssim_binding.cpp")
    jobs = []
    out = ext_path()
    if force or not _newer(out, [main_src, ssim_src] + hdrs):
        jobs.append(("_C", [main_src, ssim_src], out, ()))
    out2 = ssim_ext_path()
    if force or not _newer(out2, [ssim_src] + hdrs):
        jobs.append(("fused_ssim_cuda", [ssim_src], out2, ("-DGSR_SSIM_STANDALONE",)))
    with ThreadPoolExecutor(max_workers=2) as ex:
        list(ex.map(lambda j: _compile_ext(*j), jobs))
    return out


def build_all(force=False, verbose=False):
    lib = build_lib(force, verbose)
    ext = build_ext(force)
    return lib, ext


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv, verbose="-v" in sys.argv))
