"""Builds libgsr.so (the sm_100a CUDA kernels + C ABI of include/gsr.h) in-tree with nvcc.

No torch C++ extension is involved: the Python packages in this directory bind the C ABI with
ctypes, so the build is five nvcc calls (seconds each) and the .so travels with the source tree.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libgsr.so")
SOURCES = ["gsr_api.cu", "gsr_preprocess.cu", "gsr_sort.cu", "gsr_binning.cu", "gsr_blend.cu", "gsr_knn.cu", "gsr_ssim.cu", "gsr_adam.cu", "gsr_densify.cu", "gsr_comm.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC"]


TORCH_EXT = os.path.join(HERE, "_gsr_torch.so")


def build_torch_binding(verbose=False):
    """The thin PyTorch C++ extension over the C ABI (csrc/torch_binding.cpp -> _gsr_torch.so, in-tree): host compiler
    only, links libgsr.so (rpath $ORIGIN) and torch's own libraries."""
    import sysconfig
    import torch
    src = os.path.join(CSRC, "torch_binding.cpp")
    if not _stale(TORCH_EXT, [src, os.path.join(HERE, "..", "include", "gsr.h"), LIB]):
        return TORCH_EXT
    ti = os.path.dirname(torch.__file__)
    cmd = [os.environ.get("CXX", "g++"), "-O2", "-std=c++17", "-fPIC", "-shared",
           f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}", "-DTORCH_EXTENSION_NAME=_gsr_torch",
           "-DTORCH_API_INCLUDE_EXTENSION_H", f"-I{ti}/include", f"-I{ti}/include/torch/csrc/api/include",
           f"-I{sysconfig.get_paths()['include']}", "-I/usr/local/cuda/include", src, "-o", TORCH_EXT, f"-L{ti}/lib",
           "-ltorch", "-ltorch_cpu", "-ltorch_cuda", "-lc10", "-lc10_cuda", "-ltorch_python", f"-L{HERE}", "-lgsr",
           "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{ti}/lib",
           # the SHARED libstdc++ torch itself runs on. (A compiler whose search path only finds libstdc++.a links a
           # private copy of the iostream / locale code into the module; two copies in one process disagree about the
           # locale facet ids and the first formatted integer in an error message crashes.)
           "-nostdlib++", "-l:libstdc++.so.6"]
    if verbose:
        print(" ".join(cmd), file=sys.stderr)
    subprocess.check_call(cmd)
    return TORCH_EXT


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, variant=None, extra_flags=()):
    """variant / extra_flags: an experimental build beside the product library (libgsr_<variant>.so, own object
    directory), selected at run time with GSR_LIB=<path>; used for A/B measurements of compile-time constants."""
    hdrs = [os.path.join(CSRC, "gsr_common.cuh"), os.path.join(HERE, "..", "include", "gsr.h")]
    objdir = os.path.join(HERE, "build" if not variant else "build_" + variant)
    lib = LIB if not variant else os.path.join(HERE, f"libgsr_{variant}.so")
    os.makedirs(objdir, exist_ok=True)
    jobs = []
    for src in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(objdir, src.replace(".cu", ".o"))
        if force or _stale(o, [s] + hdrs):
            jobs.append([NVCC] + FLAGS + list(extra_flags) + ["-c", s, "-o", o])
    def run(cmd):
        if verbose:
            print(" ".join(cmd), file=sys.stderr)
        subprocess.check_call(cmd)
    with ThreadPoolExecutor(max_workers=8) as ex:
        list(ex.map(run, jobs))
    objs = [os.path.join(objdir, s.replace(".cu", ".o")) for s in SOURCES]
    if force or jobs or _stale(lib, objs):
        run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", lib] + objs)
    return lib


if __name__ == "__main__":
    if "--variant" in sys.argv:                    # python build.py --variant sort2048 -DGSR_SORT_TILE=2048
        i = sys.argv.index("--variant")
        print(build(verbose=True, variant=sys.argv[i + 1], extra_flags=[a for a in sys.argv[i + 2:] if a.startswith("-D")]))
    else:
        print(build(force="--force" in sys.argv, verbose=True))
        if "--no-torch-binding" not in sys.argv:
            print(build_torch_binding(verbose=True))
