"""In-tree build of libguava_rasterizer_b200.so (hand-written sm_100a CUDA).

    python guava_renderer_b200/build.py [--force] [--verbose]

nvcc cross-compiles without a GPU.  The shared object is written next to the
sources (guava_renderer_b200/lib/) so that it travels with the repository
snapshot to the GPU box; it is git-ignored.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
OBJDIR = os.path.join(LIBDIR, "obj")
SO = os.path.join(LIBDIR, "libguava_rasterizer_b200.so")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

SOURCES = ["api.cu", "preprocess.cu", "binning.cu", "onesweep.cu", "render_forward.cu", "render_backward.cu", "deform.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-O2",
    "-I" + INCLUDE, "-I" + CSRC,
]


def _deps():
    out = [os.path.join(INCLUDE, "guava_rasterizer.h")]
    for f in os.listdir(CSRC):
        if f.endswith((".cuh", ".h")):
            out.append(os.path.join(CSRC, f))
    return out


def _stale(out, srcs):
    if not os.path.isfile(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(s) > t for s in srcs)


def _run(cmd, verbose):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose and r.stdout.strip():
        print(r.stdout)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("build command failed: " + " ".join(cmd))


TORCH_SO = os.path.join(LIBDIR, "_guava_torch.so")


def build_torch_binding(force: bool = False, verbose: bool = False) -> str:
    """csrc/torch_binding.cpp -> lib/_guava_torch.so: the compiled `_C` module (pybind, g++ only -- torch headers are
    kept out of the .cu files), linked against libguava_rasterizer_b200.so next to it (rpath $ORIGIN)."""
    import sysconfig
    import torch
    from torch.utils import cpp_extension as ce
    src = os.path.join(CSRC, "torch_binding.cpp")
    if not (force or _stale(TORCH_SO, [src, os.path.join(INCLUDE, "guava_rasterizer.h"), SO])):
        return TORCH_SO
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    inc = [p for p in ce.include_paths() if os.path.isdir(p)]     # no device_type argument: CPU-side headers suffice ...
    cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")   # ... plus cuda_runtime_api.h for ATen/cuda
    cmd = ["g++", "-shared", "-fPIC", "-O2", "-std=c++17", "-o", TORCH_SO, src,
           "-DTORCH_EXTENSION_NAME=_guava_torch", "-DTORCH_API_INCLUDE_EXTENSION_H",
           "-D_GLIBCXX_USE_CXX11_ABI=" + str(int(torch._C._GLIBCXX_USE_CXX11_ABI)),
           "-I" + INCLUDE, "-I" + cuda_inc, "-I" + sysconfig.get_paths()["include"]] + ["-I" + p for p in inc] + \
          ["-L" + tlib, "-L" + LIBDIR, "-lguava_rasterizer_b200", "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch",
           "-ltorch_python", "-Wl,-rpath,$ORIGIN", "-Wl,-rpath," + tlib, "-Wno-deprecated-declarations"]
    _run(cmd, verbose)
    return TORCH_SO


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJDIR, exist_ok=True)
    deps = _deps()
    jobs, objs = [], []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(OBJDIR, s + ".o")
        objs.append(obj)
        if force or _stale(obj, [src] + deps):
            cmd = ["nvcc", "-c", src, "-o", obj] + NVCC_FLAGS
            if verbose:
                cmd += ["-Xptxas", "-v"]
            jobs.append(cmd)
    with ThreadPoolExecutor(max_workers=max(1, len(jobs))) as ex:
        list(ex.map(lambda c: _run(c, verbose), jobs))
    if force or jobs or _stale(SO, objs):
        _run(["nvcc", "-shared", "-o", SO] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"], verbose)
    build_torch_binding(force, verbose)
    return SO


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
