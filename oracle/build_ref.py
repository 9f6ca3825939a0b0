#!/usr/bin/env python
"""Build the UNMODIFIED reference rasterizer (CUDA) into oracle/_ref/.

TEST / BENCH INFRASTRUCTURE ONLY.  Nothing under guava_renderer_b200/ or
diff_gaussian_rasterization_32/ may import what this script produces.

The reference (`/root/reference/submodules/diff-gaussian-rasterization-32`) is a
CUDA/C++ torch extension.  Its five translation units are compiled *where they
lie* (no source is copied into the repository) with explicit nvcc / g++
commands that mirror what its `setup.py:16-30` asks `CUDAExtension` to do
(`-std=c++17`, torch's COMMON_NVCC_FLAGS, GLM include path), plus

  * ``-gencode arch=compute_100a,code=sm_100a``  (setup.py passes no arch; it
    inherits TORCH_CUDA_ARCH_LIST, which is 10.0a for this image), and
  * ``-include cstdint`` (rasterizer_impl.h:24,40-61 uses std::uintptr_t /
    uint32_t without including <cstdint>; gcc 13 rejects that.  This is a
    command-line include, not a source edit).

Outputs (all git-ignored, all travel to the GPU box with gpurun):
  oracle/_ref/obj/*.o
  oracle/_ref/diff_gaussian_rasterization_32/_C.so
  oracle/_ref/diff_gaussian_rasterization_32/__init__.py   (build artefact: the
      reference's own 208-line Python wrapper, copied next to its _C.so exactly
      as `pip install --target` would place it)

`rasterize_points.cu` takes ~5 min under nvcc (torch headers through cudafe);
the build is skipped when outputs are newer than the sources.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get(
    "GUAVA_REFERENCE_DGR32",
    "/root/reference/submodules/diff-gaussian-rasterization-32",
)
OUT = os.path.join(HERE, "_ref")
PKG = os.path.join(OUT, "diff_gaussian_rasterization_32")
OBJ = os.path.join(OUT, "obj")

CU_SOURCES = [
    "cuda_rasterizer/rasterizer_impl.cu",
    "cuda_rasterizer/forward.cu",
    "cuda_rasterizer/backward.cu",
    "rasterize_points.cu",
]
CPP_SOURCES = ["ext.cpp"]


def reference_present() -> bool:
    return os.path.isfile(os.path.join(REF, "ext.cpp"))


def built() -> bool:
    return os.path.isfile(os.path.join(PKG, "_C.so")) and os.path.isfile(
        os.path.join(PKG, "__init__.py")
    )


def _newer(out: str, srcs: list[str]) -> bool:
    if not os.path.isfile(out):
        return False
    t = os.path.getmtime(out)
    return all(os.path.getmtime(s) <= t for s in srcs if os.path.isfile(s))


def _run(cmd: list[str]) -> None:
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout[-6000:])
        raise RuntimeError("reference build failed: " + " ".join(cmd[:4]) + " ...")


def build(verbose: bool = True) -> str | None:
    """Returns the path to put on sys.path (oracle/_ref), or None if the
    reference tree is absent and nothing prebuilt exists."""
    if not reference_present():
        return OUT if built() else None

    import torch  # noqa: F401  (only for include / lib paths)
    from torch.utils.cpp_extension import COMMON_NVCC_FLAGS, include_paths, library_paths

    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(PKG, exist_ok=True)
    incs = include_paths("cuda") + [sysconfig.get_paths()["include"]]
    inc_flags = [f"-I{p}" for p in incs] + [
        f"-I{os.path.join(REF, 'third_party/glm')}",
        f"-I{REF}",
    ]
    defs = ["-DTORCH_API_INCLUDE_EXTENSION_H", "-DTORCH_EXTENSION_NAME=_C"]
    all_hdrs = []
    for root, _d, files in os.walk(os.path.join(REF, "cuda_rasterizer")):
        all_hdrs += [os.path.join(root, f) for f in files if f.endswith(".h")]
    all_hdrs.append(os.path.join(REF, "rasterize_points.h"))

    jobs = []
    objs = []
    for s in CU_SOURCES:
        src = os.path.join(REF, s)
        obj = os.path.join(OBJ, os.path.basename(s) + ".o")
        objs.append(obj)
        if _newer(obj, [src] + all_hdrs):
            continue
        jobs.append(
            ["nvcc", "-c", src, "-o", obj, "-O3", "-std=c++17",
             "-gencode", "arch=compute_100a,code=sm_100a",
             "-include", "cstdint", "-Xcompiler", "-fPIC", "-w"]
            + COMMON_NVCC_FLAGS + defs + inc_flags
        )
    for s in CPP_SOURCES:
        src = os.path.join(REF, s)
        obj = os.path.join(OBJ, os.path.basename(s) + ".o")
        objs.append(obj)
        if _newer(obj, [src] + all_hdrs):
            continue
        jobs.append(
            ["g++", "-c", src, "-o", obj, "-O3", "-std=c++17", "-fPIC", "-w"]
            + defs + inc_flags
        )
    if jobs and verbose:
        print(f"[oracle/_ref] compiling {len(jobs)} reference TU(s) "
              "(rasterize_points.cu takes ~5 min)...", flush=True)
    with ThreadPoolExecutor(max_workers=max(1, min(len(jobs), os.cpu_count() or 1))) as ex:
        list(ex.map(_run, jobs))

    so = os.path.join(PKG, "_C.so")
    if not _newer(so, objs):
        libs = library_paths("cuda")
        _run(["g++", "-shared", "-o", so] + objs
             + [f"-L{p}" for p in libs]
             + [f"-Wl,-rpath,{p}" for p in libs]
             + ["-lc10", "-ltorch_cpu", "-ltorch", "-ltorch_python",
                "-lc10_cuda", "-ltorch_cuda", "-lcudart"])
    shutil.copyfile(
        os.path.join(REF, "diff_gaussian_rasterization_32", "__init__.py"),
        os.path.join(PKG, "__init__.py"),
    )
    if verbose:
        print(f"[oracle/_ref] reference extension ready: {so}", flush=True)
    return OUT


if __name__ == "__main__":
    p = build()
    print(p if p else "reference tree absent and no prebuilt oracle/_ref")
