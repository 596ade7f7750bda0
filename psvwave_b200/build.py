"""In-tree build of the native pieces (no GPU needed: nvcc cross-compiles sm_100a).

    python -m psvwave_b200.build            # or: python psvwave_b200/build.py

Produces, next to this file,
    libpsvwave_b200.so                       CUDA kernels + host logic + the C ABI
    __psvWave_cpp<EXT_SUFFIX>                pybind11 binding, links only against the C ABI
Both are git-ignored and travel to the GPU box with the working tree.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")
LIB = os.path.join(HERE, "libpsvwave_b200.so")
EXT = os.path.join(HERE, "__psvWave_cpp" + (sysconfig.get_config_var("EXT_SUFFIX") or ".so"))

NVCC = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
HOST_CXX = "/usr/bin/g++"  # the image's $CXX wrapper (/opt/gcc) lacks a few spec files
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]

LIB_SOURCES = ["capi.cu", "model.cu", "config.cpp"]
LIB_DEPS = LIB_SOURCES + ["model.hpp", "config.hpp", "kernels.cuh", "fused.cuh", "fused32.cuh", "pair.cuh", "window.cuh", "pointwise.cuh", "stencil_math.cuh"]


def _stale(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd: list[str], verbose: bool) -> None:
    if verbose:
        print(" ".join(cmd), flush=True)
    subprocess.run(cmd, check=True)


def build(force: bool = False, verbose: bool = False) -> None:
    deps = [os.path.join(CSRC, f) for f in LIB_DEPS] + [os.path.join(INCLUDE, "psvwave_b200.h"), __file__]
    if force or _stale(LIB, deps):
        cmd = [NVCC, "-ccbin", HOST_CXX, "-std=c++17", "-O3", "-lineinfo", *ARCH,
               "-Xcompiler", "-fPIC,-O3,-Wall,-ffp-contract=off",
               "-Xptxas", "-v" if verbose else "-O3",
               *os.environ.get("PSV_EXTRA_NVCC_FLAGS", "").split(),
               "--shared", "-cudart", "static", "-I", INCLUDE, "-o", LIB,
               *[os.path.join(CSRC, f) for f in LIB_SOURCES]]
        _run(cmd, verbose)
    ext_deps = [os.path.join(CSRC, "pybind_module.cpp"), os.path.join(INCLUDE, "psvwave_b200.h"), __file__]
    if force or _stale(EXT, ext_deps + [LIB]):
        import pybind11
        cmd = [HOST_CXX, "-std=c++17", "-O2", "-shared", "-fPIC", "-fvisibility=hidden",
               "-I", pybind11.get_include(), "-I", sysconfig.get_paths()["include"], "-I", INCLUDE,
               os.path.join(CSRC, "pybind_module.cpp"), "-o", EXT,
               "-L", HERE, "-lpsvwave_b200", "-Wl,-rpath,$ORIGIN"]
        _run(cmd, verbose)


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
    print("built", LIB, "and", EXT)
