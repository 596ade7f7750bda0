"""In-tree build of the native pieces (no GPU needed: nvcc cross-compiles sm_100a).

    python -m psvwave_b200.build            # or: python psvwave_b200/build.py [--force] [--quiet]

Produces, next to this file,
    libpsvwave_b200.so                       CUDA kernels + host logic + the C ABI
    __psvWave_cpp<EXT_SUFFIX>                pybind11 binding, links only against the C ABI
Both are git-ignored and travel to the GPU box with the working tree.  Every kernel family is its own
translation unit; the objects (psvwave_b200/_obj/, git- and gpurun-ignored) are compiled in parallel and
only re-made when a source they include changed.
"""
from __future__ import annotations

import os
import re
import shutil
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "_obj")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")
LIB = os.path.join(HERE, "libpsvwave_b200.so")
EXT = os.path.join(HERE, "__psvWave_cpp" + (sysconfig.get_config_var("EXT_SUFFIX") or ".so"))

NVCC = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
HOST_CXX = "/usr/bin/g++"  # the image's $CXX wrapper (/opt/gcc) lacks a few spec files
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]

LIB_SOURCES = ["capi.cu", "model.cu", "multi.cu", "optim.cu", "config.cpp", "launch_step.cu", "launch_pipe.cu", "launch_pair.cu",
               "launch_resident.cu", "no_alloc.cu"]


def _includes(path: str, seen: set[str]) -> None:
    """Transitive closure of the quoted #include files of one source."""
    if path in seen or not os.path.exists(path):
        return
    seen.add(path)
    for inc in re.findall(r'^\s*#include\s+"([^"]+)"', open(path).read(), re.M):
        _includes(os.path.normpath(os.path.join(os.path.dirname(path), inc)), seen)


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _run(cmd: list[str], verbose: bool) -> None:
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or r.returncode != 0:
        sys.stdout.write(r.stdout)
        sys.stderr.write(r.stderr)
    if r.returncode != 0:
        raise subprocess.CalledProcessError(r.returncode, cmd)


def build(force: bool = False, verbose: bool = False) -> None:
    os.makedirs(OBJ, exist_ok=True)
    extra = os.environ.get("PSV_EXTRA_NVCC_FLAGS", "").split()
    jobs, objects = [], []
    for src in LIB_SOURCES:
        path = os.path.join(CSRC, src)
        if not os.path.exists(path):
            continue
        obj = os.path.join(OBJ, os.path.splitext(src)[0] + ".o")
        objects.append(obj)
        deps: set[str] = set()
        _includes(path, deps)
        if force or extra or _stale(obj, list(deps) + [__file__]):
            jobs.append([NVCC, "-ccbin", HOST_CXX, "-std=c++17", "-O3", "-lineinfo", *ARCH,
                         "-Xcompiler", "-fPIC,-O3,-Wall,-ffp-contract=off", "-Xptxas", "-v" if verbose else "-O3",
                         *extra, "-I", INCLUDE, "-c", path, "-o", obj])
    if jobs:
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 2)) as pool:
            list(pool.map(lambda c: _run(c, verbose), jobs))
    if jobs or not os.path.exists(LIB):
        _run([NVCC, "-ccbin", HOST_CXX, *ARCH, "--shared", "-cudart", "static", "-o", LIB, *objects, "-ldl"], verbose)
    ext_deps = [os.path.join(CSRC, "pybind_module.cpp"), os.path.join(INCLUDE, "psvwave_b200.h"), __file__]
    if force or _stale(EXT, ext_deps + [LIB]):
        import pybind11
        _run([HOST_CXX, "-std=c++17", "-O2", "-shared", "-fPIC", "-fvisibility=hidden",
              "-I", pybind11.get_include(), "-I", sysconfig.get_paths()["include"], "-I", INCLUDE,
              os.path.join(CSRC, "pybind_module.cpp"), "-o", EXT,
              "-L", HERE, "-lpsvwave_b200", "-Wl,-rpath,$ORIGIN"], verbose)


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="--quiet" not in sys.argv)
    print("built", LIB, "and", EXT)
