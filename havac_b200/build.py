"""In-tree builds: the CUDA device client (libhavac_cuda.so, sm_100a only) and the host-side
C++ driver (libhavac_host.so). nvcc cross-compiles without a GPU. Artefacts land in
havac_b200/lib/ so they travel to the GPU box with the repo snapshot."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
LIB = os.path.join(PKG, "lib")
CSRC = os.path.join(PKG, "csrc")
HOST = os.path.join(PKG, "host")
INCLUDE = os.path.join(ROOT, "include")

CUDA_LIB = os.path.join(LIB, "libhavac_cuda.so")
HOST_LIB = os.path.join(LIB, "libhavac_host.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared", "-cudart", "static"]
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-fno-fast-math", "-Wall"]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: the CUDA device client cannot be built")


def _stale(target: str, sources) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _sources(directory: str, exts) -> list:
    return sorted(os.path.join(directory, f) for f in os.listdir(directory) if f.endswith(tuple(exts)))


def build_cuda(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(LIB, exist_ok=True)
    cu = _sources(CSRC, (".cu",))
    deps = cu + _sources(CSRC, (".cuh",)) + _sources(INCLUDE, (".h",))
    if force or _stale(CUDA_LIB, deps):
        cmd = [_nvcc(), *NVCC_FLAGS, "-I", INCLUDE, "-o", CUDA_LIB, *cu]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        subprocess.run(cmd, check=True)
    return CUDA_LIB


def build_host(force: bool = False) -> str:
    os.makedirs(LIB, exist_ok=True)
    cpp = _sources(HOST, (".cpp",))
    deps = cpp + _sources(HOST, (".hpp", ".h")) + _sources(INCLUDE, (".h",))
    if force or _stale(HOST_LIB, deps):
        cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
        # links against the CUDA device client next to it ($ORIGIN), so build that first
        cmd = [cxx, *CXX_FLAGS, "-I", INCLUDE, "-I", HOST, "-o", HOST_LIB, *cpp, "-L", LIB, "-lhavac_cuda",
               "-Wl,-rpath,$ORIGIN", "-lpthread"]
        subprocess.run(cmd, check=True)
    return HOST_LIB


def build_all(force: bool = False) -> None:
    build_cuda(force)
    if os.path.isdir(HOST):
        build_host(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv)
    print("built:", [p for p in (CUDA_LIB, HOST_LIB) if os.path.exists(p)])
