"""In-tree builds: the CUDA device client (libhavac_cuda.so, sm_100a only) and the host-side
C++ driver (libhavac_host.so). nvcc cross-compiles without a GPU. Artefacts land in
havac_b200/lib/ so they travel to the GPU box with the repo snapshot.

The SSV kernels take one step more than `nvcc -c`: after ptxas has produced the cubin, the unrolled row block of the
fast kernels is re-scheduled at SASS level (havac_b200/sass/reblock.py -- order, temporaries, .reuse flags and
scheduling-control bits that neither CUDA C++ nor PTX can pin; +25 % on the B200), and the object is finished from
the patched cubin with nvcc's own fatbinary / host-compile commands. HAVAC_NO_REBLOCK=1 skips that step (the
library then carries ptxas's schedule; used for A/B measurements)."""
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

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC"]
NVCC_LINK_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-cudart", "static", "-Xcompiler", "-fPIC"]
OBJ_DIR = os.path.join(ROOT, "build", "cuda")
RESCHEDULED = "ssv_kernels.cu"  # the translation unit whose cubin goes through havac_b200/sass/reblock.py
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


def _compile_rescheduled(nvcc, cu, obj, work, defines, verbose):
    """nvcc -c with the cubin patched between ptxas and fatbinary. The command lines of the two steps that follow
    ptxas are nvcc's own (taken from `nvcc -dryrun --keep`), re-run after the patch."""
    from havac_b200.sass import reblock

    os.makedirs(work, exist_ok=True)
    base = [nvcc, "--keep", "--keep-dir", work, *NVCC_FLAGS, *defines, "-I", INCLUDE, "-c", "-o", obj, cu]
    dry = subprocess.run(base[:1] + ["-dryrun"] + base[1:], check=True, capture_output=True, text=True).stderr.splitlines()
    env = dict(os.environ)
    steps = []
    for line in dry:
        if not line.startswith("#$ "):
            continue
        line = line[3:]
        m = __import__("re").match(r"^([A-Za-z_][A-Za-z0-9_]*)=(.*)$", line)
        if m and " " not in m.group(1) and not line.startswith(("gcc", "ptxas", "fatbinary", "cicc", "cudafe", "rm", '"')):
            env[m.group(1)] = os.path.expandvars(m.group(2).strip()) if m.group(1) != "PATH" else m.group(2).strip()
        else:
            steps.append(line)
    env["PATH"] = env.get("PATH", "") + os.pathsep + os.environ.get("PATH", "")
    subprocess.run(base + (["-Xptxas=-v"] if verbose else []), check=True)
    cubin = os.path.join(work, os.path.basename(cu)[:-3] + ".cubin")
    report = reblock.patch_cubin(cubin, cubin, verbose=verbose)
    done = [(k, v) for k, v in report.items() for b in v if b[2] == "rescheduled"]
    if not done:
        raise RuntimeError(f"reblock: no block of {cu} could be re-scheduled: {report}")
    after = [c for c in steps[next(n for n, c in enumerate(steps) if c.startswith("ptxas")) + 1:] if not c.startswith("rm ")]
    for c in after:
        subprocess.run(c, shell=True, check=True, env=env, cwd=work, executable="/bin/bash")
    return report


def build_cuda(force: bool = False, verbose: bool = False, out: str = None, reblock: bool = None, defines=()) -> str:
    """libhavac_cuda.so (or a variant at `out`): every .cu compiled to an object, the SSV kernels' cubin re-scheduled
    on the way (unless reblock=False / HAVAC_NO_REBLOCK=1), then linked."""
    os.makedirs(LIB, exist_ok=True)
    target = out or CUDA_LIB
    if reblock is None:
        reblock = os.environ.get("HAVAC_NO_REBLOCK") != "1"
    cu = _sources(CSRC, (".cu",))
    deps = cu + _sources(CSRC, (".cuh",)) + _sources(INCLUDE, (".h",)) + _sources(os.path.join(PKG, "sass"), (".py",)) + [__file__]
    if force or out or _stale(target, deps):
        nvcc = _nvcc()
        tag = os.path.basename(target).replace(".so", "")
        work = os.path.join(OBJ_DIR, tag)
        os.makedirs(work, exist_ok=True)
        objs = []
        for src in cu:
            obj = os.path.join(work, os.path.basename(src)[:-3] + ".o")
            if reblock and os.path.basename(src) == RESCHEDULED:
                _compile_rescheduled(nvcc, src, obj, os.path.join(work, "keep"), list(defines), verbose)
            else:
                cmd = [nvcc, *NVCC_FLAGS, *defines, "-I", INCLUDE, "-c", "-o", obj, src]
                if verbose:
                    cmd.insert(1, "-Xptxas=-v")
                subprocess.run(cmd, check=True)
            objs.append(obj)
        subprocess.run([nvcc, *NVCC_LINK_FLAGS, "-o", target, *objs], check=True)
    return target


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
