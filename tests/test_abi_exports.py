"""The drop-in boundary is a C-ABI: both shared libraries must load on a machine without a GPU
and export every function their headers in include/ declare. No compute call is made here."""
import ctypes as C
import os
import re

import pytest

from havac_b200 import build, client, havac

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(havac_(?:cuda|host)_[a-z0-9_]+)\s*\(", text)))


@pytest.mark.parametrize("header,libpath,listed", [
    ("havac_cuda.h", build.CUDA_LIB, client.EXPORTS),
    ("havac_host.h", build.HOST_LIB, havac.EXPORTS),
])
def test_every_declared_symbol_is_exported(header, libpath, listed):
    build.build_all()
    names = declared(header)
    assert len(names) >= 10
    lib = C.CDLL(libpath)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"{os.path.basename(libpath)} lacks {missing}"
    assert sorted(listed) == names, "the Python binding's export list drifted from the header"


def test_no_gpu_means_loud_failure_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(RuntimeError, match="CUDA"):
        client.HavacHwClient("")
    with pytest.raises(RuntimeError, match="CUDA"):
        havac.Havac(0, 0.02)


def test_product_never_touches_the_oracle():
    """Nothing under havac_b200/ may import, link or execute oracle/ (it is test infrastructure)."""
    pkg = os.path.join(ROOT, "havac_b200")
    for base, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cpp", ".hpp", ".cu", ".cuh", ".h")):
                text = open(os.path.join(base, f), errors="ignore").read()
                assert "oracle_lib" not in text and "liboracle" not in text and "oracle/" not in text, f


def test_no_alloc_sort_has_its_scratch_in_the_library_image():
    """havac_cuda_sort_hits_static allocates nothing: its scratch is a file-scope device array of the library, of the
    size include/havac_cuda.h promises (two key buffers of HAVAC_CUDA_SORT_STATIC_MAX_HITS words, the digit counts of
    that many hits' tiles, one 256-byte slot), and the entry point's source names no allocating call."""
    import shutil
    import subprocess
    build.build_all()
    header = open(os.path.join(ROOT, "include", "havac_cuda.h")).read()
    m = re.search(r"#define HAVAC_CUDA_SORT_STATIC_MAX_HITS \(1u << (\d+)\)", header)
    assert m, "the header no longer states the maximum"
    max_hits = 1 << int(m.group(1))
    tile = 256 * 16  # STHREADS * ITEMS of hit_postprocess.cu
    want = 2 * max_hits * 8 + 256 * (max_hits // tile) * 4 + 256
    src = open(os.path.join(build.CSRC, "hit_postprocess.cu")).read()
    body = src[src.index('extern "C" int havac_cuda_sort_hits_static'):]
    body = body[:body.index("\n}\n") + 3]
    for call in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cudaMallocAsync", "new ", "std::vector"):
        assert call not in body, f"{call} inside the no-allocation entry point"
    assert "cudaGetSymbolAddress" in src and "static __device__ __align__(256)" in src
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not installed")
    ptx_or_elf = subprocess.run([cuobjdump, "-elf", build.CUDA_LIB], capture_output=True, text=True).stdout
    sizes = [int(x) for x in re.findall(r"g_sortScratchE\[(\d+)\]", ptx_or_elf)]
    if not sizes:  # no PTX in the image: take the symbol's size from the ELF symbol table instead
        sizes = [int(x, 16) for x in re.findall(r"\s(0x[0-9a-f]+)\s+0x[0-9a-f]+\s+\d+\s+0x[0-9a-f]+\s+\S*g_sortScratchE", ptx_or_elf)]
    assert sizes and all(s == want for s in sizes), (sizes, want)
