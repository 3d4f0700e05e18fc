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
