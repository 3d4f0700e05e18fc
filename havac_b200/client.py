"""ctypes binding of the C-ABI device client (include/havac_cuda.h).

`HavacHwClient` mirrors the reference class of the same name method for method
(host/HavacHwClient.hpp:21-115): writeSequence, writePhmm, invokeHavacSsvAsync,
getHwState, waitForHavacSsvAsync, abort, getHitList -- with the same error
behaviour mapped onto Python exceptions (length_error -> ValueError,
logic_error -> RuntimeError subclass LogicError, runtime_error -> RuntimeError,
bad_alloc -> MemoryError). There is no CPU fallback: if the CUDA library is
missing or no sm_100 device is present, construction raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

from . import build

OK = 0
E_INVALID, E_LENGTH, E_LOGIC, E_RUNTIME, E_NOMEM, E_NODEVICE, E_OVERFLOW = -1, -2, -3, -4, -5, -6, -7
STATE_NEW, STATE_QUEUED, STATE_RUNNING, STATE_COMPLETED, STATE_ERROR, STATE_ABORT, STATE_SUBMITTED, \
    STATE_TIMEOUT, STATE_NORESPONSE = range(1, 10)
SEGMENT_SYMBOLS = 12288

EXPORTS = [
    "havac_cuda_create", "havac_cuda_destroy", "havac_cuda_write_sequence", "havac_cuda_write_phmm",
    "havac_cuda_set_segment_range", "havac_cuda_invoke_async", "havac_cuda_state", "havac_cuda_wait",
    "havac_cuda_abort", "havac_cuda_num_hits", "havac_cuda_read_hits", "havac_cuda_last_error",
    "havac_cuda_device_time_ms", "havac_cuda_device_count", "havac_cuda_run_stats", "havac_cuda_set_kernel", "havac_cuda_transfer_stats", "havac_cuda_sm_cycles",
    "havac_cuda_ingest_fasta", "havac_cuda_ingest_records", "havac_cuda_read_sequence", "havac_cuda_resolve_hits", "havac_cuda_project_phmm", "havac_cuda_read_phmm",
    "havac_cuda_reverse_complement", "havac_cuda_strand_state", "havac_cuda_tile_geometry", "havac_cuda_cell_selftest",
    "havac_cuda_trace_cells", "havac_cuda_debug_poke_phmm", "havac_cuda_sort_hits_static",
]

# one cell of the per-cell trace (include/havac_cuda.h havac_cuda_cell_record; byCellComparator.hpp:23-31)
CELL_RECORD = np.dtype([("prevValue", np.uint8), ("matchScore", np.int8), ("cellValue", np.uint8), ("phmmVector", np.int8, (4,)),
                        ("symbol", np.uint8), ("passesThreshold", np.uint8), ("pad", np.uint8, (3,))])


class LogicError(RuntimeError):
    """std::logic_error of the reference (calls made out of order)."""


class HitOverflowError(RuntimeError):
    pass


_lib = None


def load_library(path: Optional[str] = None) -> C.CDLL:
    """Load libhavac_cuda.so (no GPU needed just to load it). Fails loudly when absent."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or build.CUDA_LIB
    if not os.path.exists(p):
        raise RuntimeError(f"{p} is missing: build it with `python -m havac_b200.build` "
                           "(there is no CPU fallback for the SSV path)")
    L = C.CDLL(p)
    vp, u64, u32 = C.c_void_p, C.c_uint64, C.c_uint32
    L.havac_cuda_create.argtypes = [C.c_char_p, u32, C.POINTER(vp)]
    L.havac_cuda_destroy.argtypes = [vp]
    L.havac_cuda_write_sequence.argtypes = [vp, vp, u64]
    L.havac_cuda_write_phmm.argtypes = [vp, vp, u64]
    L.havac_cuda_project_phmm.argtypes = [vp, vp, u64, vp, vp, u32]
    L.havac_cuda_read_phmm.argtypes = [vp, vp, u64, C.POINTER(u64)]
    L.havac_cuda_read_sequence.argtypes = [vp, vp, u64, C.POINTER(u64)]
    L.havac_cuda_reverse_complement.argtypes = [vp, u64]
    L.havac_cuda_strand_state.argtypes = [vp, C.POINTER(u64)]
    L.havac_cuda_set_segment_range.argtypes = [vp, u64, u64]
    L.havac_cuda_invoke_async.argtypes = [vp]
    L.havac_cuda_state.argtypes = [vp, C.POINTER(C.c_int)]
    L.havac_cuda_wait.argtypes = [vp, u64, C.POINTER(C.c_int)]
    L.havac_cuda_abort.argtypes = [vp, C.POINTER(C.c_int)]
    L.havac_cuda_num_hits.argtypes = [vp, C.POINTER(u64)]
    L.havac_cuda_read_hits.argtypes = [vp, vp, u64, C.POINTER(u64)]
    L.havac_cuda_last_error.argtypes = [vp]
    L.havac_cuda_last_error.restype = C.c_char_p
    L.havac_cuda_device_time_ms.argtypes = [vp, C.POINTER(C.c_float)]
    L.havac_cuda_device_count.argtypes = [vp, C.POINTER(C.c_int)]
    L.havac_cuda_run_stats.argtypes = [vp, C.POINTER(u64), C.POINTER(u32)]
    L.havac_cuda_set_kernel.argtypes = [vp, C.c_char_p]
    L.havac_cuda_sm_cycles.argtypes = [vp, C.POINTER(u64)]
    L.havac_cuda_transfer_stats.argtypes = [vp, C.POINTER(u64), C.POINTER(u64)]
    L.havac_cuda_cell_selftest.argtypes = [vp, vp, vp]
    L.havac_cuda_trace_cells.argtypes = [vp, u32, u32, u64, u64, vp, u64, C.POINTER(u64)]
    L.havac_cuda_debug_poke_phmm.argtypes = [vp, u64, C.c_int8]
    L.havac_cuda_tile_geometry.argtypes = [vp, u32, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(u32), C.POINTER(u32),
                                           C.POINTER(u32), C.POINTER(u32)]
    L.havac_cuda_sort_hits_static.argtypes = [u32, vp, u64, vp, vp]
    for name in EXPORTS:
        if name != "havac_cuda_last_error":
            getattr(L, name).restype = C.c_int
    if path is None:
        _lib = L
    return L


class HavacHwClient:
    def __init__(self, xclbinFileSrc: str = "", havacKernelName: str = "HavacKernel", deviceIndex: int = 0):
        self._L = load_library()
        self._h = C.c_void_p()
        rc = self._L.havac_cuda_create((xclbinFileSrc or "").encode(), deviceIndex, C.byref(self._h))
        if rc != OK:
            msg = (self._L.havac_cuda_last_error(None) or b"").decode()
            self._h = C.c_void_p()
            self._raise(rc, msg)
        self.kernelName = havacKernelName

    # -- error mapping ---------------------------------------------------------------------
    def _raise(self, rc: int, msg: Optional[str] = None):
        if msg is None:
            msg = (self._L.havac_cuda_last_error(self._h) or b"").decode()
        if rc == E_LENGTH:
            raise ValueError(msg)  # std::length_error
        if rc == E_LOGIC:
            raise LogicError(msg)
        if rc == E_NOMEM:
            raise MemoryError(msg)
        if rc == E_OVERFLOW:
            raise HitOverflowError(msg)
        raise RuntimeError(f"[havac_cuda rc={rc}] {msg}")

    def _ck(self, rc: int):
        if rc != OK:
            self._raise(rc)

    # -- the reference client's methods ------------------------------------------------------
    def writeSequence(self, compressedSequence) -> None:
        a = np.ascontiguousarray(compressedSequence, dtype=np.uint8)
        self._ck(self._L.havac_cuda_write_sequence(self._h, a.ctypes.data, a.nbytes))

    def writePhmm(self, phmmAsFlattenedArray) -> None:
        a = np.ascontiguousarray(phmmAsFlattenedArray, dtype=np.int8).reshape(-1)
        self._ck(self._L.havac_cuda_write_phmm(self._h, a.ctypes.data, a.nbytes))

    def projectPhmm(self, negLogEmissions, modelLengths, scaleFactors) -> None:
        """Score table built on the device from [rows, 4] float emissions (havac_cuda_project_phmm)."""
        e = np.ascontiguousarray(negLogEmissions, dtype=np.float32).reshape(-1, 4)
        lens = np.ascontiguousarray(modelLengths, dtype=np.uint32)
        scale = np.ascontiguousarray(scaleFactors, dtype=np.float32)
        if len(lens) != len(scale):
            raise ValueError("one scale factor per model")
        self._ck(self._L.havac_cuda_project_phmm(self._h, e.ctypes.data, e.shape[0], lens.ctypes.data, scale.ctypes.data, len(lens)))

    def readPhmm(self) -> np.ndarray:
        n = C.c_uint64(0)
        self._ck(self._L.havac_cuda_read_phmm(self._h, None, 0, C.byref(n)))
        buf = np.zeros(n.value, dtype=np.int8)
        self._ck(self._L.havac_cuda_read_phmm(self._h, buf.ctypes.data, buf.nbytes, C.byref(n)))
        return buf.reshape(-1, 4)

    def readSequence(self) -> np.ndarray:
        """The packed sequence as the (single) device holds it, copied back."""
        n = C.c_uint64(0)
        self._ck(self._L.havac_cuda_read_sequence(self._h, None, 0, C.byref(n)))
        buf = np.zeros(n.value, dtype=np.uint8)
        self._ck(self._L.havac_cuda_read_sequence(self._h, buf.ctypes.data, buf.nbytes, C.byref(n)))
        return buf

    def reverseComplement(self, nSymbols: int) -> None:
        """Resident sequence <-> its reverse complement over the first nSymbols symbols, on the device."""
        self._ck(self._L.havac_cuda_reverse_complement(self._h, nSymbols))

    def mirroredSymbols(self) -> int:
        n = C.c_uint64(0)
        self._ck(self._L.havac_cuda_strand_state(self._h, C.byref(n)))
        return n.value

    def invokeHavacSsvAsync(self) -> None:
        self._ck(self._L.havac_cuda_invoke_async(self._h))

    def getHwState(self) -> int:
        st = C.c_int(0)
        self._ck(self._L.havac_cuda_state(self._h, C.byref(st)))
        return st.value

    def waitForHavacSsvAsync(self, timeout_ms: int = 0) -> int:
        st = C.c_int(0)
        self._ck(self._L.havac_cuda_wait(self._h, timeout_ms, C.byref(st)))
        return st.value

    def abort(self) -> int:
        st = C.c_int(0)
        self._ck(self._L.havac_cuda_abort(self._h, C.byref(st)))
        return st.value

    def getNumHits(self) -> int:
        n = C.c_uint64(0)
        self._ck(self._L.havac_cuda_num_hits(self._h, C.byref(n)))
        return n.value

    def getHitList(self) -> np.ndarray:
        n = self.getNumHits()
        out = np.zeros(n, dtype=np.uint64)
        got = C.c_uint64(0)
        self._ck(self._L.havac_cuda_read_hits(self._h, out.ctypes.data, n, C.byref(got)))
        return out[: got.value]

    # -- harness helpers (no reference counterpart) ----------------------------------------------
    def setSegmentRange(self, segBegin: int, segEnd: int = 0) -> None:
        self._ck(self._L.havac_cuda_set_segment_range(self._h, segBegin, segEnd))

    def setKernel(self, name: str) -> None:
        self._ck(self._L.havac_cuda_set_kernel(self._h, name.encode()))

    def deviceTimeMs(self) -> float:
        ms = C.c_float(0)
        self._ck(self._L.havac_cuda_device_time_ms(self._h, C.byref(ms)))
        return ms.value

    def deviceCount(self) -> int:
        n = C.c_int(0)
        self._ck(self._L.havac_cuda_device_count(self._h, C.byref(n)))
        return n.value

    def runStats(self):
        cells, launches = C.c_uint64(0), C.c_uint32(0)
        self._ck(self._L.havac_cuda_run_stats(self._h, C.byref(cells), C.byref(launches)))
        return cells.value, launches.value

    def smCycles(self) -> int:
        n = C.c_uint64(0)
        self._ck(self._L.havac_cuda_sm_cycles(self._h, C.byref(n)))
        return n.value

    def transferStats(self):
        a, b = C.c_uint64(0), C.c_uint64(0)
        self._ck(self._L.havac_cuda_transfer_stats(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def traceCells(self, rowBegin: int, rowEnd: int, colBegin: int, colEnd: int) -> np.ndarray:
        """Per-cell records [rowEnd - rowBegin, colEnd - colBegin] of the window, computed on the device."""
        out = np.zeros((rowEnd - rowBegin, colEnd - colBegin), dtype=CELL_RECORD)
        n = C.c_uint64(0)
        self._ck(self._L.havac_cuda_trace_cells(self._h, rowBegin, rowEnd, colBegin, colEnd, out.ctypes.data, out.size, C.byref(n)))
        return out

    def debugPokePhmm(self, byteOffset: int, value: int) -> None:
        self._ck(self._L.havac_cuda_debug_poke_phmm(self._h, byteOffset, value))

    def cellSelftest(self):
        """(score[2][65536], hit[2][65536]) of the fast kernel's cell update on every (prev, emission) pair."""
        score = np.zeros(2 * 65536, dtype=np.uint8)
        hit = np.zeros(2 * 65536, dtype=np.uint8)
        self._ck(self._L.havac_cuda_cell_selftest(self._h, score.ctypes.data, hit.ctypes.data))
        return score.reshape(2, 65536), hit.reshape(2, 65536)

    def tileGeometry(self, device: int = 0) -> dict:
        """How the last launch cut its work (havac_cuda_tile_geometry): for parity tests that aim at edges."""
        lb, db = C.c_int64(0), C.c_int64(0)
        nt, td, rc, br = C.c_uint32(0), C.c_uint32(0), C.c_uint32(0), C.c_uint32(0)
        self._ck(self._L.havac_cuda_tile_geometry(self._h, device, C.byref(lb), C.byref(db), C.byref(nt), C.byref(td), C.byref(rc),
                                                  C.byref(br)))
        return {"loadBegin": lb.value, "diagBegin": db.value, "nTiles": nt.value, "tileDiagonals": td.value,
                "rowChunk": rc.value, "blockRows": br.value}

    def close(self) -> None:
        if getattr(self, "_h", None) and self._h.value:
            self._L.havac_cuda_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
