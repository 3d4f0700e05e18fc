"""Python mirror of the C++ driver `class Havac` (havac_b200/host/Havac.hpp, itself keeping the
reference's host/Havac.hpp:42-107 API): same method names, argument meaning and error
behaviour, bound through the C doors of include/havac_host.h with ctypes.

Exceptions: std::length_error -> ValueError, std::logic_error -> LogicError,
std::runtime_error -> RuntimeError, std::bad_alloc -> MemoryError. There is no CPU fallback:
constructing `Havac` without an sm_100 GPU raises.

The module-level functions at the bottom need no GPU; they expose the input and output
contracts either side of the kernel (FASTA -> packed buffer, .hmm -> int8 table, raw hit ->
HavacHit) so they can be checked against the reference on a CPU box.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import List, Optional, Tuple

import numpy as np

from . import build
from .client import LogicError

HAVAC_CMD_STATE_NEW, HAVAC_CMD_STATE_QUEUED, HAVAC_CMD_STATE_RUNNING, HAVAC_CMD_STATE_COMPLETED, \
    HAVAC_CMD_STATE_ERROR, HAVAC_CMD_STATE_ABORT, HAVAC_CMD_STATE_SUBMITTED, HAVAC_CMD_STATEIMEOUT, \
    HAVAC_CMD_STATE_NORESPONSE = range(1, 10)
xclbinSrcDefault = "../device/bin/havac.xclbin"

EXPORTS = [
    "havac_host_create", "havac_host_destroy", "havac_host_load_sequence", "havac_host_load_phmm", "havac_host_run",
    "havac_host_run_async", "havac_host_wait", "havac_host_abort", "havac_host_state", "havac_host_get_hits",
    "havac_host_device_time_ms", "havac_host_last_error", "havac_host_pack_fasta", "havac_host_project_phmm",
    "havac_host_scale_factor", "havac_host_resolve_hits", "havac_host_srand", "havac_host_sequence_table",
    "havac_host_read_sequence", "havac_host_read_phmm", "havac_host_parse_phmm",
    "havac_host_model_names", "havac_host_use_reverse_complement", "havac_host_reverse_complement_packed",
]

_lib = None


def load_library(path: Optional[str] = None) -> C.CDLL:
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or build.HOST_LIB
    if not os.path.exists(p):
        raise RuntimeError(f"{p} is missing: build it with `python -m havac_b200.build`")
    L = C.CDLL(p)
    vp, u64, u32 = C.c_void_p, C.c_uint64, C.c_uint32
    L.havac_host_create.argtypes = [u32, C.c_float, C.c_char_p, C.POINTER(vp)]
    L.havac_host_destroy.argtypes = [vp]
    L.havac_host_destroy.restype = None
    for name in ("havac_host_load_sequence", "havac_host_load_phmm"):
        getattr(L, name).argtypes = [vp, C.c_char_p]
    for name in ("havac_host_run", "havac_host_run_async", "havac_host_wait", "havac_host_abort"):
        getattr(L, name).argtypes = [vp]
    L.havac_host_srand.argtypes = [C.c_uint]
    L.havac_host_srand.restype = None
    L.havac_host_sequence_table.argtypes = [vp, vp, u64, C.POINTER(u64)]
    L.havac_host_read_sequence.argtypes = [vp, vp, u64, C.POINTER(u64)]
    L.havac_host_read_phmm.argtypes = [vp, vp, u64, C.POINTER(u64)]
    L.havac_host_model_names.argtypes = [C.c_char_p, vp, u64, C.POINTER(u64)]
    L.havac_host_use_reverse_complement.argtypes = [vp, C.c_int]
    L.havac_host_reverse_complement_packed.argtypes = [vp, u64, u64]
    L.havac_host_state.argtypes = [vp, C.POINTER(C.c_int)]
    L.havac_host_get_hits.argtypes = [vp, vp, vp, vp, vp, u64, C.POINTER(u64)]
    L.havac_host_device_time_ms.argtypes = [vp, C.POINTER(C.c_float)]
    L.havac_host_last_error.argtypes = [vp]
    L.havac_host_last_error.restype = C.c_char_p
    L.havac_host_pack_fasta.argtypes = [C.c_char_p, C.c_uint, vp, u64, C.POINTER(u64), vp, u64, C.POINTER(u64)]
    L.havac_host_project_phmm.argtypes = [C.c_char_p, C.c_float, vp, u64, C.POINTER(u64), vp, u32, C.POINTER(u32)]
    L.havac_host_parse_phmm.argtypes = [C.c_char_p, vp, u64, C.POINTER(u64)]
    L.havac_host_scale_factor.argtypes = [C.c_float, C.c_float, u32, u32, C.c_float]
    L.havac_host_scale_factor.restype = C.c_float
    L.havac_host_resolve_hits.argtypes = [vp, u64, vp, u64, vp, u32, vp, vp, vp, vp, C.POINTER(u64)]
    if path is None:
        _lib = L
    return L


def _raise(rc: int, msg: str):
    if rc == -2:
        raise ValueError(msg)  # std::length_error
    if rc == -3:
        raise LogicError(msg)
    if rc == -5:
        raise MemoryError(msg)
    if rc == -1:
        raise TypeError(msg or "invalid argument")
    raise RuntimeError(msg)


@dataclass(frozen=True, order=True)
class HavacHit:
    """host/Havac.hpp:31-39; constructor order (sequencePosition, sequenceIndex, phmmPosition, phmmIndex)."""
    sequencePosition: int
    sequenceIndex: int
    phmmPosition: int
    phmmIndex: int

    def toString(self) -> str:  # host/Havac.cpp:202-207
        return (f"sequence ${self.sequenceIndex}, position {self.sequencePosition}; "
                f"phmm #{self.phmmIndex} position {self.phmmPosition}")


class Havac:
    def __init__(self, deviceIndex: int = 0, requiredPValue: float = 0.02, xclbinSrc: str = xclbinSrcDefault):
        self._L = load_library()
        self._h = C.c_void_p()
        rc = self._L.havac_host_create(deviceIndex, requiredPValue, xclbinSrc.encode(), C.byref(self._h))
        if rc:
            self._h = C.c_void_p()
            _raise(rc, (self._L.havac_host_last_error(None) or b"").decode())

    def _ck(self, rc: int):
        if rc:
            _raise(rc, (self._L.havac_host_last_error(self._h) or b"").decode())

    def loadSequence(self, fastaSrc: str) -> None:
        self._ck(self._L.havac_host_load_sequence(self._h, os.fspath(fastaSrc).encode()))

    def loadPhmm(self, phmmSrc: str) -> None:
        self._ck(self._L.havac_host_load_phmm(self._h, os.fspath(phmmSrc).encode()))

    def runHardwareClient(self) -> None:
        self._ck(self._L.havac_host_run(self._h))

    def runHardwareClientAsync(self) -> None:
        self._ck(self._L.havac_host_run_async(self._h))

    def waitHardwareClientAsync(self) -> None:
        self._ck(self._L.havac_host_wait(self._h))

    def abortHardwareClient(self) -> None:
        self._ck(self._L.havac_host_abort(self._h))

    def currentHardwareState(self) -> int:
        st = C.c_int(0)
        self._ck(self._L.havac_host_state(self._h, C.byref(st)))
        return st.value

    def getHitsFromFinishedRun(self) -> List[HavacHit]:
        a = self.getHitArraysFromFinishedRun()
        return [HavacHit(int(p), int(s), int(pp), int(pi)) for p, s, pp, pi in zip(*a)]

    def getHitArraysFromFinishedRun(self) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
        """Same hits as four parallel arrays (sequencePosition, sequenceIndex, phmmPosition, phmmIndex)."""
        n = C.c_uint64(0)
        self._ck(self._L.havac_host_get_hits(self._h, None, None, None, None, 0, C.byref(n)))
        sp = np.zeros(n.value, dtype=np.uint64)
        si, pp, pi = (np.zeros(n.value, dtype=np.uint32) for _ in range(3))
        if n.value:
            self._ck(self._L.havac_host_get_hits(self._h, sp.ctypes.data, si.ctypes.data, pp.ctypes.data,
                                                 pi.ctypes.data, n.value, C.byref(n)))
        return sp, si, pp, pi

    def sequenceEndPositions(self) -> np.ndarray:
        """FastaVector's sequenceEndPosition per loaded record (harness door)."""
        n = C.c_uint64(0)
        self._ck(self._L.havac_host_sequence_table(self._h, None, 0, C.byref(n)))
        ends = np.zeros(n.value, dtype=np.uint64)
        self._ck(self._L.havac_host_sequence_table(self._h, ends.ctypes.data, len(ends), C.byref(n)))
        return ends

    def devicePackedSequence(self) -> np.ndarray:
        """The 2-bit packed sequence as the device holds it, copied back (harness door)."""
        n = C.c_uint64(0)
        self._ck(self._L.havac_host_read_sequence(self._h, None, 0, C.byref(n)))
        buf = np.zeros(n.value, dtype=np.uint8)
        self._ck(self._L.havac_host_read_sequence(self._h, buf.ctypes.data, buf.nbytes, C.byref(n)))
        return buf

    def useReverseComplement(self, crick: bool) -> None:
        """True: search the reverse complement of the loaded sequence (hits at forward coordinates)."""
        self._ck(self._L.havac_host_use_reverse_complement(self._h, 1 if crick else 0))

    def devicePhmmScores(self) -> np.ndarray:
        """The int8 score table as the device holds it, [row][A,C,G,T], copied back (harness door)."""
        n = C.c_uint64(0)
        self._ck(self._L.havac_host_read_phmm(self._h, None, 0, C.byref(n)))
        buf = np.zeros(n.value, dtype=np.int8)
        self._ck(self._L.havac_host_read_phmm(self._h, buf.ctypes.data, buf.nbytes, C.byref(n)))
        return buf.reshape(-1, 4)

    def lastDeviceTimeMs(self) -> float:
        ms = C.c_float(0)
        self._ck(self._L.havac_host_device_time_ms(self._h, C.byref(ms)))
        return ms.value

    def close(self) -> None:
        if getattr(self, "_h", None) and self._h.value:
            self._L.havac_host_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def srand(seed: int) -> None:
    """Seed the C library's rand() of the host driver (it resolves non-ACGT symbols)."""
    load_library().havac_host_srand(seed)


# ---- device-free contracts ---------------------------------------------------------------------

def pack_fasta(fastaSrc: str, seed: int = 1) -> Tuple[np.ndarray, np.ndarray]:
    """FASTA -> (packed 2-bit buffer, record end offsets); srand(seed) first."""
    L = load_library()
    nb, ns = C.c_uint64(0), C.c_uint64(0)
    rc = L.havac_host_pack_fasta(os.fspath(fastaSrc).encode(), seed, None, 0, C.byref(nb), None, 0, C.byref(ns))
    if rc:
        _raise(rc, (L.havac_host_last_error(None) or b"").decode())
    packed = np.zeros(nb.value, dtype=np.uint8)
    ends = np.zeros(ns.value, dtype=np.uint64)
    rc = L.havac_host_pack_fasta(os.fspath(fastaSrc).encode(), seed, packed.ctypes.data, packed.nbytes, C.byref(nb),
                                 ends.ctypes.data, len(ends), C.byref(ns))
    if rc:
        _raise(rc, (L.havac_host_last_error(None) or b"").decode())
    return packed, ends


def project_phmm(phmmSrc: str, pValue: float = 0.02) -> Tuple[np.ndarray, np.ndarray]:
    """.hmm -> (int8 table [rows, 4], model lengths)."""
    L = load_library()
    nb, nm = C.c_uint64(0), C.c_uint32(0)
    rc = L.havac_host_project_phmm(os.fspath(phmmSrc).encode(), pValue, None, 0, C.byref(nb), None, 0, C.byref(nm))
    if rc:
        _raise(rc, (L.havac_host_last_error(None) or b"").decode())
    table = np.zeros(nb.value, dtype=np.int8)
    lens = np.zeros(nm.value, dtype=np.uint32)
    rc = L.havac_host_project_phmm(os.fspath(phmmSrc).encode(), pValue, table.ctypes.data, table.nbytes, C.byref(nb),
                                   lens.ctypes.data, len(lens), C.byref(nm))
    if rc:
        _raise(rc, (L.havac_host_last_error(None) or b"").decode())
    return table.reshape(-1, 4), lens


def parse_phmm(phmmSrc: str) -> np.ndarray:
    """.hmm -> match emissions as the reader parsed them, float32 [rows, 4]."""
    L = load_library()
    n = C.c_uint64(0)
    rc = L.havac_host_parse_phmm(os.fspath(phmmSrc).encode(), None, 0, C.byref(n))
    if rc:
        _raise(rc, (L.havac_host_last_error(None) or b"").decode())
    e = np.zeros(n.value, dtype=np.float32)
    rc = L.havac_host_parse_phmm(os.fspath(phmmSrc).encode(), e.ctypes.data, len(e), C.byref(n))
    if rc:
        _raise(rc, (L.havac_host_last_error(None) or b"").decode())
    return e.reshape(-1, 4)


def reverse_complement_packed(packed, nSymbols: int) -> np.ndarray:
    """Reverse complement of a 2-bit packed buffer over its first nSymbols symbols (host form)."""
    L = load_library()
    buf = np.array(packed, dtype=np.uint8, copy=True)
    rc = L.havac_host_reverse_complement_packed(buf.ctypes.data, buf.nbytes, nSymbols)
    if rc:
        _raise(rc, (L.havac_host_last_error(None) or b"").decode())
    return buf


def model_names(phmmSrc: str) -> List[Tuple[str, str]]:
    """.hmm -> [(NAME, ACC)] in file order; ACC is "" where the model has none."""
    L = load_library()
    n = C.c_uint64(0)
    rc = L.havac_host_model_names(os.fspath(phmmSrc).encode(), None, 0, C.byref(n))
    if rc:
        _raise(rc, (L.havac_host_last_error(None) or b"").decode())
    buf = C.create_string_buffer(max(1, n.value))
    rc = L.havac_host_model_names(os.fspath(phmmSrc).encode(), buf, n.value, C.byref(n))
    if rc:
        _raise(rc, (L.havac_host_last_error(None) or b"").decode())
    lines = buf.raw[:n.value].decode().split("\n")[:-1]
    return [tuple((line.split("\t") + [""])[:2]) for line in lines]


def scale_factor(mu: float, lam: float, maxLength: int, modelLength: int, pValue: float) -> float:
    return float(load_library().havac_host_scale_factor(mu, lam, maxLength, modelLength, pValue))


def resolve_hits(raw, sequenceEnd, modelLength) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
    """Raw u64 device hits -> (sequencePosition, sequenceIndex, phmmPosition, phmmIndex) arrays."""
    L = load_library()
    raw = np.ascontiguousarray(raw, dtype=np.uint64)
    ends = np.ascontiguousarray(sequenceEnd, dtype=np.uint64)
    lens = np.ascontiguousarray(modelLength, dtype=np.uint32)
    sp = np.zeros(len(raw), dtype=np.uint64)
    si, pp, pi = (np.zeros(len(raw), dtype=np.uint32) for _ in range(3))
    n = C.c_uint64(0)
    rc = L.havac_host_resolve_hits(raw.ctypes.data, len(raw), ends.ctypes.data, len(ends), lens.ctypes.data, len(lens),
                                   sp.ctypes.data, si.ctypes.data, pp.ctypes.data, pi.ctypes.data, C.byref(n))
    if rc:
        _raise(rc, (L.havac_host_last_error(None) or b"").decode())
    k = n.value
    return sp[:k], si[:k], pp[:k], pi[:k]
