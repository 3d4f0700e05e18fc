"""ctypes doors into the CHECKERS: oracle/liboracle.so (our C restatement) and,
when it has been built, oracle/_ref/libhavac_ref.so (the unmodified reference
objects). Only tests/, __graft_entry__.smoke() and bench.py's CPU legs import
this module; nothing under havac_b200/ does."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C")
_i8p = np.ctypeslib.ndpointer(np.int8, flags="C")
_u64p = np.ctypeslib.ndpointer(np.uint64, flags="C")
_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C")
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C")


def build_oracle():
    subprocess.run(["make", "-C", ORACLE_DIR, "-s"], check=True, stdout=subprocess.DEVNULL)


def _load(path):
    return C.CDLL(path) if os.path.exists(path) else None


_oracle = None
_ref = None


def oracle():
    global _oracle
    if _oracle is None:
        p = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(p):
            build_oracle()
        L = C.CDLL(p)
        L.oracle_soft_ssv.restype = C.c_int64
        L.oracle_soft_ssv.argtypes = [_u8p, C.c_uint64, _i8p, C.c_uint64, _u64p, _u32p, C.c_uint64]
        L.oracle_replay_diagonal.restype = C.c_int64
        L.oracle_replay_diagonal.argtypes = [_u8p, C.c_uint64, _i8p, C.c_uint64, C.c_int64, _u32p, C.c_uint64]
        L.oracle_unpack_2bit.argtypes = [_u8p, C.c_uint64, _u8p]
        L.oracle_pack_2bit.argtypes = [_u8p, C.c_uint64, _u8p]
        L.oracle_encode_hit.restype = C.c_uint64
        L.oracle_encode_hit.argtypes = [C.c_uint64, C.c_uint32]
        L.oracle_resolve_hit.restype = C.c_int
        L.oracle_resolve_hit.argtypes = [C.c_uint64, _u64p, C.c_uint64, _u32p, C.c_uint32,
                                         C.POINTER(C.c_uint64), C.POINTER(C.c_uint32),
                                         C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        L.oracle_trace_cells.restype = C.c_int64
        L.oracle_trace_cells.argtypes = [_u8p, C.c_uint64, _i8p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p]
        L.oracle_cell_processor.argtypes = [C.c_uint8, C.c_int8, C.POINTER(C.c_uint8), C.POINTER(C.c_int)]
        L.oracle_gumbel_invsurv.restype = C.c_double
        L.oracle_gumbel_invsurv.argtypes = [C.c_double, C.c_double, C.c_double]
        L.oracle_scale_factor.restype = C.c_float
        L.oracle_scale_factor.argtypes = [C.c_float, C.c_float, C.c_uint32, C.c_uint32, C.c_float]
        L.oracle_project_model.argtypes = [_f32p, C.c_uint32, C.c_float, _i8p]
        _oracle = L
    return _oracle


def ref():
    """The compiled reference, or None where it has not been built (GPU box
    without a prebuilt oracle/_ref)."""
    global _ref
    if _ref is None:
        p = os.path.join(ORACLE_DIR, "_ref", "libhavac_ref.so")
        if not os.path.exists(p) and os.path.isdir("/root/reference/test/softSsv"):
            build_oracle()
        L = _load(p)
        if L is None:
            return None
        L.ref_soft_ssv.restype = C.c_int64
        L.ref_soft_ssv.argtypes = [_u8p, C.c_uint64, _i8p, C.c_uint64, _u64p, _u32p, C.c_uint64]
        L.ref_scale_factor.restype = C.c_float
        L.ref_scale_factor.argtypes = [C.c_float, C.c_float, C.c_uint32, C.c_uint32, C.c_float]
        L.ref_load_phmm.restype = C.c_int64
        L.ref_load_phmm.argtypes = [C.c_char_p, C.c_float, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint32,
                                    C.POINTER(C.c_uint32)]
        L.ref_load_fasta.restype = C.c_int64
        L.ref_load_fasta.argtypes = [C.c_char_p, C.c_uint, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                     C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.ref_local_position.restype = C.c_int
        L.ref_local_position.argtypes = [_u64p, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64),
                                         C.POINTER(C.c_uint64)]
        _ref = L
    return _ref


# ---- numpy-friendly wrappers -------------------------------------------------

def soft_ssv(symbols, table, lib=None, cap=1 << 22):
    """Run the SSV recurrence; returns (seqPos u64[], row u32[])."""
    L = lib or oracle()
    fn = L.oracle_soft_ssv if hasattr(L, "oracle_soft_ssv") else L.ref_soft_ssv
    symbols = np.ascontiguousarray(symbols, dtype=np.uint8)
    table = np.ascontiguousarray(table, dtype=np.int8).reshape(-1)
    pos = np.zeros(cap, dtype=np.uint64)
    row = np.zeros(cap, dtype=np.uint32)
    n = fn(symbols, len(symbols), table, len(table) // 4, pos, row, cap)
    if n < 0 or n > cap:
        raise RuntimeError(f"soft ssv failed or overflowed: {n}")
    return pos[:n].copy(), row[:n].copy()


def encode_hits(pos, row):
    pos = np.asarray(pos, dtype=np.uint64)
    row = np.asarray(row, dtype=np.uint64)
    seg = pos // np.uint64(12288)
    cell = pos % np.uint64(12288)
    return cell | (seg << np.uint64(14)) | (row << np.uint64(40))


def ref_load_phmm(path, p):
    L = ref()
    n = C.c_uint32(0)
    rows = L.ref_load_phmm(path.encode(), p, None, 0, None, 0, C.byref(n))
    if rows < 0:
        raise RuntimeError(f"reference readP7Hmm failed on {path}: {rows}")
    table = np.zeros(rows * 4, dtype=np.int8)
    lens = np.zeros(n.value, dtype=np.uint32)
    L.ref_load_phmm(path.encode(), p, table.ctypes.data, table.nbytes, lens.ctypes.data, n.value, C.byref(n))
    return table, lens


def ref_load_fasta(path, seed=1):
    L = ref()
    ns, cnt = C.c_uint64(0), C.c_uint64(0)
    nb = L.ref_load_fasta(path.encode(), seed, None, 0, None, 0, C.byref(ns), C.byref(cnt))
    if nb < 0:
        raise RuntimeError(f"reference fastaVectorReadFasta failed on {path}: {nb}")
    packed = np.zeros(nb, dtype=np.uint8)
    ends = np.zeros(ns.value, dtype=np.uint64)
    L.ref_load_fasta(path.encode(), seed, packed.ctypes.data, nb, ends.ctypes.data, ns.value, C.byref(ns), C.byref(cnt))
    return packed, ends, cnt.value


def soft_ssv_parallel(symbols, table, threads, lib=None, cap=1 << 20):
    """The same SSV on `threads` host threads: the sequence axis is cut into chunks, each run
    with a left halo of rows-1 symbols and its halo hits discarded (exact: a cell depends only
    on its own diagonal, which starts at row 0 at most rows-1 columns to the left; SURVEY.md
    section 8e). ctypes releases the GIL during the call, so plain threads scale."""
    from concurrent.futures import ThreadPoolExecutor

    symbols = np.ascontiguousarray(symbols, dtype=np.uint8)
    table = np.ascontiguousarray(table, dtype=np.int8).reshape(-1, 4)
    n, rows = len(symbols), table.shape[0]
    threads = max(1, min(int(threads), max(1, n // 1024)))
    edges = [n * i // threads for i in range(threads + 1)]

    def work(i):
        a, b = edges[i], edges[i + 1]
        lo = max(0, a - (rows - 1))
        pos, row = soft_ssv(symbols[lo:b], table, lib=lib, cap=cap)
        pos = pos + np.uint64(lo)
        keep = pos >= np.uint64(a)
        return pos[keep], row[keep]

    if threads == 1:
        return work(0)
    with ThreadPoolExecutor(threads) as ex:
        parts = list(ex.map(work, range(threads)))
    return np.concatenate([p for p, _ in parts]), np.concatenate([r for _, r in parts])


def replay_diagonal(packed, table, d, buf=None):
    """Rows at which diagonal d (column - row) of the whole padded buffer scores a hit, replayed on
    the CPU from row max(0,-d) with a window of the packed sequence (oracle_replay_diagonal,
    mirroring attemptTraceback of test/hmmerSsvRef/hmmerSsvRef.cpp:504-586)."""
    L = oracle()
    flat = np.ascontiguousarray(table, dtype=np.int8).reshape(-1)
    rows = len(flat) // 4
    if buf is None:
        buf = np.zeros(8192, dtype=np.uint32)
    a = max(0, d)
    b = min(len(packed) * 4, d + rows)
    lo = (a // 4) * 4
    sym = np.zeros((b - lo + 3) // 4 * 4, dtype=np.uint8)
    src = np.ascontiguousarray(packed[lo // 4:(b + 3) // 4])
    L.oracle_unpack_2bit(src, len(src) * 4, sym)
    k = L.oracle_replay_diagonal(sym, b - lo, flat, rows, d - lo, buf, len(buf))
    return set(buf[:k].tolist())


DIAG_EDGE_OFFSETS = (0, 1, 2, 31, 32, 33, 62, 63, 64)  # thread edges (32 diagonals a thread), register halves


def edge_diagonals(geom, rows, n_columns, n_tiles, seed):
    """Global diagonals (column - row) that sit on the fast kernel's STRUCTURE edges rather than anywhere:
    for the first tiles (the band left of column 0), the last tile and randomly drawn ones, the first and last
    diagonals of the tile, of its threads and of their register halves, and the middle of the tile.
    `geom` is HavacHwClient.tileGeometry() of the run being checked."""
    W = geom["tileDiagonals"]
    origin = geom["loadBegin"] + geom["diagBegin"]
    g = np.random.Generator(np.random.PCG64(seed))
    nt = geom["nTiles"]
    tiles = {0, 1, nt - 1, nt - 2} | set(g.integers(0, nt, size=n_tiles).tolist())
    offs = list(DIAG_EDGE_OFFSETS) + [W // 2 - 1, W // 2] + [W - 1 - o for o in DIAG_EDGE_OFFSETS]
    out = []
    for t in sorted(x for x in tiles if 0 <= x < nt):
        for o in offs:
            d = origin + t * W + o
            if -(rows - 1) <= d < n_columns:
                out.append(d)
    return out


def set_symbols(packed, col, syms):
    """Overwrite symbols [col, col + len(syms)) of a 2-bit packed buffer."""
    for k, x in enumerate(np.asarray(syms).tolist()):
        c = col + k
        sh = 2 * (c & 3)
        packed[c >> 2] = (int(packed[c >> 2]) & ~(3 << sh) & 0xFF) | (int(x) << sh)


def plant_hit_at(packed, table, d, r_star, g):
    """Edit the packed sequence along diagonal d so that the score crosses 256 at EXACTLY row r_star: a run of
    each row's worst symbol long enough to bring any score to 0, then each row's best symbol from a start row
    chosen (with a few demotions to the second-best symbol where the sums skip) so that the running sum is
    < 256 at r_star - 1 and >= 256 at r_star. Returns the first row edited, or None if no run fits."""
    t = np.asarray(table, dtype=np.int64).reshape(-1, 4)
    best, worst = t.argmax(axis=1), t.argmin(axis=1)
    order = np.argsort(t, axis=1)
    for attempt in range(200):
        sym = best.copy()
        if attempt:  # demote a few rows near the target to their second-best symbol: moves the crossing
            for r in g.integers(max(0, r_star - 40), r_star + 1, size=1 + attempt % 4).tolist():
                sym[r] = order[r, 2]
        sc = t[np.arange(len(t)), sym]
        for a in range(r_star, max(-1, r_star - 120), -1):
            run = sc[a:r_star + 1]
            c = np.cumsum(run)
            if (c < 0).any():
                break
            if c[-1] >= 256:
                if len(c) == 1 or c[:-1].max() < 256:
                    # zeroing run in front of row a
                    z, acc = a, 0
                    while z > 0 and acc > -256:
                        z -= 1
                        acc += t[z, worst[z]]
                    if acc > -256 or d + z < 0 or d + r_star >= len(packed) * 4:
                        return None
                    syms = np.concatenate([worst[z:a], sym[a:r_star + 1]]).astype(np.uint8)
                    set_symbols(packed, d + z, syms)
                    return z
                break
    return None


def edge_targets(geom, rows, n_columns, seed, n_tiles=5):
    """(diagonal, row) pairs for plant_hit_at: diagonals on tile / thread / register-half edges crossed with rows on
    block edges (r = 0, -1 mod blockRows), row-chunk edges (r = 0, -1 mod rowChunk) and the table's last row."""
    B, TC = geom["blockRows"], geom["rowChunk"]
    g = np.random.Generator(np.random.PCG64(seed))
    row_targets = [rows - 1]
    for k in range(1, 6):
        for base in (k * TC, (rows // TC) * TC if k == 1 else 0):
            for r in (base - 1, base):
                if 200 <= r < rows:
                    row_targets.append(r)
    for _ in range(8):
        b = int(g.integers(8, max(9, rows // B))) * B
        row_targets += [r for r in (b - 1, b) if 200 <= r < rows]
    row_targets = sorted(set(row_targets))
    diags = [d for d in edge_diagonals(geom, rows, n_columns, n_tiles, seed + 1) if d >= 0 and d + rows < n_columns]
    out, used = [], []
    k = int(g.integers(0, len(row_targets)))
    for d in diags:
        for _ in range(len(row_targets)):
            r = row_targets[k % len(row_targets)]
            k += 1
            lo, hi = d + r - 400, d + r + 2  # columns a planting may touch
            if all(hi < a or lo > b for a, b in used):
                used.append((lo, hi))
                out.append((d, r))
                break
    return out


def check_hits_by_replay(packed, table, raw_hits, n_hit_diagonals, n_random_diagonals, seed, geom=None, n_edge_tiles=6):
    """Scalable verification: (1) sampled diagonals that carry GPU hits replay to
    exactly those hit rows, (2) uniformly sampled diagonals replay to exactly the GPU's hits on
    them (usually none), (3) with `geom` (HavacHwClient.tileGeometry()), the diagonals on the edges of
    tiles, threads and register halves do too. Returns the number of diagonals checked."""
    raw = np.asarray(raw_hits, dtype=np.uint64)
    rows = np.asarray(table).reshape(-1, 4).shape[0]
    cell, seg, row = raw & np.uint64(0x3fff), (raw >> np.uint64(14)) & np.uint64(0x3ffffff), raw >> np.uint64(40)
    pos = (seg * np.uint64(12288) + cell).astype(np.int64)
    diag = pos - row.astype(np.int64)
    by_diag = {}
    for d, r in zip(diag.tolist(), row.tolist()):
        by_diag.setdefault(d, set()).add(int(r))
    g = np.random.Generator(np.random.PCG64(seed))
    buf = np.zeros(1 << 16, dtype=np.uint32)
    checked = 0
    if by_diag:
        keys = np.array(sorted(by_diag))
        for d in g.choice(keys, size=min(n_hit_diagonals, len(keys)), replace=False).tolist():
            got = replay_diagonal(packed, table, d, buf)
            assert got == by_diag[d], f"diagonal {d}: GPU rows {sorted(by_diag[d])[:8]} vs replay {sorted(got)[:8]}"
            checked += 1
    for d in g.integers(-(rows - 1), len(packed) * 4, size=n_random_diagonals).tolist():
        got = replay_diagonal(packed, table, d, buf)
        assert got == by_diag.get(d, set()), f"sampled diagonal {d}"
        checked += 1
    if geom is not None:
        for d in edge_diagonals(geom, rows, len(packed) * 4, n_edge_tiles, seed + 17):
            got = replay_diagonal(packed, table, d, buf)
            assert got == by_diag.get(d, set()), f"edge diagonal {d} (tile offset {(d - geom['loadBegin'] - geom['diagBegin']) % geom['tileDiagonals']})"
            checked += 1
    return checked


# ---- per-cell triage (the reference's byCellComparator) ---------------------------------------

def trace_cells(symbols, table, row_begin, row_end, col_begin, col_end):
    """The CPU oracle's per-cell records of a window (oracle_trace_cells; fields of byCellComparator.hpp:23-31)."""
    from havac_b200.client import CELL_RECORD
    L = oracle()
    symbols = np.ascontiguousarray(symbols, dtype=np.uint8)
    flat = np.ascontiguousarray(table, dtype=np.int8).reshape(-1)
    out = np.zeros((row_end - row_begin, col_end - col_begin), dtype=CELL_RECORD)
    n = L.oracle_trace_cells(symbols, len(symbols), flat, len(flat) // 4, row_begin, row_end, col_begin, col_end, out.ctypes.data)
    if n != out.size:
        raise RuntimeError(f"oracle_trace_cells failed: {n}")
    return out


def first_differing_cell(device_records, oracle_records, row_begin=0, col_begin=0):
    """Walks the two record windows the way compareCellMaps does (test/byCellComparator/byCellComparator.cpp:47-96:
    sequence index outer, model index inner) and returns None when every cell agrees, else
    (phmmIndex, globalSequenceIndex, message naming the fields that differ)."""
    fields = ("prevValue", "matchScore", "cellValue", "phmmVector", "symbol", "passesThreshold")
    diff = np.zeros(device_records.shape, dtype=bool)
    for f in fields:
        d = device_records[f] != oracle_records[f]
        diff |= d.any(axis=-1) if d.ndim == 3 else d
    if not diff.any():
        return None
    cols = np.flatnonzero(diff.any(axis=0))
    j = int(cols[0])
    p = int(np.flatnonzero(diff[:, j])[0])
    hw, sw = device_records[p, j], oracle_records[p, j]
    what = [f for f in fields if np.any(hw[f] != sw[f])]
    show = lambda r: ", ".join(f"{f}={r[f].tolist() if f == 'phmmVector' else int(r[f])}" for f in fields)
    msg = (f"TEST FAIL: cell (phmmIndex {row_begin + p}, sequenceIndex {col_begin + j}) differs in {what}\n"
           f"\tHW: {show(hw)}\n\tSW: {show(sw)}")
    return row_begin + p, col_begin + j, msg
