"""Pins the CPU oracle (oracle/*.c, our restatement) two ways: against the UNMODIFIED reference
objects compiled into oracle/_ref (skipped where that library has not been built), and against
the committed golden vectors those objects produced (tests/golden/, always run)."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_lib as O
from havac_b200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["two_models", "three_records"]
needs_ref = pytest.mark.skipif(O.ref() is None, reason="oracle/_ref/libhavac_ref.so not built on this machine")


def hitset(pos, row):
    return set(zip(np.asarray(pos).tolist(), np.asarray(row).tolist()))


@pytest.mark.parametrize("name", CASES)
def test_soft_ssv_matches_golden(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    pos, row = O.soft_ssv(synth.unpack_2bit(g["packed"]), g["table"])
    assert len(pos) == len(g["hit_pos"]) > 10
    # same hits AND the same emission order (rows ascending, positions descending: SoftSsv.cpp:31-35)
    assert np.array_equal(pos, g["hit_pos"]) and np.array_equal(row, g["hit_row"])


@pytest.mark.parametrize("name", CASES)
def test_resolve_matches_golden(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    L = O.oracle()
    got = []
    for raw in O.encode_hits(g["hit_pos"], g["hit_row"]).tolist():
        sp, si, pp, pi = C.c_uint64(0), C.c_uint32(0), C.c_uint32(0), C.c_uint32(0)
        if L.oracle_resolve_hit(raw, g["seq_end"], len(g["seq_end"]), g["model_len"], len(g["model_len"]),
                                C.byref(sp), C.byref(si), C.byref(pp), C.byref(pi)):
            got.append((sp.value, si.value, pp.value, pi.value))
    assert np.array_equal(np.array(got, dtype=np.uint64).reshape(-1, 4), g["res"])


def test_scale_factor_matches_golden():
    g = np.load(os.path.join(GOLDEN, "scale_factors.npz"))
    L = O.oracle()
    got = np.array([L.oracle_scale_factor(float(m), float(l), int(a), int(b), float(p))
                    for m, l, a, b, p in zip(g["mu"], g["lam"], g["maxl"], g["leng"], g["p"])], dtype=np.float32)
    assert np.array_equal(got.view(np.uint32), g["scale"].view(np.uint32))  # bit-identical floats


def test_cell_processor_equals_recurrence_on_all_inputs():
    """The FPGA cell (9-bit add, carry/sign trick: device/HavacHls.cpp:370-402) and the softSsv
    form (int add, <0 -> 0, >=256 -> hit + 0: SoftSsv.cpp:38-50) agree on all 65536 inputs."""
    L = O.oracle()
    for prev in range(256):
        for match in range(-128, 128):
            s, h = C.c_uint8(0), C.c_int(0)
            L.oracle_cell_processor(prev, match, C.byref(s), C.byref(h))
            v = prev + match
            want = (0, 0) if v < 0 else ((0, 1) if v >= 256 else (v, 0))
            assert (s.value, h.value) == want


def test_pack_unpack_roundtrip_and_layout():
    sym = synth.random_symbols(9, 1001)
    L = O.oracle()
    packed = np.zeros((len(sym) + 3) // 4, dtype=np.uint8)
    L.oracle_pack_2bit(sym, len(sym), packed)
    assert np.array_equal(packed, synth.pack_2bit(sym, pad_to_segment=False))
    back = np.zeros(len(sym), dtype=np.uint8)
    L.oracle_unpack_2bit(packed, len(sym), back)
    assert np.array_equal(back, sym)
    assert packed[0] == (sym[0] | (sym[1] << 2) | (sym[2] << 4) | (sym[3] << 6))  # LSB first


def test_diagonal_replay_agrees_with_full_matrix():
    table, cons = synth.consensus_score_table(21, 300)
    sym = synth.random_symbols(21, 20_000)
    synth.plant_consensus(21, sym, cons, copies=10)
    pos, row = O.soft_ssv(sym, table)
    full = hitset(pos, row)
    assert len(full) > 5
    L = O.oracle()
    flat = np.ascontiguousarray(table).reshape(-1)
    rows = np.zeros(64, dtype=np.uint32)
    seen = set()
    for d in sorted({int(p) - int(r) for p, r in full} | set(range(-299, 20_000, 997))):
        n = L.oracle_replay_diagonal(sym, len(sym), flat, 300, d, rows, len(rows))
        seen |= {(d + int(r), int(r)) for r in rows[:n]}
    assert seen == full


def test_parallel_split_is_exact():
    table, cons = synth.consensus_score_table(22, 700)
    sym = synth.random_symbols(22, 60_000)
    synth.plant_consensus(22, sym, cons, copies=25)
    a = hitset(*O.soft_ssv(sym, table))
    b = hitset(*O.soft_ssv_parallel(sym, table, 5))
    assert a == b and len(a) > 5


# ---- against the compiled reference ----------------------------------------------------------

@needs_ref
def test_soft_ssv_vs_reference_random_tables():
    for seed in range(4):
        g = synth.rng(100 + seed, 1)
        rows = int(g.integers(1, 400))
        n = int(g.integers(1, 30_000))
        table = g.integers(-128, 128, size=(rows, 4)).astype(np.int8) if seed % 2 else synth.consensus_score_table(seed, rows)[0]
        sym = synth.random_symbols(100 + seed, n)
        p1, r1 = O.soft_ssv(sym, table)
        p2, r2 = O.soft_ssv(sym, table, lib=O.ref())
        assert np.array_equal(p1, p2) and np.array_equal(r1, r2)


@needs_ref
def test_scale_factor_and_projection_vs_reference(tmp_path):
    Lr, Lo = O.ref(), O.oracle()
    g = synth.rng(55, 1)
    for _ in range(200):
        mu, lam = np.float32(g.uniform(-12, -5)), np.float32(g.uniform(0.6, 0.8))
        leng = int(g.integers(20, 5000)); maxl = leng + int(g.integers(1, 3000))
        p = np.float32(g.choice([1e-12, 1e-7, 0.02, 0.1, 0.7]))
        a = np.float32(Lo.oracle_scale_factor(mu, lam, maxl, leng, p))
        b = np.float32(Lr.ref_scale_factor(mu, lam, maxl, leng, p))
        assert a.view(np.uint32) == b.view(np.uint32)
    models = synth.random_models(56, [41, 300])
    hmm = str(tmp_path / "m.hmm")
    synth.write_hmm(hmm, models)
    for p in (0.02, 0.1):
        want, lens = O.ref_load_phmm(hmm, p)
        got = []
        for m in models:
            s = Lo.oracle_scale_factor(np.float32(m.mu), np.float32(m.lam), m.max_length, m.length, np.float32(p))
            # the file stores mu with 4 and lambda with 5 decimals: project what the parser reads back
            s = Lo.oracle_scale_factor(np.float32(f"{m.mu:.4f}"), np.float32(f"{m.lam:.5f}"), m.max_length, m.length, np.float32(p))
            out = np.zeros(4 * m.length, dtype=np.int8)
            Lo.oracle_project_model(m.neg_log_emissions().astype(np.float32).reshape(-1), m.length, s, out)
            got.append(out)
        assert np.array_equal(np.concatenate(got), want)
