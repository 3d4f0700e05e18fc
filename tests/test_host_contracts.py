"""The host-side contracts either side of the kernel, as the product implements them
(havac_b200/host/*.cpp through include/havac_host.h), against the reference objects in
oracle/_ref where built, and against the committed golden vectors always:
  FASTA -> packed buffer + record ends   FastaVector.c:54-168, SequencePreprocessor.cpp:9-85
  .hmm  -> int8 table                     p7HmmReader.c:45, PhmmReprojection.cpp:36-145, PhmmPreprocessor.cpp:9-31
  raw hit -> HavacHit                     Havac.cpp:104-187, FastaVector.c:345-373
No GPU is needed for any of these."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_lib as O
from havac_b200 import havac, synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["two_models", "three_records"]
needs_ref = pytest.mark.skipif(O.ref() is None, reason="oracle/_ref/libhavac_ref.so not built on this machine")


@pytest.mark.parametrize("name", CASES)
def test_golden_inputs_reproduced_byte_for_byte(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    packed, ends = havac.pack_fasta(os.path.join(GOLDEN, name + ".fa"), seed=int(g["seed"]))
    assert np.array_equal(packed, g["packed"]) and np.array_equal(ends, g["seq_end"])
    assert len(packed) * 4 % 12288 == 0
    table, lens = havac.project_phmm(os.path.join(GOLDEN, name + ".hmm"), float(g["p_value"]))
    assert np.array_equal(table.reshape(-1), g["table"]) and np.array_equal(lens, g["model_len"])


@pytest.mark.parametrize("name", CASES)
def test_golden_hits_resolve_identically(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    raw = O.encode_hits(g["hit_pos"], g["hit_row"])
    sp, si, pp, pi = havac.resolve_hits(raw, g["seq_end"], g["model_len"])
    got = np.stack([sp, si.astype(np.uint64), pp.astype(np.uint64), pi.astype(np.uint64)], axis=1)
    assert np.array_equal(got, g["res"])


def test_scale_factor_golden_bit_identical():
    g = np.load(os.path.join(GOLDEN, "scale_factors.npz"))
    got = np.array([havac.scale_factor(float(m), float(l), int(a), int(b), float(p))
                    for m, l, a, b, p in zip(g["mu"], g["lam"], g["maxl"], g["leng"], g["p"])], dtype=np.float32)
    assert np.array_equal(got.view(np.uint32), g["scale"].view(np.uint32))


def test_resolve_quirks_at_record_boundaries():
    """FastaVector.c:361 tests `globalPos <= sequenceEndPosition`: the terminator symbol of record k
    AND the first symbol of record k+1 both resolve to record k (positions len_k and len_k + 1 counting
    from the record start); further into the padding the hit is dropped (Havac.cpp:169-173)."""
    ends = np.array([100, 250, 300], dtype=np.uint64)  # each counts its terminator
    lens = np.array([10, 20], dtype=np.uint32)
    pos = np.array([0, 99, 100, 101, 249, 250, 251, 300, 301, 5000], dtype=np.uint64)
    row = np.array([0, 9, 10, 29, 5, 5, 5, 5, 5, 5], dtype=np.uint32)
    sp, si, pp, pi = havac.resolve_hits(O.encode_hits(pos, row), ends, lens)
    assert sp.tolist() == [0, 99, 100, 1, 149, 150, 1, 50]
    assert si.tolist() == [0, 0, 0, 1, 1, 1, 2, 2]
    assert pp.tolist() == [0, 9, 0, 19, 5, 5, 5, 5] and pi.tolist() == [0, 0, 1, 1, 0, 0, 0, 0]
    L = O.oracle()  # and the oracle's restatement says the same
    for k, raw in enumerate(O.encode_hits(pos, row).tolist()):
        a, b, c, d = C.c_uint64(0), C.c_uint32(0), C.c_uint32(0), C.c_uint32(0)
        ok = L.oracle_resolve_hit(raw, ends, 3, lens, 2, C.byref(a), C.byref(b), C.byref(c), C.byref(d))
        assert bool(ok) == (k < 8)
        if ok:
            assert (a.value, b.value, c.value, d.value) == (sp[k], si[k], pp[k], pi[k])


def test_symbol_mapping_and_padding(tmp_path):
    fa = tmp_path / "t.fa"
    fa.write_bytes(b">r1 some header\nACGTacgt\n>r2\nTTTT\n")
    packed, ends = havac.pack_fasta(str(fa), seed=3)
    assert ends.tolist() == [9, 14] and len(packed) == 12288 // 4
    sym = synth.unpack_2bit(packed)
    assert sym[:8].tolist() == [0, 1, 2, 3, 0, 1, 2, 3] and sym[9:13].tolist() == [3, 3, 3, 3]
    assert not sym[14:].any()  # zero ('A') padding after the last terminator


def test_malformed_inputs_raise(tmp_path):
    with pytest.raises(RuntimeError):
        havac.pack_fasta(str(tmp_path / "missing.fa"))
    bad = tmp_path / "orphan.fa"
    bad.write_bytes(b"ACGT\n>late\nAC\n")
    with pytest.raises(RuntimeError):
        havac.pack_fasta(str(bad))
    hmm = tmp_path / "bad.hmm"
    hmm.write_text("HMMER3/f\nNAME x\nLENG 3\nALPH DNA\nHMM A C G T\n m->m\n  1 0.1 0.2\n")
    with pytest.raises(RuntimeError, match="not formatted correctly"):
        havac.project_phmm(str(hmm))
    with pytest.raises(RuntimeError):
        havac.project_phmm(str(tmp_path / "missing.hmm"))
    amino = tmp_path / "amino.hmm"
    amino.write_text("not an hmm at all\n")
    with pytest.raises(RuntimeError):
        havac.project_phmm(str(amino))


@needs_ref
def test_fasta_packing_vs_reference_with_odd_input(tmp_path):
    g = synth.rng(31, 1)
    fa = tmp_path / "odd.fa"
    body = bytearray()
    alphabet = b"ACGTacgtNnRYSWKMryswkmBDHVX*-"
    for rec in range(7):
        body += b">rec%d description with spaces;and a comment\n" % rec
        for _ in range(int(g.integers(1, 40))):
            n = int(g.integers(0, 90))
            line = bytes(alphabet[i] for i in g.integers(0, len(alphabet), size=n))
            body += line + (b"\r\n" if g.random() < 0.2 else b"\n")
            if g.random() < 0.1:
                body += b";comment line\n"
    body += b">empty\n>tail\nACGT"  # an empty record, and no newline at end of file
    fa.write_bytes(bytes(body))
    for seed in (1, 99):
        packed, ends = havac.pack_fasta(str(fa), seed=seed)
        rp, re_, cnt = O.ref_load_fasta(str(fa), seed=seed)
        assert np.array_equal(ends, re_) and int(ends[-1]) == cnt
        assert np.array_equal(packed, rp)


@needs_ref
def test_projection_vs_reference_many_models(tmp_path):
    lens = synth.model_lengths(77, 25, 20, 900)
    models = synth.random_models(77, lens)
    hmm = str(tmp_path / "many.hmm")
    synth.write_hmm(hmm, models)
    for p in (1e-10, 0.02, 0.05, 0.1):
        table, got_lens = havac.project_phmm(hmm, p)
        want, want_lens = O.ref_load_phmm(hmm, p)
        assert np.array_equal(got_lens, want_lens) and np.array_equal(table.reshape(-1), want)


@needs_ref
def test_record_lookup_vs_reference():
    g = synth.rng(32, 1)
    ends = np.cumsum(g.integers(1, 500, size=40)).astype(np.uint64)
    pos = g.integers(0, int(ends[-1]) + 50, size=2000).astype(np.uint64)
    sp, si, _, _ = havac.resolve_hits(O.encode_hits(pos, np.zeros(len(pos), dtype=np.uint32)), ends, np.array([5], dtype=np.uint32))
    Lr = O.ref()
    want = []
    for p in pos.tolist():
        a, b = C.c_uint64(0), C.c_uint64(0)
        if Lr.ref_local_position(ends, len(ends), p, C.byref(a), C.byref(b)):
            want.append((b.value, a.value))
    assert list(zip(sp.tolist(), si.tolist())) == want
