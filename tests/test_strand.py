"""The other strand (havac_b200/csrc/strand.cu): reverse complement of the packed
sequence on the host and on the device, the Crick-strand search against the oracle run on the
reversed symbols, and the forward-strand coordinates Havac reports for it."""
import numpy as np
import pytest

import oracle_lib as O
from havac_b200 import havac, synth

SEG = 12288


def reversed_symbols(sym, n):
    out = sym.copy()
    out[:n] = 3 - sym[:n][::-1]
    return out


@pytest.mark.parametrize("n_sym,n_bytes", [(0, 8), (1, 4), (5, 4), (15, 4), (16, 4), (17, 8), (31, 8), (33, 12), (100, 3072),
                                           (12288, 3072), (12287, 3072), (12290, 6144), (40001, 12288)])
def test_host_reverse_complement_vs_symbolwise(n_sym, n_bytes):
    g = synth.rng(91, n_sym)
    packed = g.integers(0, 256, size=n_bytes, dtype=np.uint8)
    got = havac.reverse_complement_packed(packed, n_sym)
    assert np.array_equal(synth.unpack_2bit(got), reversed_symbols(synth.unpack_2bit(packed), n_sym))
    assert np.array_equal(havac.reverse_complement_packed(got, n_sym), packed)  # an involution


def test_host_reverse_complement_rejects_a_count_beyond_the_buffer():
    with pytest.raises(Exception):
        havac.reverse_complement_packed(np.zeros(4, dtype=np.uint8), 17)


# ----------------------------------------------------------------------------------------- GPU

def _client(spec=""):
    from havac_b200.client import HavacHwClient
    return HavacHwClient(spec)


@pytest.mark.gpu
@pytest.mark.parametrize("n_sym", [1, 16, 12287, 12288, 12289, 5 * SEG - 7, 40 * SEG - 16, 40 * SEG])
def test_device_reverse_complement_equals_host(n_sym):
    g = synth.rng(92, n_sym)
    segs = (n_sym + SEG - 1) // SEG
    packed = g.integers(0, 256, size=segs * SEG // 4, dtype=np.uint8)
    c = _client()
    try:
        c.writeSequence(packed)
        assert c.mirroredSymbols() == 0
        c.reverseComplement(n_sym)
        assert c.mirroredSymbols() == n_sym
        dev = c.readSequence()
        assert np.array_equal(dev, havac.reverse_complement_packed(packed, n_sym))
        with pytest.raises(Exception):  # reversed over n_sym: another count cannot undo it
            c.reverseComplement(n_sym - 1 if n_sym > 1 else 2)
        c.reverseComplement(n_sym)
        assert c.mirroredSymbols() == 0 and np.array_equal(c.readSequence(), packed)
        c.reverseComplement(n_sym)
        c.writeSequence(packed)  # a new sequence is a forward one
        assert c.mirroredSymbols() == 0
    finally:
        c.close()


def _crick_case(seed, n, rows):
    """Forward sequence with fragments of the table's consensus planted in BOTH orientations."""
    g = synth.rng(seed, 5)
    table, cons = synth.consensus_score_table(seed, rows)
    sym = synth.random_symbols(seed, n)
    fwd_at, rev_at = [], []
    for k in range(14):
        a = int(g.integers(0, rows - 60))
        frag = cons[a:a + int(g.integers(50, min(rows - a, 160) + 1))].copy()
        mut = g.random(len(frag)) < 0.08
        frag[mut] = g.integers(0, 4, size=int(mut.sum()))
        at = int(g.integers(0, n - len(frag)))
        if k % 2:
            sym[at:at + len(frag)] = 3 - frag[::-1]
            rev_at.append((at, len(frag)))
        else:
            sym[at:at + len(frag)] = frag
            fwd_at.append((at, len(frag)))
    return sym, table, fwd_at, rev_at


@pytest.mark.gpu
def test_crick_search_equals_oracle_on_reversed_symbols():
    from havac_b200.client import STATE_COMPLETED
    n = 300_001
    sym, table, fwd_at, rev_at = _crick_case(93, n, 400)
    packed = synth.pack_2bit(sym)
    padded = synth.unpack_2bit(packed)

    def run(c):
        c.invokeHavacSsvAsync()
        assert c.waitForHavacSsvAsync() == STATE_COMPLETED
        return np.sort(c.getHitList())

    c = _client()
    try:
        c.writeSequence(packed)
        c.writePhmm(table)
        watson = run(c)
        c.reverseComplement(n)
        crick = run(c)
        c.reverseComplement(n)
        again = run(c)
    finally:
        c.close()
    pw, rw = O.soft_ssv(padded, table)
    pc, rc = O.soft_ssv(reversed_symbols(padded, n), table)
    assert np.array_equal(watson, np.sort(O.encode_hits(pw, rw))) and np.array_equal(again, watson)
    assert np.array_equal(crick, np.sort(O.encode_hits(pc, rc)))
    # the planted fragments are found on their own strand
    fw = np.unique(watson & np.uint64((1 << 40) - 1))
    wpos = ((fw >> np.uint64(14)) * np.uint64(SEG) + (fw & np.uint64(0x3fff))).astype(np.int64)
    fc = np.unique(crick & np.uint64((1 << 40) - 1))
    cpos = n - 1 - ((fc >> np.uint64(14)) * np.uint64(SEG) + (fc & np.uint64(0x3fff))).astype(np.int64)
    assert all(((wpos >= at) & (wpos < at + ln)).any() for at, ln in fwd_at)
    assert all(((cpos >= at) & (cpos < at + ln)).any() for at, ln in rev_at)
    assert len(crick) >= len(rev_at)


@pytest.mark.gpu
def test_havac_reports_crick_hits_at_forward_coordinates(tmp_path, monkeypatch):
    models = synth.random_models(94, [180, 240])
    hmm = str(tmp_path / "m.hmm")
    synth.write_hmm(hmm, models)
    sym = synth.random_symbols(94, 400_000)
    rc_source = synth.random_symbols(95, 400_000)
    synth.plant_homologs(94, sym, models, copies=6, substitution=0.08)
    synth.plant_homologs(95, rc_source, models, copies=6, substitution=0.08)
    # the second half of the file carries homologs on the other strand
    sym[200_000:] = 3 - rc_source[:200_000][::-1]
    records = synth.split_records(sym, 3)
    fa = str(tmp_path / "s.fa")
    synth.write_fasta(fa, records)

    def both(resolve_on_host):
        if resolve_on_host:
            monkeypatch.setenv("HAVAC_HOST_RESOLVE", "1")
        else:
            monkeypatch.delenv("HAVAC_HOST_RESOLVE", raising=False)
        h = havac.Havac(0, 0.02)
        try:
            havac.srand(9)
            h.loadPhmm(hmm)
            h.loadSequence(fa)
            forward_packed = h.devicePackedSequence()
            ends = h.sequenceEndPositions()
            h.runHardwareClient()
            watson = list(zip(*[a.tolist() for a in h.getHitArraysFromFinishedRun()]))
            h.useReverseComplement(True)
            h.runHardwareClient()
            crick = list(zip(*[a.tolist() for a in h.getHitArraysFromFinishedRun()]))
            h.useReverseComplement(False)
            assert np.array_equal(h.devicePackedSequence(), forward_packed)
            return watson, crick, forward_packed, ends, h.devicePhmmScores()
        finally:
            h.close()

    watson, crick, packed, ends, table = both(False)
    watson_h, crick_h, _, _, _ = both(True)
    assert watson == watson_h and crick == crick_h and len(crick) >= 3 and len(watson) >= 3
    # the oracle on the reversed symbols, mapped to forward coordinates and resolved by the host resolver
    n = int(ends[-1])
    padded = synth.unpack_2bit(packed)
    pc, rc = O.soft_ssv(reversed_symbols(padded, n), table)
    keep = pc < n
    fwd_pos = n - 1 - pc[keep].astype(np.int64)
    lens = np.array([m.length for m in models], dtype=np.uint32)
    want = havac.resolve_hits(O.encode_hits(fwd_pos, rc[keep]), ends, lens)
    want = sorted(zip(*[a.tolist() for a in want]))
    assert sorted(crick) == want
    # and a crick hit does sit in the reversed half of the file
    assert any(int(ends[h[1] - 1] if h[1] else 0) + h[0] >= 200_000 for h in crick)


@pytest.mark.gpu
def test_segment_range_clients_refuse_and_multi_gpu_clients_agree():
    from havac_b200.client import STATE_COMPLETED
    n = 20 * SEG - 5
    sym, table, _, _ = _crick_case(96, n, 300)
    packed = synth.pack_2bit(sym)
    c = _client()
    try:
        c.setSegmentRange(0, 10)
        c.writeSequence(packed)
        with pytest.raises(Exception):
            c.reverseComplement(n)
    finally:
        c.close()

    def crick(spec):
        c = _client(spec)
        try:
            c.writeSequence(packed)
            c.writePhmm(table)
            c.reverseComplement(n)
            c.invokeHavacSsvAsync()
            assert c.waitForHavacSsvAsync() == STATE_COMPLETED
            hits = np.sort(c.getHitList())
            c.reverseComplement(n)
            c.invokeHavacSsvAsync()
            assert c.waitForHavacSsvAsync() == STATE_COMPLETED
            return hits, np.sort(c.getHitList())
        finally:
            c.close()

    one_c, one_w = crick("")
    try:
        many_c, many_w = crick("cuda:all")
    except RuntimeError:
        pytest.skip("cuda:all unavailable")
    assert len(one_c) >= 3
    assert np.array_equal(one_w, many_w)
    assert np.array_equal(one_c, many_c)
