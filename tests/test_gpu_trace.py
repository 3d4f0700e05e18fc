"""Per-cell debug mode: the reference keeps a cell-by-cell comparator for the situation "the hit
sets differ -- where?" (test/byCellComparator/byCellComparator.cpp:47-96, switched by HAVAC_PER_CELL_DATA_TESTING,
device/PublicDefines.h:11, test/softSsv/SoftSsv.cpp:52-58). Here the device writes the same record for every cell of a
window (havac_cuda_trace_cells), the CPU oracle writes its own (oracle_trace_cells), and a helper names the first
cell that differs."""
import numpy as np
import pytest

import oracle_lib as O
from havac_b200 import synth
from havac_b200.client import HavacHwClient, STATE_COMPLETED

pytestmark = pytest.mark.gpu


def setup_case():
    table, cons = synth.consensus_score_table(21, 300)
    sym = synth.random_symbols(21, 2 * synth.SEGMENT_SYMBOLS - 100)
    synth.plant_consensus(21, sym, cons, copies=12, substitution=0.08)
    packed = synth.pack_2bit(sym)
    return packed, synth.unpack_2bit(packed), table


def test_trace_equals_oracle_records():
    packed, sym, table = setup_case()
    c = HavacHwClient("")
    try:
        c.writeSequence(packed)
        c.writePhmm(table)
        for (r0, r1, c0, c1) in ((0, 300, 0, 700), (250, 300, 20_000, 20_513), (0, 1, 0, 1), (17, 90, 24_000, len(sym))):
            dev = c.traceCells(r0, r1, c0, c1)
            cpu = O.trace_cells(sym, table, r0, r1, c0, c1)
            assert O.first_differing_cell(dev, cpu, r0, c0) is None
            assert dev.tobytes() == cpu.tobytes()
        # the records say what the recurrence says: hits of the window == the oracle's hits in it
        dev = c.traceCells(0, 300, 0, len(sym))
        pos, row = O.soft_ssv(sym, table)
        assert len(pos) >= 3
        rr, cc = np.nonzero(dev["passesThreshold"])
        assert sorted(zip(rr.tolist(), cc.tolist())) == sorted(zip(row.tolist(), pos.tolist()))
    finally:
        c.close()


def test_trace_names_the_corrupted_cell():
    """One byte of the table is changed ON THE DEVICE only; the fast kernel's hit set then differs from the
    oracle's, and the triage helper names the first cell (in the comparator's walk order) that uses that byte."""
    packed, sym, table = setup_case()
    want = np.sort(O.encode_hits(*O.soft_ssv(sym, table)))
    row, x = 137, int(np.argmax(table[137]))  # the consensus symbol of row 137: +score -> -100
    c = HavacHwClient("")
    try:
        c.writeSequence(packed)
        c.writePhmm(table)
        c.debugPokePhmm(4 * row + x, -100)
        c.invokeHavacSsvAsync()
        assert c.waitForHavacSsvAsync() == STATE_COMPLETED
        got = np.sort(c.getHitList())
        assert not np.array_equal(got, want), "the corruption should have changed the hit set"
        dev = c.traceCells(0, 300, 0, len(sym))
    finally:
        c.close()
    cpu = O.trace_cells(sym, table, 0, 300, 0, len(sym))
    found = O.first_differing_cell(dev, cpu)
    assert found is not None
    p, j, msg = found
    first_col = int(np.flatnonzero(sym == x)[0])  # the walk is column-major: first column whose symbol reads the byte
    assert (p, j) == (row, first_col), msg
    assert "phmmVector" in msg and "matchScore" in msg
    print(msg)
