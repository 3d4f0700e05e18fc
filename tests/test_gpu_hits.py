"""Device-side hit post-processing (havac_b200/csrc/hit_postprocess.cu): the radix
sort into the FPGA's emission order and the resolution of raw hit words to HavacHit numbers, against
the host restatements of the same steps (numpy key sort; the host resolver, which the CPU suite
pins to the reference's fastaVectorGetLocalSequencePositionFromGlobal and to the golden vectors)."""
import os

import numpy as np
import pytest

from havac_b200 import havac, sharding, synth
from havac_b200.client import HavacHwClient, STATE_COMPLETED

pytestmark = pytest.mark.gpu


def run(packed, table, spec=""):
    c = HavacHwClient(spec)
    try:
        c.writeSequence(packed)
        c.writePhmm(table)
        c.invokeHavacSsvAsync()
        assert c.waitForHavacSsvAsync() == STATE_COMPLETED
        return c.getHitList()
    finally:
        c.close()


@pytest.mark.parametrize("n_symbols,rows,hi", [(3 * 12288, 60, (20, 30)),      # a handful of hits
                                               (200 * 12288, 900, (14, 22)),     # thousands, several radix tiles
                                               (400 * 12288, 3000, (30, 40))])   # dense: ~millions of hits
def test_device_sort_is_the_emission_order(n_symbols, rows, hi):
    table, cons = synth.consensus_score_table(5, rows, hi=hi)
    sym = synth.random_symbols(5, n_symbols)
    synth.plant_consensus(5, sym, cons, copies=50, substitution=0.1)
    hits = run(synth.pack_2bit(sym), table)
    assert len(hits) > 3
    key = sharding.emission_order_key(hits)
    assert np.all(key[1:] > key[:-1]), "strictly increasing (segment, row, cell): sorted, no duplicates"


def test_no_hits_and_a_few_hits():
    table = np.full((50, 4), -20, dtype=np.int8)
    sym = synth.random_symbols(6, 12288)
    assert len(run(synth.pack_2bit(sym), table)) == 0
    table[:, :] = -20
    table[np.arange(50), 0] = 6  # only a run of 43+ 'A' reaches 256
    sym[:] = 1
    sym[1000:1043] = 0
    hits = run(synth.pack_2bit(sym), table)
    # the 43-long run is crossed by the diagonals that enter it at rows 0..7: one hit each, at rows 42..49
    assert (hits >> np.uint64(40)).tolist() == list(range(42, 50))


def havac_hits(fa, hmm, host_resolve, monkeypatch, spec=havac.xclbinSrcDefault):
    if host_resolve:
        monkeypatch.setenv("HAVAC_HOST_RESOLVE", "1")
    else:
        monkeypatch.delenv("HAVAC_HOST_RESOLVE", raising=False)
    h = havac.Havac(0, 0.05, spec)
    try:
        havac.srand(3)
        h.loadPhmm(hmm)
        h.loadSequence(fa)
        h.runHardwareClient()
        return h.getHitsFromFinishedRun()
    finally:
        h.close()


def test_device_resolution_equals_host_resolution_with_padding_hits(tmp_path, monkeypatch):
    """Many short records (every boundary quirk of FastaVector.c:361 gets exercised: hits on
    terminators and on the first symbol of the next record) and a model that likes 'A', so the
    zero padding after the last record scores hits that must be dropped."""
    g = synth.rng(71, 1)
    models = synth.random_models(71, [40, 64, 33])
    for m in models:  # poly-A models: the padding (symbol 0) matches them
        m.probs[:, :] = 0.02
        m.probs[:, 0] = 0.94
        m.consensus[:] = 0
    hmm = str(tmp_path / "polya.hmm")
    synth.write_hmm(hmm, models)
    recs = []
    for _ in range(300):
        r = g.integers(0, 4, size=int(g.integers(30, 400)), dtype=np.uint8)
        if g.random() < 0.5:
            r[-int(g.integers(20, 30)):] = 0  # an A-run into the record end
        if g.random() < 0.5:
            r[: int(g.integers(20, 30))] = 0
        recs.append(r)
    fa = str(tmp_path / "many.fa")
    synth.write_fasta(fa, recs)
    dev = havac_hits(fa, hmm, False, monkeypatch)
    host = havac_hits(fa, hmm, True, monkeypatch)
    assert len(host) > 100
    assert dev == host  # same hits in the same (emission) order
    n_records = 300
    assert all(h.sequenceIndex < n_records for h in dev)


def test_device_resolution_on_a_multi_gpu_client(tmp_path, monkeypatch):
    try:
        HavacHwClient("cuda:0,1").close()
    except RuntimeError:
        pytest.skip("fewer than two GPUs visible")
    models = synth.random_models(72, [120, 200, 77])
    hmm = str(tmp_path / "m.hmm")
    synth.write_hmm(hmm, models)
    sym = synth.random_symbols(72, 600_000)
    synth.plant_homologs(72, sym, models, copies=12, substitution=0.1)
    fa = str(tmp_path / "s.fa")
    synth.write_fasta(fa, synth.split_records(sym, 5))
    one = havac_hits(fa, hmm, False, monkeypatch)
    many = havac_hits(fa, hmm, False, monkeypatch, spec="cuda:all")
    assert len(one) >= 12 and one == many
