"""BASELINE.json configs 3-5 at reduced sequence length but FULL table shape (the table axis is
what stresses the kernel: halo = P_total - 1, thousands of row chunks, long diagonals, dense
hits). Inputs go the whole way: synthetic HMMER3 file -> host projection at the config's
p-value -> C-ABI -> raw hits. Brute force is impossible at these sizes, so parity is checked by
diagonal replay on the CPU oracle: no false positives on sampled hit diagonals,
no misses on uniformly sampled diagonals, and segment-range shards == the whole run.
Each case appends one line to gpurun_out/config_sweep.jsonl (throughput of that shape)."""
import json
import os

import numpy as np
import pytest

import oracle_lib as O
from havac_b200 import havac, sharding, synth
from havac_b200.client import HavacHwClient, STATE_COMPLETED

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def lognormal_lengths(seed, count, mean, lo, hi):
    g = synth.rng(seed, 21)
    sigma = 0.6
    x = g.lognormal(np.log(mean) - sigma * sigma / 2, sigma, size=count)
    return np.clip(np.round(x), lo, hi).astype(np.int64)


CASES = {
    # name: (symbols, model lengths, p-value, planted copies per model)
    "c3_3000_models_mean400": (20_000_000, lambda: lognormal_lengths(3, 3000, 400, 50, 3000), 0.02, 1),
    "c4_200_models_2000_4000": (20_000_000, lambda: synth.model_lengths(4, 200, 2000, 4000), 0.02, 2),
    "c5_500_models_p0.02": (50_000_000, lambda: synth.model_lengths(5, 500, 100, 800), 0.02, 3),
    "c5_500_models_p0.05": (50_000_000, lambda: synth.model_lengths(5, 500, 100, 800), 0.05, 3),
    "c5_500_models_p0.1": (50_000_000, lambda: synth.model_lengths(5, 500, 100, 800), 0.1, 3),
}


@pytest.mark.parametrize("name", list(CASES))
def test_config_shape(name, tmp_path):
    n, lengths_fn, p_value, copies = CASES[name]
    lengths = lengths_fn()
    models = synth.random_models(1000 + len(lengths), lengths)
    hmm = str(tmp_path / "models.hmm")
    synth.write_hmm(hmm, models)
    table, lens = havac.project_phmm(hmm, p_value)
    assert lens.tolist() == [int(x) for x in lengths]
    rows = table.shape[0]

    sym = synth.random_symbols(77, n)
    synth.plant_homologs(78, sym, models, copies=copies, substitution=0.15)
    packed = synth.pack_2bit(sym)
    del sym

    c = HavacHwClient(";hitcap=%d" % (64 << 20))
    try:
        c.writeSequence(packed)
        c.writePhmm(table)
        c.invokeHavacSsvAsync()
        assert c.waitForHavacSsvAsync() == STATE_COMPLETED
        ms = c.deviceTimeMs()
        whole = c.getHitList()
        n_seg = len(packed) * 4 // 12288
        parts = []
        for lo, hi in sharding.plan_shards(n_seg, 2):
            c.setSegmentRange(lo, hi)
            c.writeSequence(packed)
            c.invokeHavacSsvAsync()
            assert c.waitForHavacSsvAsync() == STATE_COMPLETED
            parts.append(c.getHitList())
    finally:
        c.close()
    assert len(whole) >= len(models) // 2, "planted homologs must produce hits"
    assert np.array_equal(sharding.merge_hits(parts), whole)
    checked = O.check_hits_by_replay(packed, table, whole, n_hit_diagonals=150, n_random_diagonals=100, seed=9)

    gcups = n * rows / (ms * 1e-3) / 1e9
    line = {"case": name, "symbols": n, "models": len(models), "rows": int(rows), "p_value": p_value,
            "hits": int(len(whole)), "hits_per_gcell": len(whole) / (n * rows / 1e9), "device_ms": ms,
            "gcups": gcups, "diagonals_replayed": checked}
    print(json.dumps(line))
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "config_sweep.jsonl"), "a") as fh:
            fh.write(json.dumps(line) + "\n")


def test_sequence_longer_than_2_32_symbols():
    """6.5 G symbols (1.6 GB packed; the boundary admits just under 4 GiB, host/HavacHwClient.cpp:92-97):
    positions, diagonals and segment numbers beyond 32 bits. Fragments of the consensus are planted on
    both sides of 2^32 and near the end; hits must sit on them, and hit diagonals plus uniformly
    sampled diagonals must replay exactly on the CPU."""
    n = 6_500_000_077
    rows = 257
    table, cons = synth.consensus_score_table(41, rows)
    packed = synth.random_packed(41, n)
    g = synth.rng(41, 3)
    planted = []
    for at in [(1 << 32) - 90, (1 << 32) + 5, (1 << 32) + 12288 * 3 - 20, 5_000_000_001, n - 400, 123_456]:
        a = int(g.integers(0, 60))
        frag = cons[a:a + 150].copy()
        mut = g.random(len(frag)) < 0.06
        frag[mut] = g.integers(0, 4, size=int(mut.sum()))
        for k, s in enumerate(frag.tolist()):
            p = at + k
            packed[p >> 2] = (int(packed[p >> 2]) & ~(3 << (2 * (p & 3)))) | (s << (2 * (p & 3)))
        planted.append((at, len(frag)))
    c = HavacHwClient("")
    try:
        c.writeSequence(packed)
        c.writePhmm(table)
        c.invokeHavacSsvAsync()
        assert c.waitForHavacSsvAsync() == STATE_COMPLETED
        hits = c.getHitList()
        ms = c.deviceTimeMs()
    finally:
        c.close()
    assert np.all(hits[1:] > 0)
    pos = ((hits >> np.uint64(14)) & np.uint64(0x3ffffff)).astype(np.int64) * 12288 + (hits & np.uint64(0x3fff)).astype(np.int64)
    assert pos.max() < len(packed) * 4
    for at, ln in planted:
        assert ((pos >= at) & (pos < at + ln)).any(), f"no hit on the fragment planted at {at}"
    assert (pos > (1 << 32)).sum() >= 3
    checked = O.check_hits_by_replay(packed, table, hits, 300, 2000, seed=41)
    print(f"6.5 G symbols x {rows} rows: {len(hits)} hits, {n * rows / ms / 1e6:.0f} GCUPS, {checked} diagonals replayed")
