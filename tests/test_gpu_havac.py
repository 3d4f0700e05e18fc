"""End to end through the public driver API on a GPU (`Havac`, mirrored from host/Havac.hpp):
FASTA + .hmm files in, HavacHit 4-tuples out, against the golden vectors the reference objects
produced and against the CPU oracle. Also the client-call semantics the reference defines:
call-order errors, size contracts, states, abort, hit-buffer overflow."""
import ctypes
import os
import time

import numpy as np
import pytest

import oracle_lib as O
from havac_b200 import havac, sharding, synth
from havac_b200.client import (STATE_ABORT, STATE_COMPLETED, HavacHwClient, HitOverflowError, LogicError)

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def tuples(arrs):
    sp, si, pp, pi = arrs
    return sorted(zip(sp.tolist(), si.tolist(), pp.tolist(), pi.tolist()))


@pytest.mark.parametrize("name", ["two_models", "three_records"])
def test_havac_files_to_hits_match_golden(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    h = havac.Havac(0, float(g["p_value"]))
    try:
        # the record terminators become rand()-chosen bases (SequencePreprocessor.cpp:62-85): seed as the
        # golden run did, through the same libc
        ctypes.CDLL(None).srand(int(g["seed"]))
        h.loadSequence(os.path.join(GOLDEN, name + ".fa"))
        h.loadPhmm(os.path.join(GOLDEN, name + ".hmm"))
        h.runHardwareClient()
        assert h.currentHardwareState() == havac.HAVAC_CMD_STATE_COMPLETED
        got = tuples(h.getHitArraysFromFinishedRun())
        hits = h.getHitsFromFinishedRun()
    finally:
        h.close()
    want = sorted(map(tuple, g["res"].tolist()))
    assert len(want) > 10 and got == want
    assert hits[0].toString().startswith("sequence $") and "; phmm #" in hits[0].toString()


def test_havac_multi_record_multi_model_vs_oracle(tmp_path):
    lens = [150, 90, 400, 33]
    models = synth.random_models(61, lens)
    sym = synth.random_symbols(61, 300_000)
    synth.plant_homologs(61, sym, models, copies=8, substitution=0.1)
    fa, hmm = str(tmp_path / "s.fa"), str(tmp_path / "m.hmm")
    synth.write_fasta(fa, synth.split_records(sym, 5))
    synth.write_hmm(hmm, models)
    for p in (0.02, 0.1):
        h = havac.Havac(0, p)
        try:
            ctypes.CDLL(None).srand(5)
            h.loadSequence(fa)
            h.loadPhmm(hmm)
            h.runHardwareClientAsync()
            h.waitHardwareClientAsync()
            got = tuples(h.getHitArraysFromFinishedRun())
        finally:
            h.close()
        # the oracle side: the same files through the device-free contract doors with the same
        # srand() seed give the same packed bytes (terminators are rand()-resolved), then softSsv
        packed, ends = havac.pack_fasta(fa, seed=5)
        table, mlens = havac.project_phmm(hmm, p)
        pos, row = O.soft_ssv(synth.unpack_2bit(packed), table)
        want = tuples(havac.resolve_hits(O.encode_hits(pos, row), ends, mlens))
        assert len(want) > 10 and got == want


def test_call_order_and_size_contracts():
    h = havac.Havac(0, 0.02)
    try:
        with pytest.raises(LogicError, match="Phmm was not loaded"):
            h.runHardwareClient()
        with pytest.raises(LogicError):
            h.currentHardwareState()  # HavacHwClient.cpp:163-170: no run object yet
    finally:
        h.close()
    c = HavacHwClient("")
    try:
        with pytest.raises(ValueError, match="multiple of the sequence segment length"):
            c.writeSequence(np.zeros(100, dtype=np.uint8))
        with pytest.raises(ValueError, match="divisible by 4"):
            c.writePhmm(np.zeros(7, dtype=np.int8))
        with pytest.raises(ValueError, match="cannot be 0"):
            c.invokeHavacSsvAsync()
        with pytest.raises(LogicError):
            c.getHwState()
    finally:
        c.close()
    with pytest.raises(RuntimeError, match="not present"):
        HavacHwClient("cuda:97")


def test_abort_stops_a_long_run_and_reports_abort():
    table, _ = synth.consensus_score_table(71, 150_000)  # 3e13 cells: about 0.9 s at 32 TCUPS
    packed = synth.random_packed(71, 200_000_000)
    c = HavacHwClient("")
    try:
        c.writeSequence(packed)
        c.writePhmm(table)
        c.invokeHavacSsvAsync()
        time.sleep(0.02)
        t0 = time.perf_counter()
        st = c.abort()
        dt = time.perf_counter() - t0
        assert st == STATE_ABORT and c.getHwState() == STATE_ABORT
        assert dt < 0.3, f"abort took {dt:.3f} s; the full run is ~0.9 s"
        # and the client is usable again afterwards
        c.setKernel("f16x2")
        small = synth.random_packed(72, 12288 * 4)
        c.writeSequence(small)
        c.invokeHavacSsvAsync()
        assert c.waitForHavacSsvAsync() == STATE_COMPLETED
    finally:
        c.close()


def test_hit_buffer_overflow_is_reported_not_silent():
    g = synth.rng(73, 1)
    table = g.integers(-10, 70, size=(200, 4)).astype(np.int8)  # positive drift: hits everywhere
    packed = synth.random_packed(73, 12288 * 8)
    c = HavacHwClient(";hitcap=1000")
    try:
        c.writeSequence(packed)
        c.writePhmm(table)
        c.invokeHavacSsvAsync()
        c.waitForHavacSsvAsync()
        with pytest.raises(HitOverflowError):
            c.getHitList()
    finally:
        c.close()


def test_hit_buffer_grows_and_the_search_repeats():
    """Without a fixed capacity (";hitcap=N+" sets only the first size) an overflowed run enlarges the
    buffers to the count the kernel reported and searches again; callers see one completed run, by
    blocking wait and by polling."""
    import time
    from havac_b200.client import STATE_COMPLETED
    g = synth.rng(74, 1)
    table = g.integers(-10, 70, size=(200, 4)).astype(np.int8)  # positive drift: hits everywhere
    packed = synth.random_packed(74, 12288 * 8)
    pos, row = O.soft_ssv(synth.unpack_2bit(packed), table, cap=1 << 24)
    want = np.sort(O.encode_hits(pos, row))
    assert len(want) > 3000
    for poll in (False, True):
        c = HavacHwClient(";hitcap=1000+")
        try:
            c.writeSequence(packed)
            c.writePhmm(table)
            c.invokeHavacSsvAsync()
            if poll:
                deadline = time.time() + 60
                while c.getHwState() != STATE_COMPLETED:
                    assert time.time() < deadline
                    time.sleep(0.001)
            else:
                assert c.waitForHavacSsvAsync() == STATE_COMPLETED
            assert np.array_equal(np.sort(c.getHitList()), want)
            c.invokeHavacSsvAsync()  # the enlarged buffer serves the next run at once
            assert c.waitForHavacSsvAsync() == STATE_COMPLETED
            assert np.array_equal(np.sort(c.getHitList()), want)
        finally:
            c.close()


def test_two_clients_on_one_gpu_run_side_by_side():
    """Two Havac objects alive at once (the library keeps no process-wide state): different inputs,
    runs in flight at the same time from two threads, each gets its own hits."""
    import threading
    cases = []
    for seed, rows, n in ((81, 300, 2_000_000), (82, 777, 1_500_000)):
        table, cons = synth.consensus_score_table(seed, rows)
        sym = synth.random_symbols(seed, n)
        synth.plant_consensus(seed, sym, cons, copies=20, substitution=0.08)
        packed = synth.pack_2bit(sym)
        pos, row = O.soft_ssv(synth.unpack_2bit(packed), table)
        cases.append((packed, table, np.sort(O.encode_hits(pos, row))))
    clients = [HavacHwClient("") for _ in cases]
    got = [None, None]
    try:
        for c, (packed, table, _) in zip(clients, cases):
            c.writeSequence(packed)
            c.writePhmm(table)

        def work(i):
            for _ in range(3):
                clients[i].invokeHavacSsvAsync()
                assert clients[i].waitForHavacSsvAsync() == STATE_COMPLETED
                got[i] = np.sort(clients[i].getHitList())

        threads = [threading.Thread(target=work, args=(i,)) for i in range(2)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
    finally:
        for c in clients:
            c.close()
    for g, (_, _, want) in zip(got, cases):
        assert len(want) >= 20 and np.array_equal(g, want)


def test_full_size_properties_c2():
    """BASELINE configs[1] size (100 Mbp x ~35 k rows), where brute force on the CPU is impossible:
    (0) hits are PLANTED so that they fall on the fast kernel's structure edges -- diagonals first/last in a
        tile, in a thread's 32 and in a register half, crossed with rows first/last in a 32-row block, in a
        2048-row chunk and the table's last row -- and every one of them must be reported,
    (1) the fast kernel's hit set == the plain kernel's on the WHOLE matrix (exact: the plain kernel is an
        independently written integer kernel that is itself pinned to the oracle at small sizes),
    (2) every sampled hit diagonal, uniformly sampled diagonal and edge diagonal replays on the CPU oracle to
        exactly the GPU's hit rows, (3) segment-range shards reproduce the whole run."""
    import bench
    packed, table, n, _ = bench.build_workload(1)
    packed = packed.copy()
    rows = table.shape[0]
    c = HavacHwClient("")
    try:
        c.writeSequence(packed)
        c.writePhmm(table)
        c.invokeHavacSsvAsync()
        assert c.waitForHavacSsvAsync() == STATE_COMPLETED
        geom = c.tileGeometry()
        # (0) plant hits on the structure edges this launch geometry has, then search again
        g = np.random.Generator(np.random.PCG64(2024))
        planted = []
        for d, r in O.edge_targets(geom, rows, n, seed=5):
            if O.plant_hit_at(packed, table, d, r, g) is not None:
                planted.append((d, r))
        assert len(planted) >= 40, f"only {len(planted)} edge hits could be planted"
        B, TC, W = geom["blockRows"], geom["rowChunk"], geom["tileDiagonals"]
        origin = geom["loadBegin"] + geom["diagBegin"]
        assert {r % B for _, r in planted} >= {0, B - 1} and {r % TC for _, r in planted} >= {0, TC - 1}
        assert {(d - origin) % W for d, _ in planted} >= {0, 1, 31, 32, W - 2, W - 1}
        c.writeSequence(packed)
        c.invokeHavacSsvAsync()
        assert c.waitForHavacSsvAsync() == STATE_COMPLETED
        whole = c.getHitList()
        fast_ms = c.deviceTimeMs()
        c.setKernel("scalar")
        c.invokeHavacSsvAsync()
        assert c.waitForHavacSsvAsync() == STATE_COMPLETED
        plain = c.getHitList()
        plain_ms = c.deviceTimeMs()
        c.setKernel("f16x2")
        n_seg = len(packed) * 4 // 12288
        parts = []
        for lo, hi in sharding.plan_shards(n_seg, 3):
            c.setSegmentRange(lo, hi)
            c.writeSequence(packed)
            c.invokeHavacSsvAsync()
            assert c.waitForHavacSsvAsync() == STATE_COMPLETED
            parts.append(c.getHitList())
    finally:
        c.close()
    assert len(whole) > 1000
    print(f"C2 full size: {len(whole)} hits, f16x2 {fast_ms:.1f} ms, scalar {plain_ms:.1f} ms, {len(planted)} edge hits planted")
    assert np.array_equal(whole, plain), "fast kernel and plain kernel disagree on the full C2 matrix"
    assert np.array_equal(sharding.merge_hits(parts), whole)
    have = set(whole.tolist())
    for d, r in planted:
        assert int(O.encode_hits([d + r], [r])[0]) in have, f"planted edge hit (diagonal {d}, row {r}) was not reported"
    checked = O.check_hits_by_replay(packed, table, whole, n_hit_diagonals=400, n_random_diagonals=300, seed=74, geom=geom,
                                     n_edge_tiles=8)
    assert checked > 700
