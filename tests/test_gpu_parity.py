"""GPU parity: the CUDA path, called through the C-ABI (include/havac_cuda.h), against the
CPU oracle (oracle/soft_ssv.c, a restatement of test/softSsv/SoftSsv.cpp:15-67) on the same
packed sequence buffer and the same int8 table. Integer work: the bar is bit-exact equality
of the raw u64 hit sets (the reference's own comparators check set membership only,
device/test/softwareTestbench.cpp:224-306)."""
import numpy as np
import pytest

import oracle_lib as O
from havac_b200 import synth
from havac_b200.client import HavacHwClient, STATE_COMPLETED

pytestmark = pytest.mark.gpu


def consensus_table(seed, n_rows, hi=(14, 22), lo=(-32, -9)):
    g = synth.rng(seed, 77)
    t = g.integers(lo[0], lo[1] + 1, size=(n_rows, 4)).astype(np.int8)
    cons = g.integers(0, 4, size=n_rows)
    t[np.arange(n_rows), cons] = g.integers(hi[0], hi[1] + 1, size=n_rows)
    return t, cons.astype(np.uint8)


def planted_sequence(seed, n, cons, copies, sub=0.1):
    g = synth.rng(seed, 78)
    sym = synth.random_symbols(seed, n)
    L = len(cons)
    for _ in range(copies):
        a = int(g.integers(0, L - 40)) if L > 80 else 0
        b = int(g.integers(a + 40, L + 1)) if L > 80 else L
        frag = cons[a:b].copy()
        mut = g.random(len(frag)) < sub
        frag[mut] = g.integers(0, 4, size=int(mut.sum()))
        if len(frag) >= n:
            continue
        at = int(g.integers(0, n - len(frag)))
        sym[at:at + len(frag)] = frag
    return sym


def gpu_hits(packed, table, kernel="auto", seg_range=None, spec=""):
    c = HavacHwClient(spec)
    try:
        c.setKernel(kernel)
        if seg_range:
            c.setSegmentRange(*seg_range)
        c.writeSequence(packed)
        c.writePhmm(table)
        c.invokeHavacSsvAsync()
        assert c.waitForHavacSsvAsync() == STATE_COMPLETED
        return c.getHitList()
    finally:
        c.close()


def oracle_hits(packed, table):
    sym = synth.unpack_2bit(packed)
    pos, row = O.soft_ssv(sym, table)
    return np.sort(O.encode_hits(pos, row))


def check(packed, table, kernels=("f16x2", "f16x2-streaming", "scalar"), min_hits=1):
    want = oracle_hits(packed, table)
    assert len(want) >= min_hits, "test input produced too few oracle hits to be meaningful"
    for k in kernels:
        got = np.sort(gpu_hits(packed, table, k))
        assert len(got) == len(want), f"{k}: {len(got)} hits vs oracle {len(want)}"
        assert np.array_equal(got, want), f"{k}: hit sets differ"
    return want


def test_small_two_models():
    t, cons = consensus_table(1, 257)
    sym = planted_sequence(1, 100_000, cons, 12)
    check(synth.pack_2bit(sym), t, min_hits=3)


def test_rows_exceed_columns_left_edge():
    """P_total > N: almost every diagonal starts on the left edge (column 0)."""
    t, cons = consensus_table(2, 20_000)
    sym = planted_sequence(2, 12_288, cons, 30)
    check(synth.pack_2bit(sym), t, min_hits=3)


def test_hit_dense_resets():
    """Positive-drift table: every diagonal hits over and over, exercising reset-after-hit."""
    g = synth.rng(3, 1)
    t = g.integers(-20, 60, size=(300, 4)).astype(np.int8)
    sym = synth.random_symbols(3, 30_000)
    want = check(synth.pack_2bit(sym), t, min_hits=10_000)
    assert len(want) > 10_000


def test_hit_dense_long_table_stage_overflow():
    """Positive drift over a table of several row chunks: every warp replays block after block, its hit stage
    (240 words in shared memory) overflows into the direct path and is flushed at the half-full mark many times,
    while the chunks are handed from warp to warp."""
    g = synth.rng(31, 1)
    t = g.integers(-20, 60, size=(4200, 4)).astype(np.int8)  # 3 row chunks
    sym = synth.random_symbols(31, 12288)
    want = check(synth.pack_2bit(sym), t, kernels=("f16x2", "scalar"), min_hits=1_000_000)
    assert len(want) > 1_000_000


def test_uneven_replay_load_lets_warps_drift_chunks_apart():
    """Only 'A' scores: stretches of poly-A make the warps that own their diagonals replay every block while the
    other warps of the CTA never do, so the busy warps fall whole chunks behind (the streaming kernels have no
    CTA-wide barrier between chunks; the last warp through a chunk refills it). 9000 rows = 5 chunks."""
    g = synth.rng(32, 1)
    rows = 9000
    t = np.full((rows, 4), -30, dtype=np.int8)
    t[:, 0] = g.integers(30, 50, size=rows)
    sym = g.integers(1, 4, size=5 * 12288).astype(np.uint8)  # no 'A' at all ...
    for at in range(3000, len(sym) - 2500, 20000):           # ... but for three stretches of 600
        sym[at:at + 600] = 0
    want = check(synth.pack_2bit(sym), t, kernels=("f16x2", "scalar"), min_hits=100_000)
    assert len(want) > 100_000


def test_extreme_scores():
    g = synth.rng(4, 1)
    t = g.choice(np.array([-128, -127, -1, 0, 1, 126, 127], dtype=np.int8), size=(500, 4))
    sym = synth.random_symbols(4, 40_000)
    check(synth.pack_2bit(sym), t, min_hits=100)


def test_exact_255_is_not_a_hit():
    """Score reaches exactly 255 (no hit) and exactly 256 (hit)."""
    t = np.full((600, 4), -128, dtype=np.int8)
    t[:, 0] = 51       # 5 x 51 = 255
    t[5::7, 0] = -128  # two rows of -128 bring any score back to 0 ...
    t[6::7, 0] = -128
    t2 = t.copy()
    t2[4::7, 0] = 52   # ... or 4 x 51 + 52 = 256 on the fifth
    sym = np.zeros(12_288, dtype=np.uint8)
    packed = synth.pack_2bit(sym)
    assert len(oracle_hits(packed, t)) == 0
    for k in ("f16x2", "scalar"):
        assert len(gpu_hits(packed, t, k)) == 0
    check(packed, t2, min_hits=1000)


def test_config1_1mbp_x_200():
    """BASELINE.json configs[0]: 1 Mbp x one 200-row model, homologs planted."""
    t, cons = consensus_table(5, 200)
    sym = planted_sequence(5, 1_000_000, cons, 40, sub=0.12)
    check(synth.pack_2bit(sym), t, kernels=("f16x2",), min_hits=5)


def test_long_model_crosses_many_chunks():
    t, cons = consensus_table(6, 5_000)
    sym = planted_sequence(6, 150_000, cons, 25)
    check(synth.pack_2bit(sym), t, kernels=("f16x2",), min_hits=5)


def test_segment_ranges_union_equals_whole():
    """Sharding contract: hits of [a,b) segment ranges, each computed with its own
    left halo, partition the hits of the whole run exactly."""
    t, cons = consensus_table(7, 3_000)
    sym = planted_sequence(7, 20 * 12_288, cons, 60)
    packed = synth.pack_2bit(sym)
    whole = np.sort(gpu_hits(packed, t))
    assert len(whole) >= 5
    parts = [gpu_hits(packed, t, seg_range=r) for r in ((0, 7), (7, 8), (8, 20))]
    assert np.array_equal(np.sort(np.concatenate(parts)), whole)
    assert np.array_equal(whole, oracle_hits(packed, t))


def test_hit_order_matches_fpga_emission_order():
    t, cons = consensus_table(8, 400)
    sym = planted_sequence(8, 60_000, cons, 30)
    hits = gpu_hits(synth.pack_2bit(sym), t)
    cell, seg, row = hits & np.uint64(0x3fff), (hits >> np.uint64(14)) & np.uint64(0x3ffffff), hits >> np.uint64(40)
    key = (seg << np.uint64(38)) | (row << np.uint64(14)) | cell
    assert len(hits) > 1 and np.all(np.diff(key.astype(np.int64)) > 0)


def test_all_gpus_of_the_box_in_one_process():
    """`cuda:all` shards the sequence over every GPU of the process (halo rows-1 per shard, no
    exchange); the gathered hits must equal the single-GPU run. Skipped on a one-GPU box."""
    try:
        probe = HavacHwClient("cuda:0,1")
    except RuntimeError:
        pytest.skip("fewer than two GPUs visible")
    probe.close()
    table, cons = consensus_table(41, 700)
    sym = planted_sequence(41, 40 * synth.SEGMENT_SYMBOLS - 5, cons, copies=60)
    packed = synth.pack_2bit(sym)
    one = gpu_hits(packed, table)
    many = gpu_hits(packed, table, spec="cuda:all")
    assert len(one) > 20 and np.array_equal(one, many)
    assert np.array_equal(np.sort(one), oracle_hits(packed, table))


def test_every_tail_length():
    """Tables ending 1..32 rows into their last 32-row block (short tables run a build of the fast path
    that computes 8, 16, 24 or 32 rows of it), hit-dense so that hits fall in the last rows."""
    g = synth.rng(9, 1)
    full = g.integers(-20, 60, size=(96, 4)).astype(np.int8)
    packed = synth.pack_2bit(synth.random_symbols(9, 2 * 12_288))
    for rows in range(33, 97):
        t = np.ascontiguousarray(full[:rows])
        t[-1, :] = 60  # whatever a diagonal holds above 195 hits in the last row
        want = oracle_hits(packed, t)
        assert (want >> np.uint64(40)).max() == rows - 1, "no hit in the last row: the case would prove nothing"
        for kernel in ("f16x2", "f16x2-streaming"):  # the short-table kernel (warps on their own) and the streaming kernel's tail build
            got = np.sort(gpu_hits(packed, t, kernel))
            assert np.array_equal(got, want), f"rows={rows} {kernel}: {len(got)} hits vs oracle {len(want)}"


def test_cell_update_exhaustive():
    """All 65,536 (previous score, emission) pairs through the fast kernel's own cell update -- one
    fma.rn.relu.f16x2 on 256 x score, an overflow to +inf standing for "hit" -- in the low and in the high fp16
    half, against the FPGA cell of device/HavacHls.cpp:370-402 (oracle_cell_processor): same next score, same
    hit flag, and the kernel's overflow test fires exactly where a half hit. This is what `dtype` in the bench
    line means by "u8 scores held as f16x2 (exact)"."""
    import ctypes as C
    L = O.oracle()
    want_score = np.zeros(65536, dtype=np.uint8)
    want_hit = np.zeros(65536, dtype=np.uint8)
    s, h = C.c_uint8(0), C.c_int(0)
    for prev in range(256):
        for e in range(-128, 128):
            L.oracle_cell_processor(prev, e, C.byref(s), C.byref(h))
            want_score[prev * 256 + e + 128] = s.value
            want_hit[prev * 256 + e + 128] = h.value
    # the recurrence of test/softSsv/SoftSsv.cpp:39-49 says the same thing
    pe = np.arange(256)[:, None] + np.arange(-128, 128)[None, :]
    assert np.array_equal(want_hit.reshape(256, 256), (pe >= 256).astype(np.uint8))
    assert np.array_equal(want_score.reshape(256, 256), np.where((pe < 0) | (pe >= 256), 0, pe).astype(np.uint8))
    c = HavacHwClient("")
    try:
        score, hit = c.cellSelftest()
    finally:
        c.close()
    assert not (hit & 0x80).any(), "has_inf() disagrees with the halves"
    for half in (0, 1):
        assert np.array_equal(hit[half] & 1, want_hit), f"hit flags differ in fp16 half {half}"
        assert np.array_equal(score[half], want_score), f"scores differ in fp16 half {half}"


def test_short_table_kernel_edges():
    """Tables of at most 512 rows take the short-table kernel: 1024 diagonals per WARP, a private selector ring per warp,
    no barrier between tiles. Row counts around its limits (1, 31, 32, 33, 511, 512; 513 is the streaming kernel's), a
    sequence of exactly one segment and of many, hit-dense so that every warp tile replays, against the oracle; and the
    planted-hit technique of the full-size test at the edges of WARP tiles (diagonal 0, 1, 31, 32, 1022, 1023 of a tile)."""
    g = synth.rng(12, 1)
    full = g.integers(-20, 60, size=(520, 4)).astype(np.int8)
    for n_seg, row_counts in ((1, (1, 31, 32, 33, 512)), (5, (2, 64, 100, 511, 512, 513))):
        packed = synth.pack_2bit(synth.random_symbols(12 + n_seg, n_seg * 12_288 - 3))
        for rows in row_counts:
            t = np.ascontiguousarray(full[:rows])
            want = oracle_hits(packed, t)
            got = np.sort(gpu_hits(packed, t, "f16x2"))
            assert np.array_equal(got, want), f"rows={rows}, {n_seg} segments: {len(got)} hits vs oracle {len(want)}"
    # sparse table, hits planted on warp-tile edges
    table, cons = consensus_table(13, 480)
    n = 40 * 12_288
    packed = synth.pack_2bit(planted_sequence(13, n, cons, 20))
    c = HavacHwClient("")
    try:
        c.writeSequence(packed)
        c.writePhmm(table)
        c.invokeHavacSsvAsync()
        assert c.waitForHavacSsvAsync() == STATE_COMPLETED
        geom = c.tileGeometry()
        assert geom["tileDiagonals"] == 1024
        gg = np.random.Generator(np.random.PCG64(5))
        planted = [(d, r) for d, r in O.edge_targets(geom, 480, n, seed=3, n_tiles=6) if O.plant_hit_at(packed, table, d, r, gg) is not None]
        assert len(planted) >= 12
        c.writeSequence(packed)
        c.invokeHavacSsvAsync()
        assert c.waitForHavacSsvAsync() == STATE_COMPLETED
        got = np.sort(c.getHitList())
    finally:
        c.close()
    want = oracle_hits(packed, table)
    assert np.array_equal(got, want)
    have = set(got.tolist())
    for d, r in planted:
        assert int(O.encode_hits([d + r], [r])[0]) in have
