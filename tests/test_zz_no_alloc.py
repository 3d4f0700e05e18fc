"""havac_cuda_sort_hits_static (include/havac_cuda.h): the emission-order sort with its scratch in a file-scope
device array of the library -- no cudaMalloc / cudaFree inside the call. It has no counterpart in the reference (added
for callers that cannot allocate), so its check is against the EXISTING entry
point: havac_cuda_read_hits returns the hits of a run sorted by the cudaMalloc'ing sort; the same hit words, shuffled,
must come back from the static entry point in exactly that order. Integer work: the bar is bit-exact.

Collected last on purpose (the file name): a failure here must not hide the
parity tests. Passes alone in a fresh process and after the other tests (it opens and closes its own client)."""
import ctypes as C

import numpy as np
import pytest

from havac_b200 import client, sharding, synth
from havac_b200.client import HavacHwClient, STATE_COMPLETED

pytestmark = pytest.mark.gpu

MAX_HITS = 1 << 20  # HAVAC_CUDA_SORT_STATIC_MAX_HITS


def sorted_hits_of_a_run(seed, n_segments, rows, hi):
    """Hit list of one search as the existing entry point (havac_cuda_read_hits) returns it: emission order."""
    table, cons = synth.consensus_score_table(seed, rows, hi=hi)
    sym = synth.random_symbols(seed, n_segments * synth.SEGMENT_SYMBOLS)
    synth.plant_consensus(seed, sym, cons, copies=50, substitution=0.1)
    c = HavacHwClient("")
    try:
        c.writeSequence(synth.pack_2bit(sym))
        c.writePhmm(table)
        c.invokeHavacSsvAsync()
        assert c.waitForHavacSsvAsync() == STATE_COMPLETED
        return np.ascontiguousarray(c.getHitList(), dtype=np.uint64)
    finally:
        c.close()


def static_sort(L, hits_in, out=None, device=0):
    a = np.ascontiguousarray(hits_in, dtype=np.uint64)
    if out is None:
        out = np.full(len(a), 0xDEADBEEFDEADBEEF, dtype=np.uint64)
    rc = L.havac_cuda_sort_hits_static(device, a.ctypes.data, len(a), out.ctypes.data, None)
    assert rc == client.OK, (rc, (L.havac_cuda_last_error(None) or b"").decode())
    return out


def test_static_sort_equals_the_allocating_sort_on_two_calls_in_a_row():
    L = client.load_library()
    want = sorted_hits_of_a_run(5, 20, 3000, (30, 40))  # 459,694 hits: 113 radix tiles
    assert 2 * 4096 < len(want) <= MAX_HITS, f"{len(want)} hits: wanted several radix tiles, within the static maximum"
    g = np.random.default_rng(20261105)
    # call 1: all hits of the run, shuffled
    got1 = static_sort(L, want[g.permutation(len(want))])
    assert np.array_equal(got1, want), "first call differs from havac_cuda_read_hits' order"
    # call 2, straight after, on the SAME scratch: a different, shorter input (every third hit; fewer tiles, so most of
    # the scratch still holds call 1's keys and counts). A subsequence of a sorted list of distinct words is sorted.
    want2 = np.ascontiguousarray(want[::3])
    got2 = static_sort(L, want2[g.permutation(len(want2))])
    assert np.array_equal(got2, want2), "second call differs: the scratch carried something over"
    # and the long one again after the short one, in place (hitsOut == hitsIn)
    buf = want[g.permutation(len(want))].copy()
    static_sort(L, buf, out=buf)
    assert np.array_equal(buf, want)


def test_static_sort_on_device_memory_of_the_caller():
    """Device pointers in and out (the caller's own allocations: torch's here), on the caller's stream."""
    torch = pytest.importorskip("torch")
    L = client.load_library()
    want = sorted_hits_of_a_run(5, 40, 600, (24, 34))  # 23,089 hits
    assert 4096 < len(want) <= MAX_HITS
    g = np.random.default_rng(7)
    for sub in (want, np.ascontiguousarray(want[1::2])):  # two calls in a row here too
        shuffled = sub[g.permutation(len(sub))]
        d_in = torch.from_numpy(shuffled.view(np.int64)).to("cuda:0")
        d_out = torch.zeros_like(d_in)
        stream = torch.cuda.current_stream(0).cuda_stream
        rc = L.havac_cuda_sort_hits_static(0, d_in.data_ptr(), len(sub), d_out.data_ptr(), C.c_void_p(stream))
        assert rc == client.OK, (L.havac_cuda_last_error(None) or b"").decode()
        assert np.array_equal(d_out.cpu().numpy().view(np.uint64), sub)
        assert np.array_equal(d_in.cpu().numpy().view(np.uint64), shuffled), "the input is left alone"


def test_static_sort_sizes_at_and_beyond_the_limit():
    L = client.load_library()
    # nothing to do / one hit
    assert L.havac_cuda_sort_hits_static(0, None, 0, None, None) == client.OK
    one = np.array([(7 << 40) | (3 << 14) | 5], dtype=np.uint64)
    assert np.array_equal(static_sort(L, one), one)
    # beyond the maximum: refused before anything is touched, with a text that names the limit
    a = np.zeros(8, dtype=np.uint64)
    rc = L.havac_cuda_sort_hits_static(0, a.ctypes.data, MAX_HITS + 1, a.ctypes.data, None)
    assert rc == client.E_LENGTH
    msg = (L.havac_cuda_last_error(None) or b"").decode()
    assert str(MAX_HITS) in msg and "maximum" in msg, msg
    # exactly the maximum: distinct synthetic hit words over every field (cell 14 bits, segment 26, row 24)
    g = np.random.default_rng(3)
    idx = g.permutation(1 << 22)[:MAX_HITS].astype(np.uint64)  # distinct keys: seg = idx >> 12, row = idx & 4095, cell varies with both
    seg, row = idx >> np.uint64(12), idx & np.uint64(4095)
    cell = (idx * np.uint64(2654435761)) % np.uint64(12288)
    hits = (row * np.uint64(4093)) << np.uint64(40) | seg << np.uint64(14) | cell  # rows spread over all 24 bits
    key = sharding.emission_order_key(hits)
    assert len(np.unique(key)) == MAX_HITS
    got = static_sort(L, hits)
    assert np.array_equal(got, hits[np.argsort(key, kind="stable")])
