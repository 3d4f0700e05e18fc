"""The N>1 launch on CPU: world_size-2 `gloo` processes each take their segment range from
havac_b200.sharding.plan_shards, search it with the left halo the device client would add
(rows-1 symbols; the CPU oracle stands in for the device here), gather the raw hit lists on
rank 0 and merge them. The merged list must equal the single-process search exactly, in
emission order. No collective touches the data path: the gather is the only exchange."""
import os
import sys

import numpy as np
import pytest

from havac_b200 import sharding


def test_plan_shards_covers_everything_once():
    for n_seg in (0, 1, 2, 7, 8139, 244141):
        for world in (1, 2, 3, 4, 8):
            plan = sharding.plan_shards(n_seg, world)
            assert plan[0][0] == 0 and plan[-1][1] == n_seg
            assert all(a[1] == b[0] for a, b in zip(plan, plan[1:]))
            sizes = [hi - lo for lo, hi in plan]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.plan_shards(4, 0)
    assert sharding.halo_symbols(1) == 0 and sharding.halo_symbols(35804) == 35803


def test_merge_hits_orders_like_the_fpga():
    import oracle_lib as O
    pos = np.array([12288 * 2 + 5, 3, 12288 + 7, 3, 12288 * 2 + 4], dtype=np.uint64)
    row = np.array([1, 9, 0, 2, 1], dtype=np.uint32)
    merged = sharding.merge_hits([O.encode_hits(pos[:2], row[:2]), np.zeros(0, np.uint64), O.encode_hits(pos[2:], row[2:])])
    want = O.encode_hits(np.array([3, 3, 12288 + 7, 12288 * 2 + 4, 12288 * 2 + 5]), np.array([2, 9, 0, 1, 1]))
    assert np.array_equal(merged, want)


def _worker(rank, world, port, out_path):
    import torch.distributed as dist
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import oracle_lib as O
    from havac_b200 import synth
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    table, cons = synth.consensus_score_table(41, 1500)
    sym = synth.random_symbols(41, 9 * sharding.SEGMENT_SYMBOLS - 100)
    synth.plant_consensus(41, sym, cons, copies=40)
    packed = synth.pack_2bit(sym)
    allsym = synth.unpack_2bit(packed)
    n_seg = len(allsym) // sharding.SEGMENT_SYMBOLS
    lo, hi = sharding.plan_shards(n_seg, world)[rank]
    a, b = lo * sharding.SEGMENT_SYMBOLS, hi * sharding.SEGMENT_SYMBOLS
    start = max(0, a - sharding.halo_symbols(len(table)))
    pos, row = O.soft_ssv(allsym[start:b], table)
    pos = pos + np.uint64(start)
    keep = pos >= np.uint64(a)
    mine = O.encode_hits(pos[keep], row[keep])
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(mine, gathered, dst=0)
    # the gather bench.py's e2e leg ends with: each rank's list sorted into emission order (the device does that),
    # sent to rank 0 through host memory and laid end to end -- which must already BE the merged order
    mine_sorted = sharding.merge_hits([mine])
    whole_on_0, moved = sharding.gather_hit_lists(mine_sorted, rank, world)
    if rank == 0:
        merged = sharding.merge_hits(gathered)
        assert np.array_equal(whole_on_0, merged) and moved == 8 * (len(merged) - len(mine))
        wp, wr = O.soft_ssv(allsym, table)
        whole = sharding.merge_hits([O.encode_hits(wp, wr)])
        np.savez(out_path, merged=merged, whole=whole, per_rank=np.array([len(g) for g in gathered]))
    else:
        assert whole_on_0 is None and moved == 0
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_gloo_union_equals_whole(tmp_path):
    import torch.multiprocessing as mp
    out = str(tmp_path / "r.npz")
    port = 29600 + os.getpid() % 300
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    r = np.load(out)
    assert len(r["whole"]) > 10 and np.array_equal(r["merged"], r["whole"])
    assert (r["per_rank"] > 0).all(), "both shards should contribute hits for the test to mean something"
