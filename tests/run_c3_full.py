"""BASELINE.json configs[2] at FULL size: 3 Gbp synthetic DNA x 3,000 random pHMMs (log-normal
lengths, mean ~400, P_total ~1.19 M rows) at p = 0.02, sharded over the ranks of this launch
(one rank per GPU under torch.distributed.run; a plain `python tests/run_c3_full.py` is the 1-GPU run).

Every rank builds the same inputs from fixed seeds, searches its segment range (left halo
rows-1, no exchange) and reports count / xor / sum of its raw hit words per 1/8th of the segment
axis, so the 1-GPU and 8-GPU hit sets can be compared without moving 2 GB of hits around
(SURVEY.md 8c-3). Rank 0 also replays sampled diagonals on the CPU oracle (test harness code).
One JSON line to stdout and to gpurun_out/c3_full_<N>gpu.json."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))  # this file is test harness: it uses the CPU oracle as the checker
from havac_b200 import havac, sharding, synth  # noqa: E402
from havac_b200.client import STATE_COMPLETED, HavacHwClient  # noqa: E402

N_SYMBOLS = int(float(os.environ.get("C3_SYMBOLS", "3e9")))
N_MODELS = 3000
PARTS = 8


def lognormal_lengths(seed, count, mean, lo, hi):
    g = synth.rng(seed, 21)
    sigma = 0.6
    x = g.lognormal(np.log(mean) - sigma * sigma / 2, sigma, size=count)
    return np.clip(np.round(x), lo, hi).astype(np.int64)


def main():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    t0 = time.perf_counter()
    hmm = "/tmp/c3_models.hmm"
    models = synth.random_models(3000, lognormal_lengths(3, N_MODELS, 400, 50, 3000))
    if rank == 0:
        synth.write_hmm(hmm + ".tmp", models)
        os.replace(hmm + ".tmp", hmm)
    if dist is not None:
        dist.barrier()
    table, lens = havac.project_phmm(hmm, 0.02)
    rows = table.shape[0]
    packed = synth.random_packed(3, N_SYMBOLS)
    g = synth.rng(3, 9)
    for m in models:  # one 15 %-mutated consensus copy per model
        frag = m.consensus.copy()
        mut = g.random(len(frag)) < 0.15
        frag[mut] = g.integers(0, 4, size=int(mut.sum()))
        at = int(g.integers(0, N_SYMBOLS - len(frag) - 8)) & ~3
        nb = (len(frag) + 3) // 4
        window = synth.unpack_2bit(packed[at // 4: at // 4 + nb])
        window[: len(frag)] = frag
        packed[at // 4: at // 4 + nb] = synth.pack_2bit(window, pad_to_segment=False)
    t_inputs = time.perf_counter() - t0

    n_seg = len(packed) * 4 // synth.SEGMENT_SYMBOLS
    seg_lo, seg_hi = sharding.plan_shards(n_seg, world)[rank]
    hitcap = (400 << 20) // world + (8 << 20)
    c = HavacHwClient(f";hitcap={hitcap}", deviceIndex=local)
    try:
        c.setSegmentRange(seg_lo, seg_hi)
        t1 = time.perf_counter()
        c.writeSequence(packed)
        c.writePhmm(table)
        t_write = time.perf_counter() - t1
        if dist is not None:
            dist.barrier()
        t2 = time.perf_counter()
        c.invokeHavacSsvAsync()
        st = c.waitForHavacSsvAsync()
        t_run = time.perf_counter() - t2
        if st != STATE_COMPLETED:
            raise RuntimeError(f"rank {rank}: run ended in state {st}")
        ms = c.deviceTimeMs()
        t3 = time.perf_counter()
        hits = c.getHitList()
        t_hits = time.perf_counter() - t3
    finally:
        c.close()

    # checksums per 1/PARTS of the segment axis
    seg = (hits >> np.uint64(14)) & np.uint64(0x3FFFFFF)
    sums = []
    for lo, hi in sharding.plan_shards(n_seg, PARTS):
        sel = hits[(seg >= np.uint64(lo)) & (seg < np.uint64(hi))]
        sums.append([int(len(sel)), int(np.bitwise_xor.reduce(sel)) if len(sel) else 0,
                     int(np.add.reduce(sel, dtype=np.uint64)) if len(sel) else 0])
    my_symbols = min(N_SYMBOLS, seg_hi * synth.SEGMENT_SYMBOLS) - seg_lo * synth.SEGMENT_SYMBOLS
    mine = {"rank": rank, "segments": [seg_lo, seg_hi], "device_ms": ms, "hits": int(len(hits)), "run_wall_s": t_run,
            "write_s": t_write, "get_hits_s": t_hits, "symbols": int(my_symbols), "checksums": sums}

    checked = 0
    if rank == 0:
        import oracle_lib as O
        t4 = time.perf_counter()
        # a replay needs ALL hits of a diagonal: keep diagonals that lie wholly in this rank's columns
        col_lo, col_hi = seg_lo * synth.SEGMENT_SYMBOLS, seg_hi * synth.SEGMENT_SYMBOLS
        cell, row = hits & np.uint64(0x3FFF), (hits >> np.uint64(40)).astype(np.int64)
        diag = (seg * np.uint64(12288) + cell).astype(np.int64) - row
        inside = (diag >= col_lo) & (diag + rows <= col_hi)
        ds, rs = diag[inside], row[inside]
        order = np.argsort(ds, kind="stable")
        ds, rs = ds[order], rs[order]

        def gpu_rows_on(d):
            a, b = np.searchsorted(ds, [d, d + 1])
            return set(rs[a:b].tolist())

        gg = np.random.Generator(np.random.PCG64(12))
        picks = ds[gg.integers(0, len(ds), size=300)].tolist() if len(ds) else []   # diagonals that carry hits
        picks += gg.integers(col_lo, col_hi - rows, size=200).tolist()            # uniformly sampled diagonals
        buf = np.zeros(1 << 16, dtype=np.uint32)
        for d in picks:
            want = O.replay_diagonal(packed, table, int(d), buf)
            got = gpu_rows_on(int(d))
            assert want == got, f"diagonal {d}: GPU {sorted(got)[:6]} vs replay {sorted(want)[:6]}"
            checked += 1
        mine["replay_s"] = time.perf_counter() - t4

    if dist is not None:
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
    else:
        gathered = [mine]
    if rank == 0:
        worst = max(r["device_ms"] for r in gathered)
        cells = float(N_SYMBOLS) * rows
        total = [[0, 0, 0] for _ in range(PARTS)]
        for r in gathered:
            for k, (n, x, s) in enumerate(r["checksums"]):
                total[k][0] += n
                total[k][1] ^= x
                total[k][2] = (total[k][2] + s) % (1 << 64)
        out = {"config": "configs[2]: 3 Gbp x 3000 pHMMs, p=0.02", "n_gpus": world, "symbols": N_SYMBOLS, "models": N_MODELS,
               "rows": int(rows), "cells": cells, "device_ms_max": worst, "gcups": cells / (worst * 1e-3) / 1e9,
               "hits": int(sum(r["hits"] for r in gathered)), "diagonals_replayed_rank0": checked,
               "inputs_s": t_inputs, "per_rank": [{k: v for k, v in r.items() if k != "checksums"} for r in gathered],
               "checksums_per_eighth": total}
        line = json.dumps(out)
        print(line, flush=True)
        d = os.path.join(ROOT, "gpurun_out")
        if os.path.isdir(d):
            with open(os.path.join(d, f"c3_full_{world}gpu.json"), "w") as fh:
                fh.write(line + "\n")
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
