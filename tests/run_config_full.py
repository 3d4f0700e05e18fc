"""BASELINE.json configs[2], [3] and [4] at FULL size, sharded over the ranks of this launch (one rank per GPU
under torch.distributed.run; a plain `python tests/run_config_full.py --config c4` is the 1-GPU run):

  c3       3 Gbp synthetic DNA x 3,000 random pHMMs (log-normal lengths, mean ~400, P_total ~1.19 M rows), p = 0.02
  c4       1 Gbp x 200 pHMMs of 2,000-4,000 positions (P_total ~0.59 M rows: long diagonals, 0.59 M-symbol halos)
  c5:P     1 Gbp x 500 pHMMs of 100-800 positions at p = P (0.02 / 0.05 / 0.1: hit-report throughput)

Every rank builds the same inputs from fixed seeds, searches its segment range (left halo
rows-1, no exchange) and reports count / xor / sum of its raw hit words per 1/8th of the segment
axis, so the 1-GPU and 8-GPU hit sets can be compared without moving 2 GB of hits around
(SURVEY.md 8c-3). Rank 0 also replays diagonals on the CPU oracle (test harness code): sampled hit-carrying
ones, uniformly sampled ones, and the ones on the edges of the launch's tiles / threads / register halves.
One JSON line to stdout and to gpurun_out/<config>_full_<N>gpu.json."""
import argparse
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

PARTS = 8


def lognormal_lengths(seed, count, mean, lo, hi):
    g = synth.rng(seed, 21)
    sigma = 0.6
    x = g.lognormal(np.log(mean) - sigma * sigma / 2, sigma, size=count)
    return np.clip(np.round(x), lo, hi).astype(np.int64)


def config(name):
    """-> (label, symbols, model lengths, p-value, seed, planted copies per model, hit capacity of the whole job)"""
    if name == "c3":
        return ("configs[2]: 3 Gbp x 3000 pHMMs, p=0.02", 3_000_000_000, lognormal_lengths(3, 3000, 400, 50, 3000), 0.02, 3, 1,
                400 << 20)
    if name == "c4":
        return ("configs[3]: 1 Gbp x 200 pHMMs of 2000-4000, p=0.02", 1_000_000_000, synth.model_lengths(4, 200, 2000, 4000), 0.02,
                4, 20, 64 << 20)
    if name.startswith("c5"):
        p = float(name.split(":")[1]) if ":" in name else 0.02
        return (f"configs[4]: 1 Gbp x 500 pHMMs of 100-800, p={p:g}", 1_000_000_000, synth.model_lengths(5, 500, 100, 800), p, 5, 20,
                {0.02: 64 << 20, 0.05: 128 << 20}.get(p, 320 << 20))
    raise SystemExit(f"unknown config {name}")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c3")
    ap.add_argument("--symbols", type=float, default=0, help="override the sequence length (smoke runs)")
    args = ap.parse_args()
    label, N_SYMBOLS, lengths, p_value, seed, copies, hitcap_total = config(args.config)
    if args.symbols:
        N_SYMBOLS = int(args.symbols)
    N_MODELS = len(lengths)
    tag = args.config.replace(":", "_p")
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
    hmm = f"/tmp/{tag}_models.hmm"
    models = synth.random_models(3000 if args.config == "c3" else 1000 + N_MODELS, lengths)
    if rank == 0:
        synth.write_hmm(hmm + ".tmp", models)
        os.replace(hmm + ".tmp", hmm)
    if dist is not None:
        dist.barrier()
    table, lens = havac.project_phmm(hmm, p_value)
    rows = table.shape[0]
    packed = synth.random_packed(seed, N_SYMBOLS)
    g = synth.rng(seed, 9)
    for m in [m for m in models for _ in range(copies)]:  # 15 %-mutated consensus copies of every model
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
    hitcap = hitcap_total // world + (8 << 20)
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
        geom = c.tileGeometry()
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
        n_sampled = len(picks)
        picks += [d for d in O.edge_diagonals(geom, rows, col_hi, 6, 99) if col_lo <= d and d + rows <= col_hi]
        mine["diagonals_edge"] = len(picks) - n_sampled
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
        out = {"config": label, "p_value": p_value, "n_gpus": world, "symbols": N_SYMBOLS, "models": N_MODELS,
               "rows": int(rows), "cells": cells, "device_ms_max": worst, "gcups": cells / (worst * 1e-3) / 1e9,
               "hits": int(sum(r["hits"] for r in gathered)), "hits_per_gcell": sum(r["hits"] for r in gathered) / (cells / 1e9),
               "diagonals_replayed_rank0": checked, "of_which_on_tile_thread_edges": gathered[0].get("diagonals_edge", 0),
               "inputs_s": t_inputs, "per_rank": [{k: v for k, v in r.items() if k != "checksums"} for r in gathered],
               "checksums_per_eighth": total}
        line = json.dumps(out)
        print(line, flush=True)
        d = os.path.join(ROOT, "gpurun_out")
        if os.path.isdir(d):
            with open(os.path.join(d, f"{tag}_full_{world}gpu.json"), "w") as fh:
                fh.write(line + "\n")
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
