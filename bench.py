#!/usr/bin/env python
"""bench.py -- SSV throughput (GCUPS) of the B200 path, one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A "step" is one pass of the SSV hot path over one batch: the whole packed sequence shard of
this rank against the whole concatenated score table (one `invokeHavacSsvAsync` + wait).
Workload at N=1 is BASELINE.json configs[1]: 100 Mbp synthetic DNA x 100 random models of
100-600 positions (P_total ~ 35 k rows), threshold projected to 256 as at p = 0.02. For
N>1 (launched by torch.distributed.run, one rank per GPU) the database grows with N -- each
rank owns 100 Mbp of an N x 100 Mbp sequence plus a left halo of P_total-1 symbols
 -- so per-GPU work is fixed ("weak") and no collective is on the
data path; torch.distributed only carries the barrier and the max-over-ranks of the timings.

    value     device-timed GCUPS (CUDA events on the launching stream inside the C-ABI client,
              inputs resident in HBM, L2 flushed between steps), all ranks' cells / max time.
    e2e       the same metric through the reference-facing client calls with HOST buffers:
              writeSequence + writePhmm (H2D) + invoke + wait + getHitList (D2H) every step.
    roofline  integer/half-SIMD issue roofline of the dominant kernel: the
              kernel needs one lane-instruction per cell, and an SM issues 128 lane-instructions
              per clock (4 schedulers x 32 lanes), so peak cells/s = SMs x f_clk x 128.
              An `hbm` sub-object shows how far the path is from the HBM roofline.
    cpu_baseline  the reference's own softSsvThreshold256 (oracle/_ref, compiled from
              /root/reference) or, where that is absent, the C restatement (oracle/), on the
              host cores, on a bounded slice of the same workload (250 k symbols per host thread x the
              first 8000 rows: the ratio to it moves with the host's core count, so the per-core rate
              is printed next to it). The same leg then replays sampled and tile-edge diagonals of
              rank 0's hit list on that CPU code (`parity_check`): the line certifies its own hits.
    strong    a second timed leg on ONE fixed database whatever N is: 320 Mbp x configs[2]'s table
              (3,000 log-normal models, P_total ~ 1.19 M rows), cut into N shards with a 1.19 M-symbol
              halo each -- the "sequence database sharded across the GPUs" of north_star, where
              `value` above is N replicas of configs[1]. Hit count must be the same at every N.
    e2e at N > 1 ends with ONE hit list on rank 0: every rank's device-sorted list is read back and
              sent to rank 0 (gloo, host memory), where laying them end to end in rank order is the
              global emission order.

`--impl reference` times that CPU implementation alone, on all host threads, same metric.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

from havac_b200 import sharding, synth  # noqa: E402

METRIC = "SSV GCUPS"
UNIT = "GCUPS"
SEQ_PER_GPU = 100_000_000  # configs[1]: 100 Mbp
N_MODELS, LEN_LO, LEN_HI = 100, 100, 600
SEED = 2
LANES_PER_SM_CLK = 128  # 4 schedulers x 32 lanes, one warp-instruction per scheduler per clock
SM_COUNT_B200 = 148
PUBLISHED_FPGA_GCUPS = 1739.0  # README.md:4 of the reference


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ------------------------------------------------------------------------------------------
# workload
# ------------------------------------------------------------------------------------------
def _plant(packed, n, models, copies_each, g):
    """Overwrite windows of the packed sequence with 15 %-mutated copies of every model's consensus."""
    for m in models:
        for _ in range(copies_each):
            frag = m.consensus.copy()
            mut = g.random(len(frag)) < 0.15
            frag[mut] = g.integers(0, 4, size=int(mut.sum()))
            at = int(g.integers(0, n - len(frag) - 8)) & ~3
            nb = (len(frag) + 3) // 4
            window = synth.unpack_2bit(packed[at // 4: at // 4 + nb])
            window[: len(frag)] = frag
            packed[at // 4: at // 4 + nb] = synth.pack_2bit(window, pad_to_segment=False)


def workload_hmm_path(tag="c2"):
    return os.path.join(os.environ.get("TMPDIR", "/tmp"), f"havac_bench_{tag}_{os.getuid()}_{os.getpid()}.hmm")


def build_table(lengths, model_seed, tag):
    """Random HMMER3 models -> .hmm file -> the reference's threshold-256 projection at p = 0.02 (host C++,
    bit-identical to PhmmReprojection.cpp:109-145, checked in tests/test_host_contracts.py). This is the
    path a user's table takes; the bench does not time it."""
    from havac_b200 import havac

    models = synth.random_models(model_seed, lengths)
    path = workload_hmm_path(tag)
    synth.write_hmm(path, models)
    try:
        table, lens = havac.project_phmm(path, 0.02)
    finally:
        if not os.environ.get("HAVAC_BENCH_KEEP_HMM"):
            os.unlink(path)
    assert lens.tolist() == [int(x) for x in lengths]
    return np.ascontiguousarray(table), models


def build_workload(n_gpus: int):
    """Packed N x 100 Mbp sequence and the int8 score table of 100 projected models (configs[1])."""
    lengths = synth.model_lengths(SEED, N_MODELS, LEN_LO, LEN_HI)
    table, models = build_table(lengths, 1000 + SEED, "c2")
    n = SEQ_PER_GPU * n_gpus
    packed = synth.random_packed(SEED, n)
    # one mutated copy of every model's consensus per 10 Mbp so the hit list is not empty
    _plant(packed, n, models, max(1, n // 10_000_000), synth.rng(SEED, 9))
    return packed, table, n, [int(x) for x in lengths]


STRONG_SYMBOLS = 320_000_000
STRONG_MODELS = 3000


def lognormal_lengths(seed, count, mean, lo, hi):
    g = synth.rng(seed, 21)
    sigma = 0.6
    x = g.lognormal(np.log(mean) - sigma * sigma / 2, sigma, size=count)
    return np.clip(np.round(x), lo, hi).astype(np.int64)


def build_strong_workload():
    """The fixed database of the strong-scaling leg: 320 Mbp x configs[2]'s table shape."""
    lengths = lognormal_lengths(3, STRONG_MODELS, 400, 50, 3000)
    table, models = build_table(lengths, 3000, "c3")
    packed = synth.random_packed(31, STRONG_SYMBOLS)
    _plant(packed, STRONG_SYMBOLS, models, 1, synth.rng(31, 9))
    return packed, table, STRONG_SYMBOLS


def cpu_ssv_rate(packed, table, n_symbols, rows, threads, budget_cells):
    """Times the CPU SSV on symbols [0, n_symbols) x the first `rows` rows, split over
    `threads` with a rows-1 left halo per chunk. Returns (gcups, kind, seconds, hits)."""
    import oracle_lib as O

    lib = O.ref()
    kind = "reference" if lib is not None else "port"
    sym = synth.unpack_2bit(packed[: (n_symbols + 3) // 4], n_symbols)
    tab = np.ascontiguousarray(table[:rows])
    t0 = time.perf_counter()
    pos, row = O.soft_ssv_parallel(sym, tab, threads, lib=lib)
    dt = time.perf_counter() - t0
    return n_symbols * rows / dt / 1e9, kind, dt, len(pos)


# ------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def stop(self, t_begin, t_end):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = [l for (t, l) in self.lines if t_begin <= t <= t_end] or [l for (_, l) in self.lines]
        for l in rows:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------
# arms
# ------------------------------------------------------------------------------------------
CPU_SYMBOLS_PER_THREAD = 250_000  # the CPU sample is fixed per host thread, so a step costs ~8 s whatever the core count
CPU_ROWS = 8000


def cpu_sample(threads, rows_total):
    n_s, rows_s = CPU_SYMBOLS_PER_THREAD * threads, min(CPU_ROWS, rows_total)
    text = (f"symbols [0,{n_s}) x rows [0,{rows_s}) of the workload per step ({CPU_SYMBOLS_PER_THREAD} symbols per host thread), "
            f"split over {threads} threads with a rows-1 halo")
    return n_s, rows_s, text


def run_reference(args, rank):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    packed, table, n, lengths = build_workload(max(1, args.gpus))  # the arm's own workload at this N
    rows_total = table.shape[0]
    n_s, rows_s, sample = cpu_sample(threads, rows_total)
    vals = []
    for i in range(args.warmup + args.steps):
        g, kind, dt, nh = cpu_ssv_rate(packed, table, n_s, rows_s, threads, 0)
        log(f"[reference] step {i}: {dt:.2f} s, {g:.3f} GCUPS, {nh} hits ({kind}, {threads} threads)")
        if i >= args.warmup:
            vals.append((g, dt))
    value = float(np.mean([v for v, _ in vals]))
    ms = float(np.mean([d for _, d in vals]) * 1e3)
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": value / PUBLISHED_FPGA_GCUPS, "dtype": "u8", "data": "synthetic",
        "config": {"workload": workload_name(max(1, args.gpus), rows_total), "sample": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample,
                         "per_core_value": value / threads},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)


def traffic_from_profile(symbols, rows):
    """dram__bytes_read.sum + dram__bytes_write.sum of one ssv_f16x2_kernel launch, from the committed
    `ncu --set full` capture of this workload (profiles/r2_ncu_traffic.json, else round 1's); None when the
    capture is of another shape."""
    for name in ("r2_ncu_traffic.json", "r1_ncu_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as fh:
                t = json.load(fh)
            if t.get("symbols") == symbols and t.get("rows") == rows:
                return {"bytes_per_launch": t["dram_bytes_read"] + t["dram_bytes_write"], "source": t.get("source")}
        except (OSError, ValueError, KeyError):
            pass
    return None


def workload_name(n_gpus, rows):
    return (f"C2: {n_gpus} x 100 Mbp synthetic DNA x {N_MODELS} random HMMER3 models of {LEN_LO}-{LEN_HI} positions "
            f"(P_total={rows}), written to a .hmm file and projected to threshold 256 at p=0.02 by the host driver")


def parity_self_check(packed, table, hits, geom, col_hi):
    """Rank 0's hit list against the CPU code (outside every timed region): hit-carrying, uniformly sampled and
    tile/thread-edge diagonals that lie wholly inside rank 0's columns must replay to exactly the GPU's rows."""
    import oracle_lib as O

    rows = table.shape[0]
    t0 = time.perf_counter()
    cell, seg, row = hits & np.uint64(0x3FFF), (hits >> np.uint64(14)) & np.uint64(0x3FFFFFF), (hits >> np.uint64(40)).astype(np.int64)
    diag = (seg * np.uint64(12288) + cell).astype(np.int64) - row
    inside = (diag >= 0) & (diag + rows <= col_hi)
    ds, rs = diag[inside], row[inside]
    order = np.argsort(ds, kind="stable")
    ds, rs = ds[order], rs[order]
    g = np.random.Generator(np.random.PCG64(7))
    picks = ds[g.integers(0, len(ds), size=60)].tolist() if len(ds) else []
    picks += g.integers(0, max(1, col_hi - rows), size=40).tolist()
    picks += [d for d in O.edge_diagonals(geom, rows, col_hi, 3, 11) if d >= 0 and d + rows <= col_hi]
    buf = np.zeros(1 << 16, dtype=np.uint32)
    bad = 0
    for d in picks:
        a, b = np.searchsorted(ds, [d, d + 1])
        if O.replay_diagonal(packed, table, int(d), buf) != set(rs[a:b].tolist()):
            bad += 1
    return {"diagonals_replayed": len(picks), "mismatches": bad, "seconds": round(time.perf_counter() - t0, 2),
            "checker": "oracle/soft_ssv.c replay_diagonal (test/hmmerSsvRef/hmmerSsvRef.cpp:504-586 walk)"}


def run_gpu(args, rank, world, local_rank):
    import torch

    from havac_b200.client import STATE_COMPLETED, HavacHwClient

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the SSV path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    gloo = None
    stdout_fd = None
    if world > 1:
        # NCCL writes its version banner to file descriptor 1; stdout must carry the one JSON line
        # only, so everything below goes to stderr until rank 0 prints the result
        sys.stdout.flush()
        stdout_fd = os.dup(1)
        os.dup2(2, 1)
        import torch.distributed as dist  # noqa: F811
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        gloo = dist.new_group(backend="gloo")  # host-memory traffic: the hit lists gathered on rank 0

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(x, op):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=getattr(dist.ReduceOp, op))
        return float(t.item())

    def gather_hits(h):
        """One hit list in rank 0's host memory (sharding.gather_hit_lists: GPU to GPU over NCCL, one copy down; rank order =
        emission order)."""
        return sharding.gather_hit_lists(h, rank, world, gloo, device=(torch.device("cuda", local_rank) if world > 1 else None))

    t_setup = time.perf_counter()
    packed, table, n_total, lengths = build_workload(world)
    rows = table.shape[0]
    n_seg = len(packed) * 4 // synth.SEGMENT_SYMBOLS
    seg_lo, seg_hi = sharding.plan_shards(n_seg, world)[rank]
    my_symbols = min(n_total, seg_hi * synth.SEGMENT_SYMBOLS) - seg_lo * synth.SEGMENT_SYMBOLS  # real ones
    log(f"[rank {rank}] workload built in {time.perf_counter() - t_setup:.1f} s: {n_total} symbols, {rows} rows, "
        f"segments [{seg_lo},{seg_hi})")

    # host buffers of the e2e leg live in pinned memory (the copies are counted by the client)
    packed = torch.from_numpy(packed).pin_memory().numpy()
    table = torch.from_numpy(table).pin_memory().numpy()
    t_open = time.perf_counter()
    client = HavacHwClient("", deviceIndex=local_rank)
    device_open_s = reduce(time.perf_counter() - t_open, "MAX")
    if world > 1:
        client.setSegmentRange(seg_lo, seg_hi)
    client.writeSequence(packed)
    client.writePhmm(table)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # 2x the 126 MB L2

    def device_step():
        flush.zero_()
        torch.cuda.synchronize()
        client.invokeHavacSsvAsync()
        st = client.waitForHavacSsvAsync()
        if st != STATE_COMPLETED:
            raise RuntimeError(f"run ended in state {st}")
        return client.deviceTimeMs()

    for _ in range(args.warmup):
        device_step()
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    t_begin = time.perf_counter()
    ms_list = [device_step() for _ in range(args.steps)]
    barrier()
    t_end = time.perf_counter()
    clocks = sampler.stop(t_begin, t_end)
    n_hits = client.getNumHits()
    geom = client.tileGeometry()
    _, launches_per_step = client.runStats()
    # clock64 ticks of the kernel's first CTA over the launch's CUDA-event time (last step), as MHz: the SM clock times
    # the share of the launch that CTA was alive for (persistent CTAs leave when the tiles run out: < 1 means a tail)
    cta0_mhz = client.smCycles() / (ms_list[-1] * 1e3) if ms_list and ms_list[-1] > 0 else None
    ms_dev = reduce(float(np.mean(ms_list)), "MAX")
    cells_all = reduce(float(my_symbols) * rows, "SUM")
    hits_all = reduce(float(n_hits), "SUM")
    value = cells_all / (ms_dev * 1e-3) / 1e9

    # ---- end to end through the client calls with host buffers, ONE hit list on rank 0 -----
    phases = {}
    gathered = {"bytes": 0, "hits": 0, "h2d": 0, "d2h": 0}

    def e2e_step():
        t = [time.perf_counter()]
        client.writeSequence(packed); t.append(time.perf_counter())
        client.writePhmm(table); t.append(time.perf_counter())
        client.invokeHavacSsvAsync()
        if client.waitForHavacSsvAsync() != STATE_COMPLETED:
            raise RuntimeError("e2e run did not complete")
        t.append(time.perf_counter())
        h = client.getHitList(); t.append(time.perf_counter())
        whole, moved = gather_hits(h); t.append(time.perf_counter())
        for name, a, b in zip(("writeSequence", "writePhmm", "run", "getHitList", "gatherOnRank0"), t, t[1:]):
            phases[name] = round((b - a) * 1e3, 3)
        gathered["bytes"] += moved
        if world > 1:  # the gather's own PCIe traffic: every rank's list goes up again, the whole list comes down once
            gathered["h2d"] += int(h.nbytes)
            gathered["d2h"] += int(whole.nbytes) if whole is not None else 0
        if whole is not None:
            gathered["hits"] = int(len(whole))
        return h, whole

    # untimed e2e passes: NCCL sets up its point-to-point channels on first use and the pinned blocks the gathered
    # list lands in are allocated on the first two passes (a block stays referenced until the next list replaces it)
    for _ in range(max(1, args.warmup)):
        hits, whole = e2e_step()
    gathered["bytes"] = gathered["h2d"] = gathered["d2h"] = 0
    h2d0, d2h0 = client.transferStats()
    barrier()
    t0 = time.perf_counter()
    step_ms = []
    slowest = (0.0, {})
    for _ in range(args.steps):
        ts = time.perf_counter()
        hits, whole = e2e_step()
        step_ms.append(round((time.perf_counter() - ts) * 1e3, 2))
        if step_ms[-1] > slowest[0]:
            slowest = (step_ms[-1], dict(phases))
    barrier()
    log(f"[rank {rank}] e2e wall per step (ms): {step_ms}; phases of the slowest step: {slowest[1]}")
    e2e_s = reduce((time.perf_counter() - t0) / args.steps, "MAX")
    h2d1, d2h1 = client.transferStats()
    log(f"[rank {rank}] e2e phases of the last step (ms): {phases}")
    e2e_val = cells_all / e2e_s / 1e9
    h2d = reduce(float(h2d1 - h2d0 + gathered["h2d"]), "SUM") / args.steps
    d2h = reduce(float(d2h1 - d2h0 + gathered["d2h"]), "SUM") / args.steps
    if rank == 0 and whole is not None and len(whole) > 1:  # the gathered list is in emission order as it stands
        key = sharding.emission_order_key(whole).astype(np.int64)
        if not np.all(np.diff(key) > 0):
            raise RuntimeError("gathered hit list is not in emission order")

    # ---- strong scaling: ONE fixed database cut into `world` shards ---------------------------
    strong = None
    if not args.no_strong:
        s_packed, s_table, s_n = build_strong_workload()
        s_rows = s_table.shape[0]
        s_seg = len(s_packed) * 4 // synth.SEGMENT_SYMBOLS
        s_lo, s_hi = sharding.plan_shards(s_seg, world)[rank]
        s_mine = min(s_n, s_hi * synth.SEGMENT_SYMBOLS) - s_lo * synth.SEGMENT_SYMBOLS
        if world > 1:
            client.setSegmentRange(s_lo, s_hi)
        client.writeSequence(s_packed)
        client.writePhmm(s_table)
        device_step()  # one warm-up pass of this shape; the clocks are up from the leg above
        barrier()
        s_ms = [device_step() for _ in range(args.strong_steps)]
        barrier()
        s_ms_dev = reduce(float(np.mean(s_ms)), "MAX")
        s_hits = reduce(float(client.getNumHits()), "SUM")
        t0 = time.perf_counter()
        s_list, _ = gather_hits(client.getHitList())
        s_gather = reduce(time.perf_counter() - t0, "MAX")
        s_value = float(s_n) * s_rows / (s_ms_dev * 1e-3) / 1e9
        strong = {"value": s_value, "unit": UNIT, "ms_per_step": s_ms_dev, "steps": args.strong_steps, "warmup": 1,
                  "workload": f"{s_n // 1_000_000} Mbp x {STRONG_MODELS} log-normal models (P_total={s_rows}), p=0.02 projection, "
                              f"ONE database cut into {world} shard(s)",
                  "symbols": s_n, "rows": s_rows, "halo_symbols": s_rows - 1 if world > 1 else 0,
                  "hits": int(s_hits), "hits_on_rank0_after_gather": (int(len(s_list)) if s_list is not None else None),
                  "read_and_gather_hits_s": s_gather,
                  "efficiency_vs_n1_hint": s_value / (world * (float(my_symbols) * rows / (ms_dev * 1e-3) / 1e9)),
                  "efficiency_hint_basis": "strong GCUPS / (N x this run's per-GPU rate on the weak leg)",
                  "shard_overhead_model": 1.0 / (1.0 + (s_rows * world / (2.0 * s_n) if world > 1 else 0.0))}
        del s_packed, s_list
    client.close()

    if rank == 0:
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
                peaks = json.load(fh)
        except OSError:
            pass
        sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
        f_max = (clocks.get("sm_max_mhz") or peaks.get("sm_max_mhz") or 1965.0) * 1e6
        f_run = (clocks.get("sm_mhz") or 0) * 1e6
        per_gpu_cells_s = float(my_symbols) * rows / (ms_dev * 1e-3)
        peak = sms * f_max * LANES_PER_SM_CLK
        hbm_bytes = my_symbols / 4 + 8 * rows + 8 * n_hits  # sequence + fp16 table + hits, per launch
        hbm_peak = peaks.get("hbm_gbs")
        roofline = {
            "bound": "int_alu_issue", "kernel": "ssv_f16x2_kernel",
            "achieved": per_gpu_cells_s / 1e9, "peak": peak / 1e9, "unit": "Gcell/s (1 lane-instruction per cell)",
            "frac": per_gpu_cells_s / peak, "peak_basis": f"{sms} SMs x {f_max / 1e6:.0f} MHz (max SM clock) x {LANES_PER_SM_CLK} lanes/clk",
            "frac_at_sampled_clock": (per_gpu_cells_s / (sms * f_run * LANES_PER_SM_CLK)) if f_run else None,
            "first_cta_cycles_per_us": cta0_mhz,
            "first_cta_alive_share": (cta0_mhz * 1e6 / f_run) if (cta0_mhz and f_run) else None,
            "traffic": traffic_from_profile(my_symbols, rows),
            "hbm": {"achieved": hbm_bytes / (ms_dev * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                    "frac": (hbm_bytes / (ms_dev * 1e-3) / 1e9 / hbm_peak) if hbm_peak else None,
                    "peak_source": "MEASURED_PEAKS.json (measured)" if hbm_peak else "absent",
                    "algorithmic_bytes_per_launch": hbm_bytes},
        }
        cores = os.cpu_count() or 1
        n_s, rows_s, sample = cpu_sample(cores, rows)
        g, kind, dt, nh = cpu_ssv_rate(packed, table, n_s, rows_s, cores, 0)
        g1, _, dt1, _ = cpu_ssv_rate(packed, table, CPU_SYMBOLS_PER_THREAD, rows_s, 1, 0)
        cpu = {"value": g, "unit": UNIT, "cores": cores, "kind": kind, "seconds": dt, "per_core_value": g / cores,
               "single_thread_value": g1,
               "sample": sample + f"; single thread on [0,{CPU_SYMBOLS_PER_THREAD}) x the same rows",
               "parity_check": parity_self_check(packed, table, hits, geom, seg_hi * synth.SEGMENT_SYMBOLS)}
        if cpu["parity_check"]["mismatches"]:
            raise RuntimeError(f"bench hits disagree with the CPU replay: {cpu['parity_check']}")
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "weak",
            # the reference publishes 1,739 GCUPS for this metric (ONE Alveo U50), so the
            # contract's value / published is the whole job against one card; the per-GPU figure is next to it
            "vs_baseline": value / PUBLISHED_FPGA_GCUPS, "vs_baseline_per_gpu": value / world / PUBLISHED_FPGA_GCUPS,
            "vs_baseline_note": f"{world} GPU(s) against the published figure of ONE Alveo U50 card",
            "dtype": "u8 scores held as f16x2 (exact)", "data": "synthetic",
            "config": {"workload": workload_name(world, rows), "l2": "256 MiB written between steps (L2 flush)",
                       "symbols_per_gpu": SEQ_PER_GPU, "rows": rows, "hits": int(hits_all),
                       "parallelism": f"sequence axis sharded over {world} GPU(s), halo rows-1, no collective"},
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "ms_per_step": e2e_s * 1e3, "ms_per_step_median": float(np.median(step_ms)), "ms_steps": step_ms,
                    "phases_ms": phases, "phases_ms_of_slowest_step": slowest[1],
                    "hits_on_rank0_after_gather": gathered["hits"],
                    "host_gather_bytes_per_step": int(gathered["bytes"] / args.steps),
                    "gather_road": ("NCCL point-to-point into rank 0's GPU, one copy to its pinned host memory" if world > 1 else "none (one rank)"),
                    "ends_with": "one hit list in emission order in rank 0's host memory"},
            "device_open_s": device_open_s,
            "gpu_launches": int(launches_per_step * args.steps * world),
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
            "published_fpga_gcups": PUBLISHED_FPGA_GCUPS,
        }
        if strong is not None:
            out["strong"] = strong
        if stdout_fd is not None:
            sys.stdout.flush()
            os.dup2(stdout_fd, 1)
        print(json.dumps(out), flush=True)
        if stdout_fd is not None:
            os.dup2(2, 1)  # whatever the teardown prints goes to stderr too
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="graft", choices=["graft", "reference"])
    ap.add_argument("--strong-steps", type=int, default=2, help="timed passes of the strong-scaling leg")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling leg (profiling runs)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world == 1 and args.gpus > 1:
        # launched without torchrun: re-launch ourselves one rank per GPU
        port = 29500 + os.getpid() % 2000
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__),
               "--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup", str(args.warmup),
               "--strong-steps", str(args.strong_steps)] + (["--no-strong"] if args.no_strong else [])
        raise SystemExit(subprocess.call(cmd))
    run_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
