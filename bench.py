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
              host cores, on a bounded slice of the same workload.

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
def build_workload(n_gpus: int):
    """Packed N x 100 Mbp sequence and the int8 score table of 100 models."""
    lengths = synth.model_lengths(SEED, N_MODELS, LEN_LO, LEN_HI)
    tables, conses = [], []
    for i, L in enumerate(lengths):
        t, c = synth.consensus_score_table(1000 + i, int(L))
        tables.append(t)
        conses.append(c)
    table = np.concatenate(tables, axis=0)
    n = SEQ_PER_GPU * n_gpus
    packed = synth.random_packed(SEED, n)
    # one mutated copy of every model's consensus per 10 Mbp so the hit list is not empty
    g = synth.rng(SEED, 9)
    for c in conses:
        for _ in range(max(1, n // 10_000_000)):
            frag = c.copy()
            mut = g.random(len(frag)) < 0.15
            frag[mut] = g.integers(0, 4, size=int(mut.sum()))
            at = int(g.integers(0, n - len(frag) - 8)) & ~3
            nb = (len(frag) + 3) // 4
            window = synth.unpack_2bit(packed[at // 4: at // 4 + nb])
            window[: len(frag)] = frag
            packed[at // 4: at // 4 + nb] = synth.pack_2bit(window, pad_to_segment=False)
    return packed, table, n, [int(x) for x in lengths]


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
def run_reference(args, rank):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    packed, table, n, lengths = build_workload(max(1, args.gpus))  # the arm's own workload at this N
    rows_total = table.shape[0]
    # bounded sample: 250 k symbols per host thread (4 Mbp on 16 threads) x the first 8000 rows of the same
    # table per step, ~8 s a step whatever the core count, so the default 13 steps end within two minutes
    n_s, rows_s = 250_000 * threads, min(8000, rows_total)
    vals = []
    for i in range(args.warmup + args.steps):
        g, kind, dt, nh = cpu_ssv_rate(packed, table, n_s, rows_s, threads, 0)
        log(f"[reference] step {i}: {dt:.2f} s, {g:.3f} GCUPS, {nh} hits ({kind}, {threads} threads)")
        if i >= args.warmup:
            vals.append((g, dt))
    value = float(np.mean([v for v, _ in vals]))
    ms = float(np.mean([d for _, d in vals]) * 1e3)
    sample = f"symbols [0,{n_s}) x rows [0,{rows_s}) of the workload per step, split over {threads} threads with a rows-1 halo"
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": value / PUBLISHED_FPGA_GCUPS, "dtype": "u8", "data": "synthetic",
        "config": {"workload": workload_name(max(1, args.gpus), rows_total), "sample": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)


def traffic_from_profile(symbols, rows):
    """dram__bytes_read.sum + dram__bytes_write.sum of one ssv_f16x2_kernel launch, from the committed
    `ncu --set full` capture of this workload (profiles/r1_ncu_traffic.json); None when the capture
    is of another shape."""
    try:
        with open(os.path.join(ROOT, "profiles", "r1_ncu_traffic.json")) as fh:
            t = json.load(fh)
        if t.get("symbols") == symbols and t.get("rows") == rows:
            return {"bytes_per_launch": t["dram_bytes_read"] + t["dram_bytes_write"], "source": t.get("source")}
    except (OSError, ValueError, KeyError):
        pass
    return None


def workload_name(n_gpus, rows):
    return (f"C2: {n_gpus} x 100 Mbp synthetic DNA x {N_MODELS} random models of {LEN_LO}-{LEN_HI} positions "
            f"(P_total={rows}), threshold 256 (p=0.02 projection)")


def run_gpu(args, rank, world, local_rank):
    import torch

    from havac_b200.client import STATE_COMPLETED, HavacHwClient

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the SSV path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    stdout_fd = None
    if world > 1:
        # NCCL writes its version banner to file descriptor 1; stdout must carry the one JSON line
        # only, so everything below goes to stderr until rank 0 prints the result
        sys.stdout.flush()
        stdout_fd = os.dup(1)
        os.dup2(2, 1)
        import torch.distributed as dist  # noqa: F811
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_max(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def reduce_sum(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

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
    client = HavacHwClient("", deviceIndex=local_rank)
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
    _, launches_per_step = client.runStats()
    ms_dev = reduce_max(float(np.mean(ms_list)))
    cells_all = reduce_sum(float(my_symbols) * rows)
    hits_all = reduce_sum(float(n_hits))
    value = cells_all / (ms_dev * 1e-3) / 1e9

    # ---- end to end through the client calls with host buffers ----------------------------
    phases = {}

    def e2e_step():
        t = [time.perf_counter()]
        client.writeSequence(packed); t.append(time.perf_counter())
        client.writePhmm(table); t.append(time.perf_counter())
        client.invokeHavacSsvAsync()
        if client.waitForHavacSsvAsync() != STATE_COMPLETED:
            raise RuntimeError("e2e run did not complete")
        t.append(time.perf_counter())
        h = client.getHitList(); t.append(time.perf_counter())
        for name, a, b in zip(("writeSequence", "writePhmm", "run", "getHitList"), t, t[1:]):
            phases[name] = round((b - a) * 1e3, 3)
        return h

    e2e_step()
    h2d0, d2h0 = client.transferStats()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        hits = e2e_step()
    barrier()
    e2e_s = reduce_max((time.perf_counter() - t0) / args.steps)
    h2d1, d2h1 = client.transferStats()
    log(f"[rank {rank}] e2e phases of the last step (ms): {phases}")
    e2e_val = cells_all / e2e_s / 1e9
    h2d = reduce_sum(float(h2d1 - h2d0)) / args.steps
    d2h = reduce_sum(float(d2h1 - d2h0)) / args.steps
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
            "traffic": traffic_from_profile(my_symbols, rows),
            "hbm": {"achieved": hbm_bytes / (ms_dev * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                    "frac": (hbm_bytes / (ms_dev * 1e-3) / 1e9 / hbm_peak) if hbm_peak else None,
                    "peak_source": "MEASURED_PEAKS.json (measured)" if hbm_peak else "absent",
                    "algorithmic_bytes_per_launch": hbm_bytes},
        }
        cores = os.cpu_count() or 1
        n_s, rows_s = 4_000_000, min(8000, rows)  # ~10 s of host work on 16-24 cores
        g, kind, dt, nh = cpu_ssv_rate(packed, table, n_s, rows_s, cores, 0)
        g1, _, dt1, _ = cpu_ssv_rate(packed, table, 250_000, rows_s, 1, 0)
        cpu = {"value": g, "unit": UNIT, "cores": cores, "kind": kind, "seconds": dt,
               "single_thread_value": g1,
               "sample": f"symbols [0,{n_s}) x rows [0,{rows_s}) of the workload, {cores} threads, rows-1 halo per chunk; "
                         f"single thread on [0,250000) x the same rows"}
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "weak",
            # the reference publishes 1,739 GCUPS for this metric (one Alveo U50)
            "vs_baseline": value / PUBLISHED_FPGA_GCUPS, "dtype": "u8 scores held as f16x2 (exact)", "data": "synthetic",
            "config": {"workload": workload_name(world, rows), "l2": "256 MiB written between steps (L2 flush)",
                       "symbols_per_gpu": SEQ_PER_GPU, "rows": rows, "hits": int(hits_all),
                       "parallelism": f"sequence axis sharded over {world} GPU(s), halo rows-1, no collective"},
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "ms_per_step": e2e_s * 1e3, "phases_ms": phases},
            "gpu_launches": int(launches_per_step * args.steps * world),
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
            "published_fpga_gcups": PUBLISHED_FPGA_GCUPS,
        }
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
               "--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup", str(args.warmup)]
        raise SystemExit(subprocess.call(cmd))
    run_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
