"""Host-side logic of the one-process-per-GPU launch: how the packed sequence is cut into
per-rank segment ranges, and how the ranks' raw hit lists are merged.

Every diagonal of the SSV matrix is an independent scan (device/HavacHls.cpp:370-402), so ranks
exchange nothing while searching: rank r searches segments [lo_r, hi_r) with a left halo of
rows-1 symbols that the device client adds itself (havac_cuda_set_segment_range). The only
cross-rank step is gathering the hits, which is a host-side concatenation; no NCCL collective
is on the data path.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np

SEGMENT_SYMBOLS = 12288  # device/PublicDefines.h:18-22


def plan_shards(n_segments: int, world_size: int) -> List[Tuple[int, int]]:
    """Contiguous, near-equal segment ranges [lo, hi) for ranks 0..world_size-1 (some may be
    empty when there are fewer segments than ranks)."""
    if n_segments < 0 or world_size < 1:
        raise ValueError("n_segments must be >= 0 and world_size >= 1")
    return [(n_segments * r // world_size, n_segments * (r + 1) // world_size) for r in range(world_size)]


def halo_symbols(n_rows: int) -> int:
    """Symbols a shard needs left of its first column for exact hits: a cell (p, i) depends on
    (p-k, i-k) down to row 0, and rows never reset between concatenated models
    (test/softSsv/SoftSsv.cpp:31-62), so the halo is the TOTAL table length minus one."""
    return max(0, n_rows - 1)


def emission_order_key(hits: np.ndarray) -> np.ndarray:
    """Sort key reproducing the FPGA's emission order: segment, then model row, then cell
    (device/HavacHls.cpp:131-162 loop nest; hit word layout device/HitReporting.cpp:421-430)."""
    hits = np.asarray(hits, dtype=np.uint64)
    cell = hits & np.uint64(0x3FFF)
    seg = (hits >> np.uint64(14)) & np.uint64(0x3FFFFFF)
    row = hits >> np.uint64(40)
    return (seg << np.uint64(38)) | (row << np.uint64(14)) | cell


def merge_hits(parts: Sequence[np.ndarray]) -> np.ndarray:
    """Concatenate the ranks' raw hit lists and put them in emission order."""
    parts = [np.asarray(p, dtype=np.uint64) for p in parts if len(p)]
    if not parts:
        return np.zeros(0, dtype=np.uint64)
    allhits = np.concatenate(parts)
    return allhits[np.argsort(emission_order_key(allhits), kind="stable")]


def gather_hit_lists(hits: np.ndarray, rank: int, world_size: int, group=None, device=None):
    """Every rank's hit list to rank 0. Each rank's list is already in emission order (the device sorts it) and ranks
    hold increasing segment ranges, so laying the lists end to end in rank order IS the global emission order -- no
    merge on rank 0. Returns (the whole list on rank 0 / None elsewhere, bytes received by rank 0).

    Two roads, same result: through HOST memory over `group` (gloo; the CPU test and the fallback), or -- with
    `device` a CUDA device of this rank and the default process group on NCCL -- GPU to GPU over NVLink into one
    buffer on rank 0's GPU and from there in ONE copy into rank 0's pinned host memory (at 8 ranks the host road
    spends 6 ms in eight serial TCP receives, this one about 1 ms)."""
    import torch
    import torch.distributed as dist

    hits = np.ascontiguousarray(hits, dtype=np.uint64)
    if world_size == 1:
        return hits, 0
    if device is not None and dist.get_backend() == "nccl":
        mine = torch.from_numpy(hits.view(np.int64)).to(device, non_blocking=True)
        counts_dev = torch.empty(world_size, dtype=torch.int64, device=device)
        dist.all_gather_into_tensor(counts_dev, torch.tensor([len(hits)], dtype=torch.int64, device=device))
        counts = counts_dev.tolist()
        if rank != 0:
            if counts[rank]:
                for w in dist.batch_isend_irecv([dist.P2POp(dist.isend, mine, 0)]):
                    w.wait()
            torch.cuda.synchronize(device)
            return None, 0
        whole = torch.empty(sum(counts), dtype=torch.int64, device=device)
        whole[: counts[0]] = mine
        ops, at = [], counts[0]
        for r in range(1, world_size):
            if counts[r]:
                ops.append(dist.P2POp(dist.irecv, whole[at: at + counts[r]], r))
            at += counts[r]
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        out = torch.empty(sum(counts), dtype=torch.int64, pin_memory=True)
        out.copy_(whole, non_blocking=True)
        torch.cuda.synchronize(device)
        return out.numpy().view(np.uint64), 8 * (len(out) - counts[0])
    counts = [None] * world_size
    dist.all_gather_object(counts, int(len(hits)), group=group)
    if rank != 0:
        if len(hits):
            dist.send(torch.from_numpy(hits.view(np.int64)), dst=0, group=group)
        return None, 0
    out = np.empty(sum(counts), dtype=np.uint64)
    out[: counts[0]] = hits
    at = counts[0]
    for r in range(1, world_size):
        if counts[r]:
            dist.recv(torch.from_numpy(out[at: at + counts[r]].view(np.int64)), src=r, group=group)
        at += counts[r]
    return out, 8 * (len(out) - counts[0])
