"""Read sharding across GPUs: reads are independent, so a batch is
cut into contiguous read ranges, one per rank; every rank compresses its own
shard, and the only exchange is an all-gather of each rank's per-read byte
positions, from which every rank derives the byte offset of every read in ONE
contiguous container (rank-major).  Decompression needs no exchange.

The data path of this module is torch.distributed only (NCCL on GPUs, gloo in the
CPU tests); the index itself is computed on the device by
press_gpu_global_offsets (csrc/press_gpu.cu) -- `global_index` below is its host
twin, used by the tests and for host-side bookkeeping.
"""
from __future__ import annotations

import numpy as np


def partition_reads(lengths: np.ndarray, nranks: int) -> list[tuple[int, int]]:
    """Contiguous read ranges [r0, r1) per rank with balanced sample counts
    (cut where the prefix sum of the lengths crosses k/nranks of the total)."""
    lengths = np.asarray(lengths, dtype=np.int64)
    cum = np.concatenate([[0], np.cumsum(lengths)])
    total = int(cum[-1])
    cuts = [0]
    for k in range(1, nranks):
        c = int(np.searchsorted(cum, total * k / nranks, side="left"))
        cuts.append(min(max(c, cuts[-1]), lengths.size))
    cuts.append(int(lengths.size))
    return [(cuts[k], cuts[k + 1]) for k in range(nranks)]


def equal_count_ranges(nreads: int, nranks: int) -> list[tuple[int, int]]:
    """Equal read counts per rank (the last ranks may get one read fewer): the
    layout the fixed-size all-gather uses."""
    per = -(-nreads // nranks)
    return [(min(nreads, k * per), min(nreads, (k + 1) * per)) for k in range(nranks)]


def global_index(gathered: np.ndarray) -> np.ndarray:
    """Host twin of press_gpu_global_offsets: gathered[k] = rank k's out_offsets
    ([nreads + 1], last = its total bytes) -> byte offset of read r of rank k at
    [k * nreads + r], total container bytes at [nranks * nreads]."""
    gathered = np.asarray(gathered, dtype=np.uint64)
    nranks, n1 = gathered.shape
    nreads = n1 - 1
    totals = gathered[:, nreads]
    base = np.concatenate([[0], np.cumsum(totals)]).astype(np.uint64)
    out = np.empty(nranks * nreads + 1, dtype=np.uint64)
    for k in range(nranks):
        out[k * nreads:(k + 1) * nreads] = base[k] + gathered[k, :nreads]
    out[nranks * nreads] = base[nranks]
    return out


def allgather_offsets(local_offsets, world: int):
    """All-gather of one rank's out_offsets tensor ([nreads + 1] int64, the same
    length on every rank) -> [world * (nreads + 1)] on every rank.  NCCL for CUDA
    tensors, gloo for CPU tensors."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return local_offsets.clone()
    out = torch.empty(world * local_offsets.numel(), dtype=local_offsets.dtype, device=local_offsets.device)
    if local_offsets.is_cuda:
        dist.all_gather_into_tensor(out, local_offsets)
    else:
        parts = [torch.empty_like(local_offsets) for _ in range(world)]
        dist.all_gather(parts, local_offsets)
        out = torch.cat(parts)
    return out
