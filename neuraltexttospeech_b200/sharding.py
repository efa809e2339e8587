"""Utterance sharding across GPUs.

The generator forward pass has no batch-coupled op (reference HiFiGAN_TF/model.py:155-173), so N GPUs
simply take N disjoint sets of utterances; there is no collective on the data path.  This module holds
the host-side logic: which utterances a rank takes, and an optional off-path gather of the waveforms to
one rank over `torch.distributed` (NCCL on GPUs, gloo in the CPU tests)."""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np


def shard_utterances(lengths: Sequence[int], world_size: int, rank: int, balance: bool = False) -> List[int]:
    """Indices of the utterances rank `rank` synthesises.

    balance=False: contiguous shards whose sizes differ by at most one (what bench.py's fixed-length
    batches use).  balance=True: ragged batches -- sort by length (longest first, stable) and deal
    round-robin so every rank gets a similar number of frames."""
    n = len(lengths)
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError(f"bad rank/world_size {rank}/{world_size}")
    if not balance:
        base, extra = divmod(n, world_size)
        start = rank * base + min(rank, extra)
        return list(range(start, start + base + (1 if rank < extra else 0)))
    order = sorted(range(n), key=lambda i: (-int(lengths[i]), i))
    return sorted(order[rank::world_size])


def group_by_length(lengths: Sequence[int], indices: Sequence[int]) -> List[List[int]]:
    """Groups a rank's utterances into equal-length batches (Generator.call takes a dense [B,T,80])."""
    groups = {}
    for i in indices:
        groups.setdefault(int(lengths[i]), []).append(i)
    return [groups[t] for t in sorted(groups, reverse=True)]


def gather_waveforms(local: np.ndarray, indices: Sequence[int], total: int, dst: int = 0) -> Optional[np.ndarray]:
    """Off-path gather of equal-length waveforms [b_local, L, 1] to rank `dst`, placed at their original
    utterance indices.  Needs an initialised torch.distributed process group; returns None on other ranks."""
    import torch
    import torch.distributed as dist
    world, rank = dist.get_world_size(), dist.get_rank()
    payload = (list(int(i) for i in indices), np.ascontiguousarray(local, dtype=np.float32))
    gathered = [None] * world if rank == dst else None
    dist.gather_object(payload, gathered, dst=dst)
    if rank != dst:
        return None
    L = max(p[1].shape[1] for p in gathered if len(p[0]))
    out = np.zeros((total, L, 1), dtype=np.float32)
    for idx, wav in gathered:
        if len(idx):
            out[np.asarray(idx)] = wav
    return out
