"""Slot-range sharding plumbing shared by bench.py (NCCL, one rank per GPU) and the CPU tests (gloo).

The hot path shards by contiguous slot range: rank r mirrors slots
[r*ceil(N/W), ...), searches its shard, and the per-rank top-k record lists (32 bytes x k) are
all-gathered; every rank then merges W sorted lists with the reference comparator + slot index.
No data-path collective other than that gather exists.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(total_slots: int, world: int, rank: int) -> tuple[int, int]:
    """(first slot, slot count) of `rank`; contiguous, ceil-divided, last shard may be short or empty."""
    per = -(-total_slots // world)
    lo = min(rank * per, total_slots)
    hi = min(lo + per, total_slots)
    return lo, hi - lo


def gather_lists(local_records: torch.Tensor, local_count: torch.Tensor, world: int):
    """all-gather one [k, 32]-byte record list and its count from every rank.

    Returns (gathered [world, k, 32] uint8, counts [world] int32) on the same device as the inputs."""
    k = local_records.shape[0]
    # concatenated along dim 0 (the form both NCCL and gloo accept), viewed as [world, k, 32]
    flat = torch.empty((world * k, local_records.shape[1]), dtype=local_records.dtype, device=local_records.device)
    counts = torch.empty(world, dtype=local_count.dtype, device=local_count.device)
    if world == 1:
        flat.copy_(local_records)
        counts.copy_(local_count)
    else:
        dist.all_gather_into_tensor(flat, local_records)
        dist.all_gather_into_tensor(counts, local_count)
    return flat.view(world, k, local_records.shape[1]), counts
