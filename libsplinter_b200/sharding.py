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


def gather_blocks(local_block: torch.Tensor, local_counts: torch.Tensor, world: int, out_block=None, out_counts=None):
    """all-gather one rank's batch of results: local_block [nq*k, 32] uint8 (nq record lists of k), local_counts [nq].

    Returns (gathered [world*nq*k, 32], counts [world*nq]) laid out [rank][query][k] / [rank][query] -- the layout
    spl_gpu_merge_topk_batch_device merges in one launch.  Pass preallocated outputs to avoid per-call allocation."""
    if out_block is None:
        out_block = torch.empty((world * local_block.shape[0], local_block.shape[1]), dtype=local_block.dtype,
                                device=local_block.device)
    if out_counts is None:
        out_counts = torch.empty(world * local_counts.shape[0], dtype=local_counts.dtype, device=local_counts.device)
    if world == 1:
        out_block.copy_(local_block)
        out_counts.copy_(local_counts)
    else:
        dist.all_gather_into_tensor(out_block, local_block)
        dist.all_gather_into_tensor(out_counts, local_counts)
    return out_block, out_counts
