"""Slot-range sharding across GPUs, one process per GPU; used by bench.py (NCCL) and the CPU tests (gloo).

Rank r mirrors slots [r*ceil(N/W), ...) and searches only those.  Per batch of queries the ranks exchange ONE packed
block each -- the [nq][k] result records followed by the [nq] hit counts (include/splinter_gpu.h,
SPL_GPU_PACKED_BLOCK_BYTES) -- with a single all-gather, and every rank merges the W blocks on its own device with
the reference's comparator + global slot index (spl_gpu_merge_topk_packed_device).  There is no other data-path
collective.  A query that some shard could not answer in the requested mode (SPL_GPU_COUNT_UNANSWERED: bf16
selection not provable, candidate list overflow) comes out of the merge marked the same way on EVERY rank -- the
merge is a pure function of the gathered bytes -- so all ranks agree, without further communication, on which
queries to run again on the fp32 scan.
"""
from __future__ import annotations

import copy

import numpy as np
import torch
import torch.distributed as dist

from .binding import COUNT_UNANSWERED, MODE_FAST, packed_block_bytes, packed_counts_offset


def shard_range(total_slots: int, world: int, rank: int) -> tuple[int, int]:
    """(first slot, slot count) of `rank`; contiguous, ceil-divided, last shard may be short or empty."""
    per = -(-total_slots // world)
    lo = min(rank * per, total_slots)
    hi = min(lo + per, total_slots)
    return lo, hi - lo


class ShardedSearcher:
    """One rank's end of a sharded search with fixed (nq, k, order).

    search_fn(params, q, block)  enqueue the shard search of `q` ([nq, dim] fp32 on the device) writing records to the
                                 start of `block` (uint8 [block_bytes]) and counts at packed_counts_offset(nq, k)
    merge_fn(gathered, nq, fin, fcnt)  enqueue the merge of `world` packed blocks into fin [nq, k, 32] / fcnt [nq]
    Both default to the mirror's device-pointer entry points on `stream` (a cudaStream_t as int); the CPU tests inject
    oracle-backed stand-ins, so that the block layout, the single gather and the re-issue logic run under gloo.
    """

    def __init__(self, mirror, world: int, nq: int, k: int, order: int, device, stream: int = 0, search_fn=None, merge_fn=None):
        self.m, self.world, self.nq, self.k, self.order, self.stream = mirror, world, nq, k, order, stream
        self.device = device
        self._search = search_fn or self._search_device
        self._merge = merge_fn or self._merge_device
        self._bufs = {}
        self.fin, self.fcnt = self._buffers(nq)[2:]
        self.reissued = 0          # queries answered by the fp32 re-issue so far

    def _buffers(self, nq):
        if nq not in self._bufs:
            nbytes = packed_block_bytes(nq, self.k)
            z = lambda *shape, dtype=torch.uint8: torch.zeros(shape, dtype=dtype, device=self.device)  # noqa: E731
            self._bufs[nq] = (z(nbytes), z(self.world * nbytes), z(nq, self.k, 32), z(nq, dtype=torch.int32))
        return self._bufs[nq]

    # ---- defaults: the C ABI's device-pointer calls ----
    def _search_device(self, params, q, block):
        nq = int(params.nq)
        self.m.search_device(params, q.data_ptr(), block.data_ptr(), self.k, block.data_ptr() + packed_counts_offset(nq, self.k),
                             self.stream)

    def _merge_device(self, gathered, nq, fin, fcnt):
        self.m.merge_topk_packed_device(gathered.data_ptr(), self.world, nq, self.k, self.order, self.k, fin.data_ptr(), self.k,
                                        fcnt.data_ptr(), self.stream)

    def enqueue(self, params, q):
        """shard search -> one all-gather -> merge, all enqueued on the current stream; returns (fin, fcnt) device tensors"""
        nq = int(params.nq)
        block, gathered, fin, fcnt = self._buffers(nq)
        self._search(params, q, block)
        if self.world > 1:
            dist.all_gather_into_tensor(gathered, block)
        else:
            gathered = block
        self._merge(gathered, nq, fin, fcnt)
        return fin, fcnt

    def search(self, params, q):
        """enqueue + wait + re-issue what no mode but the fp32 scan can answer.  Returns (records [nq, k] structured
        numpy array, counts [nq] uint32), identical on every rank."""
        from .binding import RESULT_DTYPE
        fin, fcnt = self.enqueue(params, q)
        counts = fcnt.cpu().numpy().astype(np.uint32)          # (synchronises with the stream the merge ran on)
        recs = fin.cpu().numpy().reshape(int(params.nq), self.k * 32).view(RESULT_DTYPE).copy()
        redo = np.nonzero(counts == COUNT_UNANSWERED)[0]
        if len(redo):
            p1 = copy.copy(params)
            p1.nq, p1.mode = 1, MODE_FAST
            for qi in redo:                                     # the same list on every rank: see the module docstring
                f1, c1 = self.enqueue(p1, q[int(qi):int(qi) + 1])
                counts[qi] = np.uint32(int(c1.cpu()[0]))
                recs[qi] = f1.cpu().numpy().reshape(self.k * 32).view(RESULT_DTYPE)
            self.reissued += len(redo)
        return recs, counts


# ---- the two-collective helpers of round 1, kept for callers that hold separate record / count tensors ----
def gather_lists(local_records: torch.Tensor, local_count: torch.Tensor, world: int):
    """all-gather one [k, 32]-byte record list and its count from every rank.

    Returns (gathered [world, k, 32] uint8, counts [world] int32) on the same device as the inputs."""
    k = local_records.shape[0]
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
