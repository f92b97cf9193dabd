"""world_size-2 (and 3) CPU tests of the multi-GPU plumbing over gloo: slot-range shards, all-gather of
the per-rank top-k record lists, k-way merge = the single-store result.  The per-rank search here is the
CPU oracle (no GPU in this container); on the GPU box bench.py drives the same plumbing with NCCL."""
import os
import socket
import subprocess
import sys
import textwrap
from pathlib import Path

import numpy as np
import pytest

from libsplinter_b200.sharding import shard_range

ROOT = Path(__file__).resolve().parent.parent

WORKER = textwrap.dedent("""
    import os, sys
    import numpy as np, torch, torch.distributed as dist
    sys.path.insert(0, {root!r}); sys.path.insert(0, {root!r} + "/tests")
    import support
    from libsplinter_b200.sharding import shard_range, gather_lists
    from libsplinter_b200.binding import RESULT_DTYPE
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    N, DIM, K = 10_000, 256, 25
    lo, cnt = shard_range(N, world, rank)
    E, bloom, flags = support.synth_dense(1234, 99, lo, cnt, DIM)      # this rank's shard only
    q = support.synth_queries(4321, 3, DIM)
    ok = True
    for order, mask in ((0, 0), (0, 3), (1, 1)):
        ref_local, rc = support.oracle_search_dense(q, E, bloom, flags, k=K, order=order, bloom_mask=mask)
        for qi in range(3):
            rec = np.zeros(K, RESULT_DTYPE)
            n = int(rc[qi])
            rec["slot_idx"][:n] = ref_local[qi]["idx"][:n] + lo            # GLOBAL slot indices
            rec["similarity"][:n] = ref_local[qi]["similarity"][:n]
            rec["distance"][:n] = ref_local[qi]["distance"][:n]
            rec["has_embedding"][:n] = 1
            t = torch.from_numpy(rec.view(np.uint8).reshape(K, 32).copy())
            g, c = gather_lists(t, torch.tensor([n], dtype=torch.int32), world)
            lists = g.numpy().reshape(world * K, 32).view(RESULT_DTYPE).reshape(world, K)
            # merge on every rank with the oracle's k-way merge
            flat = np.zeros((world, K), support.SPLO_RESULT)
            for w in range(world):
                flat[w]["idx"] = lists[w]["slot_idx"]; flat[w]["similarity"] = lists[w]["similarity"]
                flat[w]["distance"] = lists[w]["distance"]
            out = np.zeros(K, support.SPLO_RESULT)
            counts = c.numpy().astype(np.uint32)
            m = support.oracle().splo_merge_topk(support._p(flat), support._p(counts), world, K, order, K, support._p(out))
            # the single-store answer
            Eg, bg, fg = support.synth_dense(1234, 99, 0, N, DIM)
            full, fc = support.oracle_search_dense(q[qi], Eg, bg, fg, k=K, order=order, bloom_mask=mask)
            ok &= (m == fc[0]) and np.array_equal(out[:m]["idx"], full[0][:m]["idx"]) \\
                and np.array_equal(out[:m]["similarity"].view(np.uint32), full[0][:m]["similarity"].view(np.uint32))
    # the batched form: every rank contributes a [nq][K] block, gathered as [rank][query][K]
    from libsplinter_b200.sharding import gather_blocks
    ref_local, rc = support.oracle_search_dense(q, E, bloom, flags, k=K, bloom_mask=1)
    blk = np.zeros((3, K), RESULT_DTYPE)
    for qi in range(3):
        n = int(rc[qi])
        blk[qi]["slot_idx"][:n] = ref_local[qi]["idx"][:n] + lo
        blk[qi]["similarity"][:n] = ref_local[qi]["similarity"][:n]
        blk[qi]["distance"][:n] = ref_local[qi]["distance"][:n]
    gb, gc = gather_blocks(torch.from_numpy(blk.view(np.uint8).reshape(3 * K, 32).copy()),
                           torch.from_numpy(rc.astype(np.int32)), world)
    gb = gb.numpy().view(RESULT_DTYPE).reshape(world, 3, K)
    gc = gc.numpy().reshape(world, 3)
    Eg, bg, fg = support.synth_dense(1234, 99, 0, N, DIM)
    full, fc = support.oracle_search_dense(q, Eg, bg, fg, k=K, bloom_mask=1)
    for qi in range(3):
        flat = np.zeros((world, K), support.SPLO_RESULT)
        for w in range(world):
            flat[w]["idx"] = gb[w, qi]["slot_idx"]; flat[w]["similarity"] = gb[w, qi]["similarity"]
            flat[w]["distance"] = gb[w, qi]["distance"]
        out = np.zeros(K, support.SPLO_RESULT)
        m = support.oracle().splo_merge_topk(support._p(flat), support._p(gc[:, qi].astype(np.uint32).copy()), world, K, 0, K,
                                             support._p(out))
        ok &= (m == fc[qi]) and np.array_equal(out[:m]["idx"], full[qi][:m]["idx"])
    # ---- ShardedSearcher: packed blocks, ONE gather per batch, and the re-issue of a query a shard could not answer ----
    import ctypes as C
    from libsplinter_b200.sharding import ShardedSearcher
    from libsplinter_b200.binding import (COUNT_UNANSWERED, MODE_BATCH_BF16, MODE_FAST, SearchParams, packed_block_bytes,
                                          packed_counts_offset)
    calls = dict(search=0, gathers=0)
    real_gather = dist.all_gather_into_tensor
    def counting_gather(*a, **kw):
        calls["gathers"] += 1
        return real_gather(*a, **kw)
    dist.all_gather_into_tensor = counting_gather
    def search_fn(params, qt, block):
        nq = int(params.nq)
        calls["search"] += 1
        loc, lc = support.oracle_search_dense(qt.numpy(), E, bloom, flags, k=K, bloom_mask=int(params.bloom_mask))
        raw = block.numpy()
        raw[:] = 0
        recs = raw[:nq * K * 32].view(RESULT_DTYPE).reshape(nq, K)
        for i in range(nq):
            n = int(lc[i])
            recs[i]["slot_idx"][:n] = loc[i]["idx"][:n] + lo
            recs[i]["similarity"][:n] = loc[i]["similarity"][:n]
            recs[i]["distance"][:n] = loc[i]["distance"][:n]
        cnts = raw[packed_counts_offset(nq, K):packed_counts_offset(nq, K) + 4 * nq].view(np.uint32)
        cnts[:] = lc
        if int(params.mode) != MODE_FAST and rank == world - 1:
            cnts[1] = COUNT_UNANSWERED          # pretend this shard's bf16 pass could not prove query 1
    def merge_fn(gathered, nq, fin, fcnt):
        blk_bytes = packed_block_bytes(nq, K)
        g = gathered.numpy().reshape(world, blk_bytes)
        outr = fin.numpy().reshape(nq, K * 32).view(RESULT_DTYPE).reshape(nq, K)
        for i in range(nq):
            cs = np.array([g[w][packed_counts_offset(nq, K):].view(np.uint32)[i] for w in range(world)], np.uint32)
            if (cs == COUNT_UNANSWERED).any():
                fcnt[i] = -1                    # the marker, as the int32 the tensor holds
                continue
            flat = np.zeros((world, K), support.SPLO_RESULT)
            for w in range(world):
                r = g[w][:nq * K * 32].view(RESULT_DTYPE).reshape(nq, K)[i]
                flat[w]["idx"] = r["slot_idx"]; flat[w]["similarity"] = r["similarity"]; flat[w]["distance"] = r["distance"]
            mo = np.zeros(K, support.SPLO_RESULT)
            mm = support.oracle().splo_merge_topk(support._p(flat), support._p(cs), world, K, 0, K, support._p(mo))
            outr[i]["slot_idx"][:mm] = mo["idx"][:mm]; outr[i]["similarity"][:mm] = mo["similarity"][:mm]
            outr[i]["distance"][:mm] = mo["distance"][:mm]
            fcnt[i] = int(mm)
    S = ShardedSearcher(None, world, 3, K, 0, "cpu", search_fn=search_fn, merge_fn=merge_fn)
    prm = SearchParams(C.sizeof(SearchParams), 3, DIM, K, 0, MODE_BATCH_BF16, 1, 0.0, 0.0, 0xFFFFFFFF, 0)
    recs, cnts = S.search(prm, torch.from_numpy(q))
    for qi in range(3):
        ok &= int(cnts[qi]) == int(fc[qi]) and np.array_equal(recs[qi][:cnts[qi]]["slot_idx"], full[qi][:fc[qi]]["idx"])
    ok &= S.reissued == 1 and calls["search"] == 2        # one batched pass + one fp32 re-issue of query 1, on every rank
    ok &= calls["gathers"] == (2 if world > 1 else 0)     # ONE collective per pass
    flag = torch.tensor([1 if ok else 0])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("SHARDING_OK" if int(flag) == 1 else "SHARDING_MISMATCH")
    dist.destroy_process_group()
""")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_search_merges_to_the_single_store_answer(world, tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=str(ROOT)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), str(script)],
                       capture_output=True, text=True, timeout=300, env={**os.environ, "OMP_NUM_THREADS": "1"})
    assert r.returncode == 0, r.stderr[-2000:]
    assert "SHARDING_OK" in r.stdout, r.stdout[-500:] + r.stderr[-500:]


def test_shard_ranges_cover_the_slot_space_exactly():
    for total in (1, 7, 4096, 1 << 20, (1 << 26) + 5):
        for world in (1, 2, 3, 4, 8):
            seen = 0
            for r in range(world):
                lo, cnt = shard_range(total, world, r)
                assert lo == min(seen, total) or cnt == 0
                seen += cnt
            assert seen == total
