"""The multi-GPU path on real GPUs: 2 ranks over NCCL (skipped on a one-GPU box; the gloo tests cover the plumbing
on CPU).  Each rank mirrors its slot range, searches it through the C ABI on its own stream, the packed result blocks
cross NVLink in ONE all-gather per batch, every rank merges on its device -- and the merged answer must be the CPU
oracle's over the whole corpus: single-query fp32 scans (k = 100, config C4's shape), a tensor-core batch, and a batch
with a crowd of near-duplicates, where the shards hand up SPL_GPU_COUNT_UNANSWERED and all ranks re-issue the same
queries on the fp32 scan without talking to each other."""
import os
import socket
import subprocess
import sys
import textwrap
from pathlib import Path

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent

WORKER = textwrap.dedent("""
    import os, sys
    import numpy as np, torch, torch.distributed as dist
    sys.path.insert(0, {root!r}); sys.path.insert(0, {root!r} + "/tests")
    import libsplinter_b200 as sg, support
    from libsplinter_b200.sharding import ShardedSearcher, shard_range
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    N, DIM = 120_000, 768
    E, bloom, flags = support.synth_dense(1234, 99, 0, N, DIM)
    rng = np.random.default_rng(5)
    crowd = rng.choice(np.arange(1, N), 5000, replace=False)          # near-copies of row 0, spread over both shards
    E[crowd] = E[0][None, :] + (1e-3 / np.sqrt(DIM)) * rng.standard_normal((5000, DIM)).astype(np.float32)
    flags[crowd] = 7; flags[0] = 7
    lo, cnt = shard_range(N, world, rank)
    m = sg.Mirror.open_dense(cnt, DIM, global_row0=lo, device=local, flags=sg.OPT_BF16_MIRROR)
    m.dense_upload(0, E[lo:lo + cnt], bloom[lo:lo + cnt], flags[lo:lo + cnt])
    st = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(st)
    ok = True
    def check(recs, counts, q, k, **kw):
        ref, rc = support.oracle_search_dense(q, E, bloom, flags, k=k, order=kw.get("order", 0), bloom_mask=kw.get("bloom_mask", 0))
        for i in range(q.shape[0]):
            assert counts[i] == rc[i], (counts[i], rc[i])
            support.assert_topk_matches(recs[i][:counts[i]], ref[i][:rc[i]], order=kw.get("order", 0))
    # 1. single-query fp32 scans, top-100 (C4's shape)
    q = support.synth_queries(4321, 3, DIM)
    S = ShardedSearcher(m, world, 1, 100, sg.ORDER_COSINE, dev, st.cuda_stream)
    for i in range(3):
        for mask in (0, 0x1):
            recs, counts = S.search(m.params(1, 100, mode=sg.MODE_FAST, bloom_mask=mask), torch.from_numpy(q[i:i + 1]).to(dev))
            check(recs, counts, q[i:i + 1], 100, bloom_mask=mask)
    # 2. a tensor-core batch, far from the crowd: answered by the bf16 path on both shards
    qb = support.synth_queries(77, 16, DIM)
    SB = ShardedSearcher(m, world, 16, 10, sg.ORDER_COSINE, dev, st.cuda_stream)
    recs, counts = SB.search(m.params(16, 10, mode=sg.MODE_BATCH_BF16), torch.from_numpy(qb).to(dev))
    check(recs, counts, qb, 10)
    assert SB.reissued == 0
    # 3. queries inside the crowd: the shards cannot prove their bf16 selection, the merge hands the marker on,
    #    every rank re-issues the same queries on the fp32 scan
    qc = qb.copy()
    qc[:4] = E[0][None, :] + (0.2 / np.sqrt(DIM)) * rng.standard_normal((4, DIM)).astype(np.float32)
    recs, counts = SB.search(m.params(16, 10, mode=sg.MODE_BATCH_BF16), torch.from_numpy(qc).to(dev))
    check(recs, counts, qc, 10)
    assert SB.reissued >= 4, SB.reissued
    t = torch.tensor([SB.reissued], device=dev)
    lst = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(lst, t)
    assert len({{int(x) for x in lst}}) == 1            # the same decision everywhere, with no extra communication
    m.close()
    dist.barrier()
    if rank == 0:
        print("NCCL_SHARDING_OK")
    dist.destroy_process_group()
""")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (run with gpurun --gpus 2)")
def test_two_ranks_over_nccl_give_the_single_store_answer(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=str(ROOT)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), str(script)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    assert "NCCL_SHARDING_OK" in r.stdout, r.stdout[-500:] + r.stderr[-1500:]
