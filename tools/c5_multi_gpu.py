"""C5 across GPUs: one reference-format store in /dev/shm, every rank mirrors its own slot range
(fp32 + bf16), writer processes mutate the store through the reference API with C5's mix, every rank keeps syncing
its range, and batched searches (nq queries, cosine top-k, tensor-core path) run across all ranks: shard search ->
NCCL all-gather of the [nq][k] blocks -> one batched merge launch.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
      tools/c5_multi_gpu.py [SLOTS] [FILL] [SECONDS] [SYNC_PERIOD_MS]
"""
import json
import mmap
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
import torch.distributed as dist

import libsplinter_b200 as sg
import support
from libsplinter_b200.sharding import gather_blocks, shard_range

slots = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
fill = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
seconds = float(sys.argv[3]) if len(sys.argv) > 3 else 4.0
period = (float(sys.argv[4]) if len(sys.argv) > 4 else 100.0) / 1e3
DIM, MAXVAL, NQ, K = 768, 64, 64, 10

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local_rank = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local_rank)
dev = torch.device("cuda", local_rank)
dist.init_process_group("nccl", device_id=dev)
name = f"/splb200_c5_{os.environ.get('MASTER_PORT', '0')}"
path = "/dev/shm" + name
size = support.oracle().splo_image_size(slots, MAXVAL, DIM)


def avail_bytes():
    for ln in open("/proc/meminfo"):
        if ln.startswith("MemAvailable"):
            return int(ln.split()[1]) * 1024
    return 0


out = {"gpus": world, "slots": slots, "fill": fill, "store_bytes": size, "nq": NQ, "k": K}
img = None
if rank == 0:
    assert avail_bytes() > size * 1.3, f"host has {avail_bytes() >> 30} GiB available, the store needs {size >> 30} GiB"
    if os.path.exists(path):
        os.unlink(path)
    with open(path, "wb") as f:
        f.truncate(size)
    fd = os.open(path, os.O_RDWR)
    mm = mmap.mmap(fd, size)
    img = np.frombuffer(mm, np.uint8)
    t0 = time.perf_counter()
    assert support.oracle().splo_image_init(support._p(img), slots, MAXVAL, DIM) == 0
    placed = support.image_fill_synthetic(img, int(slots * fill))
    out["keys_placed"] = int(placed)
    out["host_build_s"] = round(time.perf_counter() - t0, 2)
dist.barrier()

try:
    lo, cnt = shard_range(slots, world, rank)
    cores = os.cpu_count() or 8
    t0 = time.perf_counter()
    m = sg.Mirror.attach(name, slot_begin=lo, slot_count=cnt, flags=sg.OPT_BF16_MIRROR, device=local_rank,
                         sync_threads=max(2, cores // world))
    attach_s = time.perf_counter() - t0
    dist.barrier()

    tstream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    q_host = support.synth_queries(777, NQ, DIM)
    dq = torch.from_numpy(q_host).to(dev)
    blk = torch.zeros((NQ * K, 32), dtype=torch.uint8, device=dev)
    bcnt = torch.zeros(NQ, dtype=torch.int32, device=dev)
    g_blk = torch.zeros((world * NQ * K, 32), dtype=torch.uint8, device=dev)
    g_cnt = torch.zeros(world * NQ, dtype=torch.int32, device=dev)
    fin = torch.zeros((NQ, K, 32), dtype=torch.uint8, device=dev)
    fcnt = torch.zeros(NQ, dtype=torch.int32, device=dev)
    p = m.params(NQ, K, mode=sg.MODE_BATCH_BF16, order=sg.ORDER_COSINE)

    def batched_search():
        m.search_device(p, dq.data_ptr(), blk.data_ptr(), K, bcnt.data_ptr(), stream)
        gather_blocks(blk, bcnt, world, g_blk, g_cnt)
        m.merge_topk_batch_device(g_blk.data_ptr(), g_cnt.data_ptr(), world, NQ, K, sg.ORDER_COSINE, K,
                                  fin.data_ptr(), K, fcnt.data_ptr(), stream)
        tstream.synchronize()

    for _ in range(3):
        batched_search()
    dist.barrier()

    syncs, stop = [], threading.Event()

    def syncer():
        while not stop.is_set():
            m.wait(100)
            t0 = time.perf_counter()
            s = m.sync()
            dt = time.perf_counter() - t0
            syncs.append((dt, s.rows_uploaded, s.skipped_busy))
            if period > dt:
                time.sleep(period - dt)

    procs = []
    if rank == 0:
        writer = str(support.REF_DIR / "ref_writer")
        nkeys, lanes = int(slots * fill), 4
        procs = [subprocess.Popen([writer, name, str(i * (nkeys // lanes)), str((i + 1) * (nkeys // lanes)), str(77 + i),
                                   str(int(seconds * 1e3))], stdout=subprocess.PIPE, text=True) for i in range(lanes)]
    th = threading.Thread(target=syncer)
    th.start()
    flag = torch.ones(1, dtype=torch.int32, device=dev)
    batches, t_search, t_begin = 0, 0.0, time.perf_counter()
    while True:
        if rank == 0:
            flag.fill_(1 if any(pr.poll() is None for pr in procs) else 0)
        dist.broadcast(flag, 0)
        if int(flag.item()) == 0:
            break
        t0 = time.perf_counter()
        batched_search()
        t_search += time.perf_counter() - t0
        batches += 1
    wall = time.perf_counter() - t_begin
    stop.set()
    th.join()
    ops = sum(int(pr.communicate()[0].split()[0]) for pr in procs) if rank == 0 else 0
    m.sync()                      # quiesced: one more sweep and every shard equals its range of the store
    dist.barrier()
    batched_search()

    mine = {"rank": rank, "attach_s": round(attach_s, 3), "syncs": len(syncs),
            "sync_ms_median": round(float(np.median([x[0] for x in syncs])) * 1e3, 2) if syncs else None,
            "rows_uploaded": int(sum(x[1] for x in syncs)), "skipped_busy": int(sum(x[2] for x in syncs))}
    allr = [None] * world
    dist.all_gather_object(allr, mine)
    if rank == 0:
        res = fin.cpu().numpy().view(sg.RESULT_DTYPE).reshape(NQ, K)
        rc = fcnt.cpu().numpy()
        for i in (0, 17, 63):     # merged cross-GPU answer == the CPU oracle over the whole store image
            ref, _ = support.oracle_search_image(img, q_host[i], K)
            assert rc[i] == len(ref)
            support.assert_topk_matches(res[i][:rc[i]], ref)
        out.update({"parity_after_quiesce": True, "writer_processes": 4, "writer_ops_per_s": round(ops / seconds),
                    "writer_mix": "80% set_embedding, 10% retrain_slot, 5% set/unset_label, 5% unset+set",
                    "sync_period_ms": period * 1e3, "batched_searches": batches,
                    "ms_per_batch": round(t_search / max(batches, 1) * 1e3, 3),
                    "queries_per_s": round(NQ * batches / max(t_search, 1e-9)),
                    "queries_per_s_wall": round(NQ * batches / wall), "per_rank": allr})
        print(json.dumps(out))
    m.close()
finally:
    dist.barrier()
    if rank == 0 and os.path.exists(path):
        os.unlink(path)
    dist.destroy_process_group()
