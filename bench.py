#!/usr/bin/env python
"""bench.py -- headline benchmark of the similarity-search hot path (BASELINE.json).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the hot path over one query: label predicate + fused cosine/L2 scoring
+ top-k over the whole (sharded) slot mirror.

  N = 1  workload C2: 2^20 slots x 768 fp32, single-query cosine top-10, bloom mask 0x3
         (the all-matching mask and the unfiltered scan ride along as `variants`).
  N > 1  workload C4: 2^26 slots x 768 fp32 sharded by slot range over N GPUs, single-query
         cosine top-100, per-GPU top-k all-gathered over NCCL and merged on the device.

metric = scanned GB/s, whole job (algorithmic bytes of SURVEY 8(d): 9 B per slot + 4*D per slot
that passes the label predicate + the query), which is the BASELINE.json quantity that is
comparable across shard sizes; queries/s is printed beside it.  `value` is measured with the
query and the result buffers resident in HBM (CUDA events on the launching stream); `e2e` is
the same metric through the host-buffer C-ABI call (pinned H2D of the query + D2H of the hits
inside the timed region).  `--impl reference` times the reference's own CPU code
(oracle/_ref: its cosine_similarity / euclidean_distance / compare_results + qsort, split
over all host cores) on the same config.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

DIM = 768
SEED_CORPUS, SEED_BLOOM, SEED_QUERY = 1234, 99, 4321
C2_ROWS, C2_K, C2_MASK = 1 << 20, 10, 0x3
C4_ROWS, C4_K = 1 << 26, 100
N_QUERIES = 64


def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.FIELDS}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.t.start()
        except OSError:
            self.proc = None
        return self

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            self.t.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6 or not f[0].isdigit():
                continue
            sm.append(int(f[0])); mx.append(int(f[1]))
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": int(statistics.median(sm)), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def algorithmic_bytes(n_rows, n_pass):
    """8 (bloom) + 1 (flags) per slot, 4*D per slot passing the predicate, + the query."""
    return 9 * n_rows + 4 * DIM * n_pass + 4 * DIM


# ------------------------------------------------------------------------------------------ reference arm
def run_reference(args):
    import support
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    if not support.have_ref():
        # the reference did compile in the build container; if oracle/_ref did not travel, fall back to the port
        kind = "port"
    else:
        kind = "reference"
    cores = os.cpu_count() or 1
    single = args.gpus == 1
    k, mask = (C2_K, C2_MASK) if single else (C4_K, 0)
    n = C2_ROWS  # C2 in full; for C4 a 2^20-row sample of the 2^26 (GB/s is size-independent)
    E, bloom, flags = support.synth_dense(SEED_CORPUS, SEED_BLOOM, 0, n, DIM)
    q = support.synth_queries(SEED_QUERY, N_QUERIES, DIM)
    need = support.F_ENUM if mask else support.F_LIST
    cand = (((flags & need) != 0) & ((bloom & np.uint64(mask)) == np.uint64(mask))).astype(np.uint8)
    n_pass = int(cand.sum())
    nbytes = algorithmic_bytes(n, n_pass)

    def step(i, threads):
        qi = q[i % N_QUERIES]
        if kind == "reference":
            oi, os_, od = np.zeros(k, np.uint32), np.zeros(k, np.float32), np.zeros(k, np.float32)
            support.ref_search().ref_topk_dense(support._p(qi), support._p(E), n, DIM, support._p(cand), 0.0, 0.0, k,
                                                threads, support._p(oi), support._p(os_), support._p(od))
        else:
            support.oracle_search_dense(qi, E, bloom, flags, k=k, bloom_mask=mask, threads=threads)

    for i in range(args.warmup):
        step(i, cores)
    t0 = time.perf_counter()
    for i in range(args.steps):
        step(i, cores)
    dt = time.perf_counter() - t0
    gbs = nbytes * args.steps / dt / 1e9
    qps = args.steps / dt * (n / (C2_ROWS if single else C4_ROWS))
    sample = ("every query over the full 2^20-slot corpus" if single else
              "2^20-slot sample of the 2^26-slot corpus; queries/s scaled by 2^20/2^26, GB/s is size-independent")
    line = {
        "impl": "reference", "metric": "scanned_GB_per_s", "value": round(gbs, 3), "unit": "GB/s",
        "queries_per_s": round(qps, 4), "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(dt / args.steps * 1e3, 3), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.gpus),
        "cpu_baseline": {"value": round(gbs, 3), "unit": "GB/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": round(gbs, 3), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def workload_config(n_gpus):
    if n_gpus == 1:
        return {"workload": "C2: 2^20 slots x 768-d fp32, single-query cosine top-10, bloom-label pre-filter mask 0x3 "
                            "(12.2% of slots pass); mirror resident in HBM (3.2 GB > 126 MB L2, no flush needed)",
                "slots": C2_ROWS, "dim": DIM, "k": C2_K, "bloom_mask": "0x3", "parallelism": "1 GPU",
                "l2_policy": "inputs larger than L2",
                "pipelining": "`value`: a query loop on one stream, searches flagged SPL_SEARCH_PIPELINED (a scan may start while "
                              "the previous one merges its per-SM lists; results in stream order); `latency_us` and `e2e`: one "
                              "search at a time, nothing overlapped"}
    return {"workload": f"C4: 2^26 slots x 768-d fp32 sharded by slot range over {n_gpus} GPUs, single-query cosine "
                        "top-100, per-GPU top-k all-gathered over NCCL + device merge",
            "slots": C4_ROWS, "dim": DIM, "k": C4_K, "bloom_mask": "0", "parallelism": f"slot-range shards x{n_gpus}",
            "l2_policy": "inputs larger than L2",
            "note": "the total (2^26 slots, 206 GB) is fixed for N = 2, 4, 8 (strong scaling among them); it does not fit one "
                    "GPU, so N = 1 runs C2 (2^20 slots) -- compare GB/s per GPU, not queries/s, across that step"}


# ------------------------------------------------------------------------------------------ our arm
def cpu_baseline_leg(k, mask, seconds=12.0):
    """The reference's own CPU search (oracle/_ref) on the box's host cores, bounded sample."""
    import support
    cores = os.cpu_count() or 1
    n = 1 << 18  # 2^18-slot sample (805 MB): generation stays in seconds; GB/s is size-independent
    E, bloom, flags = support.synth_dense(SEED_CORPUS, SEED_BLOOM, 0, n, DIM)
    q = support.synth_queries(SEED_QUERY, 8, DIM)
    need = support.F_ENUM if mask else support.F_LIST
    cand = (((flags & need) != 0) & ((bloom & np.uint64(mask)) == np.uint64(mask))).astype(np.uint8)
    nbytes = algorithmic_bytes(n, int(cand.sum()))
    kind = "reference" if support.have_ref() else "port"
    out = {}
    for label, threads in (("all_cores", cores), ("one_thread", 1)):
        t0 = time.perf_counter()
        steps = 0
        while True:
            qi = q[steps % 8]
            if kind == "reference":
                oi, os_, od = np.zeros(k, np.uint32), np.zeros(k, np.float32), np.zeros(k, np.float32)
                support.ref_search().ref_topk_dense(support._p(qi), support._p(E), n, DIM, support._p(cand), 0.0, 0.0,
                                                    k, threads, support._p(oi), support._p(os_), support._p(od))
            else:
                support.oracle_search_dense(qi, E, bloom, flags, k=k, bloom_mask=mask, threads=threads)
            steps += 1
            dt = time.perf_counter() - t0
            if dt > seconds / 2 and steps >= 3:
                break
        out[label] = (nbytes * steps / dt / 1e9, steps / dt * n)
    return {
        "value": round(out["all_cores"][0], 3), "unit": "GB/s", "cores": cores, "kind": kind,
        "sample": f"2^18-slot sample of the workload, ~{seconds / 2:.0f} s per leg; the reference's scorers + qsort "
                  "split over all host cores (the shipped reference is single-threaded)",
        "slots_per_s": round(out["all_cores"][1]),
        "one_thread": {"value": round(out["one_thread"][0], 3), "unit": "GB/s", "slots_per_s": round(out["one_thread"][1])},
    }


def run_ours(args):
    import torch
    import torch.distributed as dist

    import libsplinter_b200 as sg
    import support

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch multi-GPU runs with torch.distributed.run (one rank per GPU)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    # the contract is ONE JSON line on stdout: whatever libraries print while they initialise (NCCL announces its
    # version there) goes to stderr; stdout comes back just before the line is printed
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    single = world == 1
    total_rows = C2_ROWS if single else C4_ROWS
    k = C2_K if single else C4_K
    from libsplinter_b200.sharding import shard_range
    row0, rows = shard_range(total_rows, world, rank)
    peak, peak_src = measured_peaks()

    m = sg.Mirror.open_dense(rows, DIM, global_row0=row0, device=local_rank)
    m.fill_synthetic(SEED_CORPUS, SEED_BLOOM)
    _, bloom_h, flags_h, _ = m.read_mirror(row0, rows, want_E=False)

    q_host = support.synth_queries(SEED_QUERY, N_QUERIES, DIM)
    q_dev = torch.from_numpy(q_host).to(dev)
    out_dev = torch.zeros((k, 32), dtype=torch.uint8, device=dev)
    cnt_dev = torch.zeros(1, dtype=torch.int32, device=dev)
    q_pin = torch.from_numpy(q_host).pin_memory()
    res_pin = torch.zeros((k, 32), dtype=torch.uint8).pin_memory()
    # A dedicated (non-default) stream: the kernels, the NCCL ops and the timing events must all sit on
    # the SAME stream -- torch.cuda.Event only sees torch's current stream.
    tstream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    assert stream != 0
    from libsplinter_b200.sharding import ShardedSearcher
    # N > 1: shard scan -> ONE all-gather of the packed {records, count} block per query -> merge on every rank
    sharded = ShardedSearcher(m, world, 1, k, sg.ORDER_COSINE, dev, stream) if world > 1 else None
    fin_dev = sharded.fin if sharded else None

    def n_pass_of(mask):
        need = sg.ROW_ENUM if mask else sg.ROW_LIST
        return int((((flags_h & need) != 0) & ((bloom_h & np.uint64(mask)) == np.uint64(mask))).sum())

    def device_step(i, mask, qsrc=None):
        # a query loop: each search directly follows another search of this context on this stream
        p = m.params(1, k, mode=sg.MODE_AUTO, bloom_mask=mask, flags=sg.SEARCH_PIPELINED)
        qv = qsrc if qsrc is not None else q_dev[i % N_QUERIES]
        if world > 1:
            sharded.enqueue(p, qv)      # (fp32 scan: always answers, no re-issue to wait for)
        else:
            m.search_device(p, qv.data_ptr(), out_dev.data_ptr(), k, cnt_dev.data_ptr(), stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_device(mask, steps, warmup):
        for i in range(warmup):
            device_step(i, mask)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = m.launch_count()
        e0.record()
        for i in range(steps):
            device_step(i, mask)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        launches = m.launch_count() - l0
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms / steps, launches

    def timed_e2e(mask, steps, warmup):
        """host query in pinned memory -> H2D, scan (+ gather/merge), D2H of the hits, every step"""
        qbuf = torch.empty(DIM, dtype=torch.float32, device=dev)
        hp = m.params(1, k, mode=sg.MODE_AUTO, bloom_mask=mask)
        h_out, h_cnt = np.zeros((1, k), sg.RESULT_DTYPE), np.zeros(1, np.uint32)
        # what a C host holds: plain addresses (numpy's .ctypes accessors cost about a microsecond each)
        import ctypes
        hp_ref, q_addr = ctypes.byref(hp), [q_host[i].ctypes.data for i in range(N_QUERIES)]
        out_addr, cnt_addr = h_out.ctypes.data, h_cnt.ctypes.data

        def one(i):
            if single:
                # the host-buffer C-ABI call itself: query in host memory in, records in host memory out
                m.search_ptrs(hp_ref, q_addr[i % N_QUERIES], out_addr, k, cnt_addr)
            else:
                qbuf.copy_(q_pin[i % N_QUERIES], non_blocking=True)
                device_step(i, mask, qsrc=qbuf)
                res_pin.copy_(fin_dev.view(k, 32), non_blocking=True)
                torch.cuda.current_stream().synchronize()
        for i in range(warmup):
            one(i)
        barrier()
        t0 = time.perf_counter()
        for i in range(steps):
            one(i)
        barrier()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        return dt / steps * 1e3

    head_mask = C2_MASK if single else 0
    n_pass_local = n_pass_of(head_mask)
    if world > 1:
        t = torch.tensor([n_pass_local], device=dev, dtype=torch.int64)
        dist.all_reduce(t)
        n_pass_total = int(t.item())
    else:
        n_pass_total = n_pass_local
    job_bytes = algorithmic_bytes(total_rows, n_pass_total)

    with ClockSampler(local_rank) as clk:
        time.sleep(0.25)  # nvidia-smi needs a moment before its first sample; keep it out of the timed region
        ms_step, launches = timed_device(head_mask, args.steps, args.warmup)
        if ms_step * args.steps < 300.0:  # short runs: keep the GPU under the same load until a few samples exist
            timed_device(head_mask, max(10, int(300.0 / max(ms_step, 1e-3))), 0)
    clocks = clk.summary()
    ms_e2e = timed_e2e(head_mask, args.steps, args.warmup)

    # per-query latency distribution (median + p99): one event between consecutive launches
    latency = None
    if single:
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
        barrier()
        evs[0].record()
        for i in range(args.steps):
            device_step(i, head_mask)
            evs[i + 1].record()
        barrier()
        per = np.array([evs[i].elapsed_time(evs[i + 1]) * 1e3 for i in range(args.steps)])
        latency = {"median": round(float(np.median(per)), 2), "p99": round(float(np.percentile(per, 99)), 2),
                   "max": round(float(per.max()), 2), "n": int(args.steps)}

    variants = {}
    if single:
        for name, mask in (("mask_all_match_bit63", 1 << 63), ("unfiltered_mask_0", 0)):
            ms_v, _ = timed_device(mask, args.steps, args.warmup)
            b = algorithmic_bytes(total_rows, n_pass_of(mask))
            variants[name] = {"ms_per_step": round(ms_v, 4), "queries_per_s": round(1e3 / ms_v, 2),
                              "scanned_GB_per_s": round(b / ms_v / 1e6, 1), "roofline_frac": round(b / ms_v / 1e6 / peak, 4),
                              "algorithmic_bytes": b}

    # ---- small batch: 4 queries per call share one pass over the store (fp32 scan kernel, NQ = 4) ----
    small, small8 = None, None
    if single:
        small, small8 = {}, {}
        for nqs, dest in ((4, small), (8, small8)):
            q4 = torch.stack([q_dev[i] for i in range(nqs)]).contiguous()
            out4 = torch.zeros((nqs * k, 32), dtype=torch.uint8, device=dev)
            cnt4 = torch.zeros(nqs, dtype=torch.int32, device=dev)
            for name, mask in (("mask_0x3", C2_MASK), ("mask_all_match_bit63", 1 << 63)):
                p4 = m.params(nqs, k, mode=sg.MODE_FAST, bloom_mask=mask, flags=sg.SEARCH_PIPELINED)   # a loop of calls on one stream
                for _ in range(args.warmup):
                    m.search_device(p4, q4.data_ptr(), out4.data_ptr(), k, cnt4.data_ptr(), stream)
                barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(args.steps):
                    m.search_device(p4, q4.data_ptr(), out4.data_ptr(), k, cnt4.data_ptr(), stream)
                e1.record()
                barrier()
                ms4 = e0.elapsed_time(e1) / args.steps
                b = algorithmic_bytes(total_rows, n_pass_of(mask))
                dest[name] = {"queries_per_call": nqs, "ms_per_call": round(ms4, 4), "queries_per_s": round(nqs * 1e3 / ms4, 1),
                              "hbm_GB_per_s": round(b / ms4 / 1e6, 1), "speedup_vs_one_at_a_time": None}

    # ---- the reference's DEFAULT invocation: no --limit, every survivor sorted and returned (_search.c:412-416) ----
    unlimited = None
    if single and not args.skip_extra:
        unlimited = {}
        m.set_timing(True)
        # caller-owned, reused host buffers, as a C host has them (a fresh 33 MB numpy array per call would add its page
        # faults -- 3x the call itself -- to the number)
        big_out = np.zeros(rows * 32, np.uint8)
        big_out[:] = 1
        big_cnt = np.zeros(1, np.uint32)
        for name, mask in (("mask_0x3", C2_MASK), ("unfiltered_mask_0", 0)):
            p0 = m.params(1, 0, mode=sg.MODE_AUTO, bloom_mask=mask)
            for _ in range(2):
                m.search_into(p0, q_host[0], big_out, rows, big_cnt)
            scans, walls, n_hits = [], [], 0
            for i in range(7):
                t0 = time.perf_counter()
                m.search_into(p0, q_host[i % len(q_host)], big_out, rows, big_cnt)
                walls.append((time.perf_counter() - t0) * 1e3)
                scans.append(m.last_scan_ms())
                n_hits = int(big_cnt[0])
            b = algorithmic_bytes(total_rows, n_pass_of(mask))
            scan_ms, wall_ms = float(np.median(scans)), float(np.median(walls))
            unlimited[name] = {"results": n_hits, "scan_ms": round(scan_ms, 4), "scan_GB_per_s": round(b / scan_ms / 1e6, 1),
                               "scan_roofline_frac": round(b / scan_ms / 1e6 / peak, 4), "e2e_ms": round(wall_ms, 3),
                               "d2h_bytes": 32 * n_hits + 4,
                               "note": "e2e = spl_gpu_search with reused pageable host buffers: H2D query + key-per-row scan + compaction + sort of the survivors + D2H of every record"}
        m.set_timing(False)

    # ---- C3: batched 256-query L2 top-32 on the tensor cores (rides along at N = 1) ----
    c3 = c3_batched_leg(sg, support, torch, dev, stream, args) if single else None

    # ---- parity gate on the timed configuration (rank-local shard sample vs oracle) ----
    parity = parity_gate(m, sg, support, rows, row0, k, head_mask)
    if world > 1:   # ... and on the MERGED answer every rank ends up with
        parity.update(merged_parity_gate(sharded, m, sg, support, torch, dist, q_host, q_dev, k, head_mask, world, rank, dev))

    # ---- same-workload single-GPU anchor for the C4 scaling curve (N = 1 only): the largest shard that fits ----
    anchor = c4_anchor_leg(sg, support, torch, dev, stream, peak, local_rank) if single and not args.skip_extra else None
    # ---- C5: mixed workload across all ranks (N = 8 by default; SPL_BENCH_C5_SLOTS forces it at any N > 1) ----
    c5 = None
    c5_slots = int(os.environ.get("SPL_BENCH_C5_SLOTS", str(1 << 24) if world == 8 else "0"))
    if world > 1 and c5_slots and not args.skip_extra:
        c5 = c5_mixed_leg(sg, support, torch, dist, c5_slots, world, rank, local_rank, dev, tstream)

    # ---- the same C4 query through the one-process group call of the C ABI (N > 1: rank 0 drives all GPUs) ----
    group_leg = None
    if world > 1 and not args.skip_extra:
        m.close()
        m = None
        torch.cuda.empty_cache()
        dist.barrier()
        torch.cuda.synchronize()
        # The other ranks must leave their GPUs ALONE while rank 0's single process drives all of them: an NCCL barrier
        # would park a spinning kernel on every GPU and the two processes would be time-sliced (measured: 2x slower).
        # They wait on the rendezvous store instead -- a CPU-side wait.
        import datetime
        store = dist.distributed_c10d._get_default_store()
        if rank == 0:
            try:
                group_leg = group_call_leg(sg, support, world, k, peak)
            finally:
                store.set("spl_group_leg_done", "1")
        else:
            store.wait(["spl_group_leg_done"], datetime.timedelta(seconds=900))
        dist.barrier()

    if rank == 0:
        gbs = job_bytes / ms_step / 1e6
        per_gpu_bytes = job_bytes / world
        traffic, traffic_note, traffic_warm = None, None, None
        tp = ROOT / "profiles" / "roofline_traffic.json"
        if tp.exists():
            tj = json.loads(tp.read_text())
            if single:
                # DRAM bytes per launch from the committed ncu capture -- but only if that capture was taken from the
                # kernel sources this run was built from
                sys.path.insert(0, str(ROOT / "tools"))
                from roofline_traffic import kernel_sha16
                cap = tj.get("c2_mask3_capture", {})
                if cap.get("kernel_source_sha16") == kernel_sha16():
                    traffic = tj.get("c2_mask3_dram_bytes_per_launch")
                    # the same launches without ncu's cache flush between them (the predicate bytes are kept in L2)
                    traffic_warm = tj.get("c2_mask3_dram_bytes_per_launch_warm_l2")
                    traffic_note = f"ncu --set full, {cap.get('report')}, kernel sources {cap.get('kernel_source_sha16')}"
                else:
                    traffic_note = ("the committed capture belongs to other kernel sources (" + str(cap.get("kernel_source_sha16")) +
                                    " vs " + kernel_sha16() + "): not reported")
            else:
                traffic = tj.get("c4_dram_bytes_per_launch_by_gpus", {}).get(str(world))
                traffic_note = tj.get("_source_c4", "")[:120]
        line = {
            "metric": "scanned_GB_per_s", "value": round(gbs, 2), "unit": "GB/s",
            "queries_per_s": round(1e3 / ms_step, 2),
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_step, 5),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(world),
            "e2e": {"value": round(job_bytes / ms_e2e / 1e6, 2), "unit": "GB/s", "queries_per_s": round(1e3 / ms_e2e, 2),
                    "ms_per_step": round(ms_e2e, 5), "h2d_bytes_per_step": 4 * DIM, "d2h_bytes_per_step": 32 * k + 4},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": round(per_gpu_bytes / ms_step / 1e6, 2), "peak": peak, "unit": "GB/s",
                         "frac": round(per_gpu_bytes / ms_step / 1e6 / peak, 4), "traffic": traffic, "traffic_source": traffic_note,
                         "traffic_warm_l2": traffic_warm,
                         "peak_source": peak_src, "algorithmic_bytes_per_launch": int(per_gpu_bytes),
                         "kernel": "spl::scan_*_kernel (one launch per query per GPU)"},
            "parity": parity,
        }
        if latency:
            line["latency_us"] = latency
        if variants:
            line["variants"] = variants
        for key, sm in (("small_batch_4q", small), ("small_batch_8q", small8)):
            if sm:
                sm["mask_0x3"]["speedup_vs_one_at_a_time"] = round(sm["mask_0x3"]["queries_per_s"] * ms_step / 1e3, 3)
                sm["mask_all_match_bit63"]["speedup_vs_one_at_a_time"] = round(
                    sm["mask_all_match_bit63"]["queries_per_s"] * variants["mask_all_match_bit63"]["ms_per_step"] / 1e3, 3)
                line[key] = sm
        if c3:
            line["c3_batched"] = c3
        if unlimited:
            line["unlimited_k0"] = unlimited
        if anchor:
            line["c4_single_gpu_anchor"] = anchor
        if c5:
            line["c5_mixed"] = c5
        if group_leg:
            line["group_call"] = group_leg
        if single and not args.skip_cpu:
            line["cpu_baseline"] = cpu_baseline_leg(k, head_mask)
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        print(json.dumps(line), flush=True)
    if m is not None:
        m.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def c3_batched_leg(sg, support, torch, dev, stream, args):
    """Config C3: 2^20 slots x 1024-d, 256 queries, L2 top-32, bf16 tensor-core contraction with fp32 re-rank."""
    n, dim, nq, k = 1 << 20, 1024, 256, 32
    peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
    tf_peak = float(peaks.get("bf16_tflops", 1590.0))
    m = sg.Mirror.open_dense(n, dim, flags=sg.OPT_BF16_MIRROR)
    try:
        m.fill_synthetic(SEED_CORPUS, SEED_BLOOM)
        q_host = support.synth_queries(SEED_QUERY, nq, dim)
        q = torch.from_numpy(q_host).to(dev)
        out = torch.zeros((nq, k, 32), dtype=torch.uint8, device=dev)
        cnt = torch.zeros(nq, dtype=torch.int32, device=dev)
        p = m.params(nq, k, order=sg.ORDER_L2, mode=sg.MODE_BATCH_BF16)
        steps = max(5, min(args.steps, 20))
        for _ in range(3):
            m.search_device(p, q.data_ptr(), out.data_ptr(), k, cnt.data_ptr(), stream)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            m.search_device(p, q.data_ptr(), out.data_ptr(), k, cnt.data_ptr(), stream)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        # e2e through the host-buffer call
        t0 = time.perf_counter()
        for _ in range(steps):
            res, counts = m.search(q_host, k, order=sg.ORDER_L2, mode=sg.MODE_BATCH_BF16)
        ms_e2e = (time.perf_counter() - t0) / steps * 1e3
        # parity: 8 of the 256 queries against the fp32 CPU oracle over a 2^17-slot prefix mirror
        sample = 1 << 17
        E, bloom, flags = support.synth_dense(SEED_CORPUS, SEED_BLOOM, 0, sample, dim)
        with sg.Mirror.open_dense(sample, dim, flags=sg.OPT_BF16_MIRROR) as ms_:
            ms_.fill_synthetic(SEED_CORPUS, SEED_BLOOM)
            got, gc = ms_.search(q_host[:8], k, order=sg.ORDER_L2, mode=sg.MODE_BATCH_BF16)
        ref, rc = support.oracle_search_dense(q_host[:8], E, bloom, flags, k=k, order=support.ORDER_L2)
        ok, ident = True, True
        try:
            for i in range(8):
                support.assert_topk_matches(got[i][:gc[i]], ref[i][:rc[i]], order=support.ORDER_L2)
                ident &= bool(np.array_equal(got[i][:gc[i]]["slot_idx"], ref[i][:rc[i]]["idx"]))
        except AssertionError:
            ok = False
        flops = 2.0 * nq * n * dim
        cnts = m.counters()
        tf_sustained = float(peaks.get("bf16_tflops_sustained", 1400.0))
        return {"workload": "C3: 2^20 slots x 1024-d, 256 queries, L2 top-32, bf16 tcgen05 contraction + fp32 re-rank + selection proof",
                "selection_proof": {"queries_through_host_call": int(cnts.batch_queries), "answered_by_fp32_fallback": int(cnts.batch_unproven),
                                    "rows_rescored_beyond_2k": int(cnts.batch_rescored_extra),
                                    "device_call_unanswered": int((cnt == -1).sum().item())},
                "frac_of_sustained_peak": round(flops / ms / 1e9 / tf_sustained, 4),
                "ms_per_batch": round(ms, 4), "queries_per_s": round(nq / ms * 1e3, 1),
                "e2e_ms_per_batch": round(ms_e2e, 4), "e2e_queries_per_s": round(nq / ms_e2e * 1e3, 1),
                "roofline": {"bound": "tensor", "achieved": round(flops / ms / 1e9, 1), "peak": tf_peak, "unit": "TFLOP/s",
                             "frac": round(flops / ms / 1e9 / tf_peak, 4),
                             "note": "algorithmic flops 2*nq*n*D per batch (the 1/64 sampling pass is extra work, not counted); peak = the "
                                     "measured cuBLAS bf16 BURST; the batch loop runs power-capped at ~1.4 GHz, where cuBLAS itself "
                                     "sustains bf16_tflops_sustained -- frac_of_sustained_peak is against that"},
                "hbm_GB_per_s": round((2.0 * n * dim + 13.0 * n) / ms / 1e6, 1),
                "parity": {"ok": ok, "indices_identical": ident, "checked_queries": 8, "checked_slots": sample},
                "all_counts_full": bool(int(cnt.min()) == k)}
    finally:
        m.close()


def parity_gate(m, sg, support, rows, row0, k, mask):
    """Re-score what the GPU returned with the oracle (rows regenerated from the counter RNG) and run one
    full oracle search over the first 2^17 slots of this rank's shard."""
    sample = min(rows, 1 << 17)
    E, bloom, flags = support.synth_dense(SEED_CORPUS, SEED_BLOOM, row0, sample, DIM)
    q = support.synth_queries(SEED_QUERY, 1, DIM)
    with sg.Mirror.open_dense(sample, DIM, global_row0=row0, device=m.info().device) as ms:
        ms.fill_synthetic(SEED_CORPUS, SEED_BLOOM)
        out, cnt = ms.search(q, k, bloom_mask=mask)
    ref, rc = support.oracle_search_dense(q, E, bloom, flags, k=k, bloom_mask=mask)
    got = out[0][:cnt[0]].copy()
    got["slot_idx"] -= row0
    try:
        support.assert_topk_matches(got, ref[0][:rc[0]], exact=False)
        idx_equal = bool(np.array_equal(got["slot_idx"], ref[0][:rc[0]]["idx"]))
        return {"checked_slots": sample, "ok": True, "indices_identical": idx_equal,
                "max_abs_dsim": float(np.max(np.abs(got["similarity"] - ref[0][:rc[0]]["similarity"]))) if len(got) else 0.0,
                "max_rel_ddist": float(np.max(np.abs(got["distance"] - ref[0][:rc[0]]["distance"]) / ref[0][:rc[0]]["distance"])) if len(got) else 0.0}
    except AssertionError as e:
        return {"checked_slots": sample, "ok": False, "error": str(e)[:200]}


def _oracle_rescore(support, q, global_idx):
    """(similarity, distance) of the synthetic corpus row `global_idx` against q, by the oracle's scorers"""
    E, _, flags = support.synth_dense(SEED_CORPUS, SEED_BLOOM, int(global_idx), 1, DIM, threads=1)
    if not (flags[0] & support.F_EMB):
        return 0.0, float(np.finfo(np.float32).max)
    return (float(support.oracle().splo_cosine(support._p(q), support._p(E[0]), DIM)),
            float(support.oracle().splo_euclid(support._p(q), support._p(E[0]), DIM)))


def merged_parity_gate(sharded, m, sg, support, torch, dist, q_host, q_dev, k, mask, world, rank, dev):
    """One query through scan -> all-gather -> merge, then: every rank re-scores ITS local hits with the oracle (rows
    regenerated from the counter RNG); rank 0 re-scores the k merged hits, checks their order, and checks that the merged
    list is exactly the k best of all ranks' local hits under the reference comparator + global slot index."""
    from libsplinter_b200.binding import packed_block_bytes
    p = m.params(1, k, mode=sg.MODE_AUTO, bloom_mask=mask)
    fin, fcnt = sharded.enqueue(p, q_dev[0:1])
    torch.cuda.synchronize()
    block, gathered = sharded._buffers(1)[:2]
    bb = packed_block_bytes(1, k)
    g = gathered.cpu().numpy().reshape(world, bb)
    lists = [g[w][:k * 32].view(sg.RESULT_DTYPE) for w in range(world)]
    counts = [int(g[w][k * 32:k * 32 + 4].view(np.uint32)[0]) for w in range(world)]
    merged = fin.cpu().numpy().reshape(k * 32).view(sg.RESULT_DTYPE)[:int(fcnt.cpu()[0])]
    q = q_host[0]

    def rescored_ok(recs):
        for r in recs:
            sim, dist_ = _oracle_rescore(support, q, r["slot_idx"])
            if abs(sim - float(r["similarity"])) > support.TOL_SIM:
                return False
            if dist_ < 3e38 and abs(dist_ - float(r["distance"])) > support.TOL_DIST_REL * dist_ + 1e-7:
                return False
        return True

    local_ok = rescored_ok(lists[rank][:counts[rank]])
    t = torch.tensor([1 if local_ok else 0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    out = {"local_hits_rescored_ok_all_ranks": bool(int(t.item()))}
    if rank == 0:
        pool = np.concatenate([lists[w][:counts[w]] for w in range(world)])
        # reference comparator (similarity desc, distance asc) + global slot index asc
        order = np.lexsort((pool["slot_idx"], pool["distance"], -pool["similarity"].astype(np.float64)))
        want = pool[order][:k]
        out["merged_is_topk_of_all_local_hits"] = bool(want.tobytes() == merged.tobytes())
        out["merged_hits_rescored_ok"] = bool(rescored_ok(merged))
        out["merged_ok"] = bool(out["merged_is_topk_of_all_local_hits"] and out["merged_hits_rescored_ok"]
                                and out["local_hits_rescored_ok_all_ranks"] and len(merged) == k)
        out["merged_from_ranks"] = sorted({int(r["slot_idx"]) * world // C4_ROWS for r in merged})
    return out


def c4_anchor_leg(sg, support, torch, dev, stream, peak, device_index):
    """The C4 workload (unfiltered cosine top-100) on ONE GPU over the largest power-of-two shard that fits (2^25 rows,
    103 GB): the like-for-like N = 1 point for the N = 2/4/8 lines, whose headline N = 1 run is C2."""
    rows, k = 1 << 25, C4_K
    free, _ = torch.cuda.mem_get_info()
    while rows * (4 * DIM + 17) * 1.05 > free and rows > (1 << 20):
        rows >>= 1
    m = sg.Mirror.open_dense(rows, DIM, device=device_index)
    try:
        m.fill_synthetic(SEED_CORPUS, SEED_BLOOM)
        q = torch.from_numpy(support.synth_queries(SEED_QUERY, 8, DIM)).to(dev)
        out = torch.zeros((k, 32), dtype=torch.uint8, device=dev)
        cnt = torch.zeros(1, dtype=torch.int32, device=dev)
        p = m.params(1, k, mode=sg.MODE_AUTO)
        for i in range(3):
            m.search_device(p, q[i].data_ptr(), out.data_ptr(), k, cnt.data_ptr(), stream)
        torch.cuda.synchronize()
        steps = 12
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            m.search_device(p, q[i % 8].data_ptr(), out.data_ptr(), k, cnt.data_ptr(), stream)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        _, bloom_h, flags_h, _ = m.read_mirror(0, rows, want_E=False)
        b = algorithmic_bytes(rows, int(((flags_h & sg.ROW_LIST) != 0).sum()))
        return {"workload": f"C4's query (unfiltered cosine top-{k}) over one {rows}-slot shard on one GPU",
                "slots": rows, "ms_per_step": round(ms, 4), "queries_per_s": round(1e3 / ms, 2),
                "scanned_GB_per_s": round(b / ms / 1e6, 1), "roofline_frac": round(b / ms / 1e6 / peak, 4),
                "queries_per_s_scaled_to_2p26": round(1e3 / ms * rows / C4_ROWS, 3)}
    finally:
        m.close()


def group_call_leg(sg, support, world, k, peak):
    """spl_gpu_group_search: ONE process, one context per GPU over slot ranges, peer copies + device merge -- what a C host
    calls.  Host buffers in, host buffers out (an end-to-end number), 2^22 slots per GPU, C4's query (unfiltered top-100)."""
    rows = (1 << 22) * world
    with sg.Group.open_dense(rows, DIM, list(range(world))) as g:
        g.fill_synthetic(SEED_CORPUS, SEED_BLOOM)
        q = support.synth_queries(SEED_QUERY, 8, DIM)
        for i in range(3):
            out, cnt = g.search(q[i], k, mode=sg.MODE_FAST)
        t0 = time.perf_counter()
        steps = 20
        for i in range(steps):
            out, cnt = g.search(q[i % 8], k, mode=sg.MODE_FAST)
        ms = (time.perf_counter() - t0) / steps * 1e3
        # parity: the merged hits re-scored by the oracle from regenerated rows, and sorted by the reference comparator
        ok = int(cnt[0]) == k
        for r in out[0][:k]:
            sim, dist_ = _oracle_rescore(support, q[(steps - 1) % 8], r["slot_idx"])
            ok &= abs(sim - float(r["similarity"])) <= support.TOL_SIM
        key = np.lexsort((out[0][:k]["slot_idx"], out[0][:k]["distance"], -out[0][:k]["similarity"].astype(np.float64)))
        ok &= bool(np.array_equal(key, np.arange(k)))
        b = 9 * rows + 4 * DIM * int(rows * 0.999) + 4 * DIM
        return {"workload": f"C4's query through spl_gpu_group_search: one process, {world} GPUs, {rows} slots, host buffers",
                "slots": rows, "e2e_ms_per_query": round(ms, 4), "queries_per_s": round(1e3 / ms, 2),
                "scanned_GB_per_s": round(b / ms / 1e6, 1), "per_gpu_frac_of_measured_hbm": round(b / ms / 1e6 / world / peak, 4),
                "parity_ok": bool(ok)}


def _unit_identity_violations(recs):
    """hits of unit queries over unit-or-zero rows: distance^2 = 2 - 2 similarity, or (0, 1) for a zero row, or
    (0, FLT_MAX, has_embedding 0) for a slot whose first copy is still pending"""
    sim = recs["similarity"].astype(np.float64)
    dist_ = recs["distance"].astype(np.float64)
    ok = np.abs(dist_ * dist_ - (2.0 - 2.0 * sim)) <= 2e-4
    ok |= (sim == 0.0) & (np.abs(dist_ - 1.0) <= 1e-5)
    ok |= (recs["has_embedding"] == 0) & (sim == 0.0)
    ok &= recs["epoch"] % 2 == 0
    return int(np.count_nonzero(~ok))


def c5_mixed_leg(sg, support, torch, dist, slots, world, rank, local_rank, dev, tstream):
    """Config C5: ONE reference-format store in /dev/shm, every rank mirrors its slot range (fp32 + bf16), writer
    PROCESSES mutate the store through the reference's API (80% set_embedding, 10% retrain_slot, 5% labels, 5% unset+set;
    splinter_chi_sao.c's disjoint key lanes), every rank keeps syncing its range, and nq = 64 batched cosine top-10
    searches run across all ranks (tensor-core shard search on a caller stream -> one all-gather -> merge).  Parity is
    sampled WHILE the writers run -- every vector in the store is a unit vector or zero, so a hit from a torn row breaks
    distance^2 = 2 - 2 similarity -- and against the CPU oracle over the whole store once they have stopped."""
    import mmap
    from libsplinter_b200.sharding import ShardedSearcher, shard_range
    NQ, K, MAXVAL, SECONDS, PERIOD = 64, 10, 64, 6.0, 0.1
    avail = 0
    for ln in open("/proc/meminfo"):
        if ln.startswith("MemAvailable"):
            avail = int(ln.split()[1]) * 1024
    while support.oracle().splo_image_size(slots, MAXVAL, DIM) * 1.3 > avail and slots > (1 << 16):
        slots >>= 1
    t = torch.tensor([slots], device=dev, dtype=torch.int64)
    dist.broadcast(t, 0)
    slots = int(t.item())
    size = support.oracle().splo_image_size(slots, MAXVAL, DIM)
    name = f"/splb200_c5_{os.environ.get('MASTER_PORT', '0')}"
    path = "/dev/shm" + name
    out = {"workload": f"C5: {slots} slots x 768-d in one store, concurrent writers + incremental re-upload + batched search",
           "slots": slots, "store_bytes": int(size), "nq": NQ, "k": K, "gpus": world}
    img = None
    if rank == 0:
        if os.path.exists(path):
            os.unlink(path)
        with open(path, "wb") as f:
            f.truncate(size)
        fd = os.open(path, os.O_RDWR)
        mm = mmap.mmap(fd, size)
        img = np.frombuffer(mm, np.uint8)
        t0 = time.perf_counter()
        assert support.oracle().splo_image_init(support._p(img), slots, MAXVAL, DIM) == 0
        out["keys_placed"] = int(support.image_fill_synthetic(img, slots // 2))
        out["host_build_s"] = round(time.perf_counter() - t0, 2)
    dist.barrier()
    m = None
    try:
        lo, cnt = shard_range(slots, world, rank)
        t0 = time.perf_counter()
        m = sg.Mirror.attach(name, slot_begin=lo, slot_count=cnt, flags=sg.OPT_BF16_MIRROR, device=local_rank,
                             sync_threads=max(2, (os.cpu_count() or 8) // world))
        attach_s = time.perf_counter() - t0
        dist.barrier()
        q_host = support.synth_queries(777, NQ, DIM)
        dq = torch.from_numpy(q_host).to(dev)
        S = ShardedSearcher(m, world, NQ, K, sg.ORDER_COSINE, dev, tstream.cuda_stream)
        p = m.params(NQ, K, mode=sg.MODE_BATCH_BF16, order=sg.ORDER_COSINE)
        for _ in range(3):
            S.search(p, dq)
        dist.barrier()
        syncs, stop = [], threading.Event()

        def syncer():
            while not stop.is_set():
                m.wait(100)
                t0 = time.perf_counter()
                st = m.sync()
                dt = time.perf_counter() - t0
                syncs.append((dt, st.rows_uploaded, st.skipped_busy))
                if PERIOD > dt:
                    time.sleep(PERIOD - dt)

        procs = []
        if rank == 0:
            writer, nkeys, lanes = str(support.REF_DIR / "ref_writer"), slots // 2, 4
            procs = [subprocess.Popen([writer, name, str(i * (nkeys // lanes)), str((i + 1) * (nkeys // lanes)), str(77 + i),
                                       str(int(SECONDS * 1e3))], stdout=subprocess.PIPE, text=True,
                                      env={**os.environ, "REF_WRITER_UNIT": "1"}) for i in range(lanes)]
        th = threading.Thread(target=syncer)
        th.start()
        flag = torch.ones(1, dtype=torch.int32, device=dev)
        batches, t_search, live_hits, live_bad, t_begin = 0, 0.0, 0, 0, time.perf_counter()
        while True:
            if rank == 0:
                flag.fill_(1 if any(pr.poll() is None for pr in procs) else 0)
            dist.broadcast(flag, 0)
            if int(flag.item()) == 0:
                break
            t0 = time.perf_counter()
            recs, counts = S.search(p, dq)        # incl. the D2H of the merged [64][10] records: an end-to-end batch
            t_search += time.perf_counter() - t0
            batches += 1
            for i in range(NQ):                   # live parity, outside the timed part
                live_bad += _unit_identity_violations(recs[i][:counts[i]])
                live_hits += int(counts[i])
        wall = time.perf_counter() - t_begin
        stop.set()
        th.join()
        ops = sum(int(pr.communicate()[0].split()[0]) for pr in procs) if rank == 0 else 0
        m.sync()                                   # quiesced: one more sweep and every shard equals its range of the store
        dist.barrier()
        recs, counts = S.search(p, dq)
        mine = {"rank": rank, "attach_s": round(attach_s, 3), "syncs": len(syncs),
                "sync_ms_median": round(float(np.median([x[0] for x in syncs])) * 1e3, 2) if syncs else None,
                "rows_uploaded": int(sum(x[1] for x in syncs)), "skipped_busy": int(sum(x[2] for x in syncs)),
                "live_hits_checked": live_hits, "live_violations": live_bad, "reissued_fp32": S.reissued}
        allr = [None] * world
        dist.all_gather_object(allr, mine)
        if rank == 0:
            ok = True
            try:
                for i in (0, 17, 63):              # merged cross-GPU answer == the CPU oracle over the whole store image
                    ref, _ = support.oracle_search_image(img, q_host[i], K)
                    assert counts[i] == len(ref)
                    support.assert_topk_matches(recs[i][:counts[i]], ref)
            except AssertionError as e:
                ok = False
                out["parity_error"] = str(e)[:200]
            out.update({"parity_after_quiesce": ok, "parity_live": {"hits_checked": sum(r["live_hits_checked"] for r in allr),
                                                                    "violations": sum(r["live_violations"] for r in allr)},
                        "writer_processes": 4, "writer_ops_per_s": round(ops / SECONDS),
                        "writer_mix": "80% set_embedding, 10% retrain_slot, 5% set/unset_label, 5% unset+set (reference API)",
                        "sync_period_ms": PERIOD * 1e3, "batched_searches": batches,
                        "ms_per_batch": round(t_search / max(batches, 1) * 1e3, 3),
                        "queries_per_s": round(NQ * batches / max(t_search, 1e-9)),
                        "queries_per_s_wall": round(NQ * batches / wall), "per_rank": allr})
            return out
        return None
    finally:
        if m is not None:
            m.close()
        dist.barrier()
        if rank == 0 and os.path.exists(path):
            os.unlink(path)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--skip-cpu", action="store_true", help="omit the cpu_baseline leg (profiling runs only)")
    ap.add_argument("--skip-extra", action="store_true", help="omit the C4 anchor / C5 legs (profiling runs only)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
