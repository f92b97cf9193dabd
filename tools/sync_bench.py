"""Boundary-layer measurement ("upload/sync time reported separately"): full first sync of a
2^20-slot reference-format store, idle re-sync, and incremental syncs while writer processes mutate the
store through the reference API; searches run in between.  Run on the GPU box."""
import json, mmap, os, subprocess, sys, threading, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, libsplinter_b200 as sg, support

slots = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
fill = float(sys.argv[2]) if len(sys.argv) > 2 else 0.98     # load factor of the reference's open-addressing table
nkeys = int(slots * fill)
dim, maxval = 768, 64
name = f"/splb200_syncbench_{os.getpid()}"
path = "/dev/shm" + name
size = support.oracle().splo_image_size(slots, maxval, dim)
with open(path, "wb") as f:
    f.truncate(size)
fd = os.open(path, os.O_RDWR)
mm = mmap.mmap(fd, size)
img = np.frombuffer(mm, np.uint8)
t0 = time.perf_counter()
assert support.oracle().splo_image_init(support._p(img), slots, maxval, dim) == 0
placed = support.image_fill_synthetic(img, nkeys)
t_build = time.perf_counter() - t0
out = {"slots": slots, "fill": fill, "keys_placed": int(placed), "store_bytes": size, "host_build_s": round(t_build, 2)}
try:
    warm = sg.Mirror.open_dense(256, dim)      # CUDA context creation is not the sync layer's cost
    warm.close()
    t0 = time.perf_counter()
    m = sg.Mirror.attach(name, sync_threads=16, flags=sg.OPT_BF16_MIRROR)   # C5 searches are batched: keep the bf16 shadow too
    t_attach = time.perf_counter() - t0
    st = m.sync()
    out["first_sync"] = {"attach_plus_full_sync_s": round(t_attach, 3), "rows": int(placed),
                         "rows_per_s": round(placed / t_attach), "GB_per_s": round(placed * 3072 / t_attach / 1e9, 2)}
    out["idle_sync_ms"] = round(st.sweep_ms + st.upload_ms, 2)
    q = support.synth_queries(4321, 4, dim)
    hits, cnt = m.search(q[0], 10, bloom_mask=3)
    ref, rc = support.oracle_search_image(img, q[0], 10, bloom_mask=3)
    support.assert_topk_matches(hits[0][:cnt[0]], ref)
    out["parity_after_full_sync"] = True
    # every slot rewritten (epoch += 2, vector negated, header epoch moved): one full re-upload, timed by the library
    slot_size = 128 + 4 * dim
    body = img[5440:5440 + slots * slot_size].reshape(slots, slot_size)
    body[:, 8:16].view(np.uint64)[...] += np.uint64(2) * (body[:, 8:16].view(np.uint64) != 0)
    body[:, 128:].view(np.float32)[...] *= np.float32(-1.0)
    img[16:24].view(np.uint64)[0] += np.uint64(2)
    t0 = time.perf_counter()
    st = m.sync()
    dt = time.perf_counter() - t0
    out["full_resync"] = {"wall_s": round(dt, 3), "rows_uploaded": int(st.rows_uploaded), "sweep_ms": round(st.sweep_ms, 1),
                          "upload_ms": round(st.upload_ms, 1), "rows_per_s": round(st.rows_uploaded / dt),
                          "GB_per_s": round(st.rows_uploaded * 3072 / dt / 1e9, 2)}
    ref, rc = support.oracle_search_image(img, q[1], 10, bloom_mask=3)
    hits, cnt = m.search(q[1], 10, bloom_mask=3)
    support.assert_topk_matches(hits[0][:cnt[0]], ref)
    if support.have_ref():
        writer = str(support.REF_DIR / "ref_writer")
        lanes = 4
        procs = [subprocess.Popen([writer, name, str(i * (nkeys // lanes)), str((i + 1) * (nkeys // lanes)), str(77 + i), "4000"],
                                  stdout=subprocess.PIPE, text=True) for i in range(lanes)]
        syncs = []
        qb = support.synth_queries(777, 64, dim)
        stop = threading.Event()
        batched = dict(calls=0, secs=0.0)

        def searcher():   # C5: nq = 64 batched cosine top-10 searches while the writers and the syncs run
            while not stop.is_set():
                t0 = time.perf_counter()
                out, cnt = m.search(qb, 10, mode=sg.MODE_BATCH_BF16)
                batched["secs"] += time.perf_counter() - t0
                batched["calls"] += 1

        th = threading.Thread(target=searcher)
        th.start()
        period = float(os.environ.get("SYNC_PERIOD_MS", "0")) / 1e3    # 0 = sweep back to back
        while any(p.poll() is None for p in procs):
            m.wait(100)                      # event bus / H->epoch gate (returns at once while writers are active)
            t0 = time.perf_counter()
            s = m.sync()
            dt = time.perf_counter() - t0
            syncs.append((dt, s.rows_uploaded, s.meta_updated, s.skipped_busy, s.aba_rechecked, s.sweep_ms, s.upload_ms))
            if period > dt:
                time.sleep(period - dt)      # a sweep reads one to three cache lines of EVERY slot: pace it on big stores
        stop.set()
        th.join()
        ops = sum(int(p.communicate()[0].split()[0]) for p in procs)
        s = m.sync()
        dts = np.array([x[0] for x in syncs]); rows = np.array([x[1] for x in syncs])
        out["under_writers"] = {"writer_processes": lanes, "writer_ops_per_s": round(ops / 4.0),
                                "syncs": len(syncs), "sync_ms_median": round(float(np.median(dts)) * 1e3, 2),
                                "rows_per_sync_median": int(np.median(rows)),
                                "sweep_ms_median": round(float(np.median([x[5] for x in syncs])), 2),
                                "upload_ms_median": round(float(np.median([x[6] for x in syncs])), 2),
                                "dirty_rows_per_s": round(float(rows.sum() / dts.sum())),
                                "skipped_busy_total": int(sum(x[3] for x in syncs)), "aba_rechecked_total": int(sum(x[4] for x in syncs)),
                                "sync_period_ms": period * 1e3, "writer_mix": "80% set_embedding, 10% retrain_slot, 5% set/unset_label, 5% unset+set (ref_writer default)",
                                "batched_search": {"nq": 64, "k": 10, "mode": "BATCH_BF16 cosine", "calls": batched["calls"],
                                                   "ms_per_call": round(batched["secs"] / max(batched["calls"], 1) * 1e3, 3),
                                                   "queries_per_s": round(64 * batched["calls"] / max(batched["secs"], 1e-9))}}
        # quiesced: mirror == store on a sample of slots
        E, bloom, flags, epoch = m.read_mirror(0, 65536)
        Es = np.zeros((65536, dim), np.float32)
        slot_size = 128 + 4 * dim
        for i in range(0, 65536, 997):
            if not (flags[i] & 3):
                continue          # not a live slot (never used, or unset): its mirror row is not a candidate and is left alone
            off = 5440 + i * slot_size + 128
            assert np.array_equal(E[i].view(np.uint32), img[off:off + 4 * dim].view(np.uint32)), i
        out["mirror_equals_store_after_quiesce"] = True
    m.close()
finally:
    os.close(fd); os.unlink(path)      # the mapping itself goes away with the process
print(json.dumps(out))
