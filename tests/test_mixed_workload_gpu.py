"""Mixed workload (config C5 in miniature): writer PROCESSES mutate a live shm store through the
reference's own API (splinter_set_embedding / retrain / labels / unset+set, oracle/_ref/ref_writer.c, in
the style of the reference's MRMW stress tool splinter_chi_sao.c) while this process keeps syncing the
HBM mirror and searching it.  Checks, like splinter_stress.c's integrity_fail counter:
  * no torn row ever reaches the mirror (every row the writers produce is self-describing);
  * busy slots are skipped, not spun on, and are picked up later;
  * once the writers stop, one more sync makes the mirror byte-equal to the store, and a search over it
    equals the CPU oracle run on the store image."""
import os
import subprocess
import threading
import time

import numpy as np
import pytest

import libsplinter_b200 as sg
import support

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not support.have_ref(), reason="needs oracle/_ref")]

DIM, SLOTS, NKEYS = 768, 4096, 3000


def _consistent(rows):
    """rows written by ref_writer: v[1] = version, v[d] = version + d/1024 for d >= 2; zero rows are legal
    (retrain / fresh set); initial rows have v[1] == 0.5 marker handled by the caller."""
    ver = rows[:, 1:2]
    d = np.arange(2, DIM, dtype=np.float32)[None, :] / np.float32(1024.0)
    want = (ver + d).astype(np.float32)
    return np.all(rows[:, 2:] == want, axis=1)


def test_concurrent_writers_sync_and_search():
    ref = support.ref_store()
    name = f"/splb200_mixed_{os.getpid()}"
    path = "/dev/shm" + name
    if os.path.exists(path):
        os.unlink(path)
    assert ref.splinter_create(name.encode(), SLOTS, 64) == 0
    vec = np.zeros(DIM, np.float32)
    for k in range(NKEYS):
        key = f"doc{k:08d}".encode()
        assert ref.splinter_set(key, b"v", 1) == 0
        vec[0] = k; vec[1] = 1.0
        vec[2:] = np.float32(1.0) + np.arange(2, DIM, dtype=np.float32) / np.float32(1024.0)
        assert ref.splinter_set_embedding(key, support._p(vec)) == 0
    try:
        m = sg.Mirror.attach(name, sync_threads=4, flags=sg.OPT_BF16_MIRROR)   # syncs also maintain the bf16 shadow
        writer = str(support.REF_DIR / "ref_writer")
        procs = [subprocess.Popen([writer, name, str(lo), str(lo + NKEYS // 4), str(1000 + lo), "2500"],
                                  stdout=subprocess.PIPE, text=True) for lo in range(0, NKEYS, NKEYS // 4)]
        stats = dict(syncs=0, rows=0, busy=0, torn=0, searches=0)
        stop = threading.Event()

        def searcher():  # searches run concurrently with syncs from another thread
            q = np.ones((1, DIM), np.float32)
            qb = np.random.default_rng(3).standard_normal((20, DIM)).astype(np.float32)
            while not stop.is_set():
                out, cnt = m.search(q, 10, bloom_mask=0)
                assert cnt[0] == 10
                # the batched tensor-core path reads the bf16 shadow while scatter kernels rewrite it
                out, cnt = m.search(qb, 10, mode=sg.MODE_BATCH_BF16)
                assert np.all(cnt == 10)
                stats["searches"] += 2

        th = threading.Thread(target=searcher)
        th.start()
        while any(p.poll() is None for p in procs):
            st = m.sync()
            stats["syncs"] += 1; stats["rows"] += st.rows_uploaded; stats["busy"] += st.skipped_busy
            E, bloom, flags, epoch = m.read_mirror(0, SLOTS)
            live = (flags & sg.ROW_EMB).astype(bool) & (np.abs(E).sum(axis=1) > 0)
            bad = ~_consistent(E[live])
            if bad.any() and "example" not in stats:   # keep the first offender for the failure message
                r = E[live][bad][0]
                d = np.arange(2, DIM, dtype=np.float32) / np.float32(1024.0)
                wrong = np.nonzero(r[2:] != (r[1] + d).astype(np.float32))[0] + 2
                stats["example"] = dict(key=float(r[0]), ver=float(r[1]), n_wrong=int(wrong.size), first=int(wrong[0]),
                                        last=int(wrong[-1]), vals=[float(x) for x in r[wrong[:3]]],
                                        slot=int(np.nonzero(live)[0][np.nonzero(bad)[0][0]]),
                                        epoch=int(epoch[np.nonzero(live)[0][np.nonzero(bad)[0][0]]]))
            stats["torn"] += int(bad.sum())
        stop.set()
        th.join()
        outs = [p.communicate()[0].split() for p in procs]
        total_ops = sum(int(o[0]) for o in outs)
        assert total_ops > 1000 and stats["syncs"] >= 3 and stats["rows"] > 100 and stats["searches"] > 0
        assert stats["torn"] == 0, stats
        # quiesced: one more sync and the mirror equals the store
        st = m.sync()
        assert st.skipped_busy == 0
        img = np.frombuffer(open(path, "rb").read(), np.uint8).copy()
        support.oracle().splo_image_dim(support._p(img), img.nbytes)
        Es, bs, fs, es = support.image_to_dense(img, DIM)
        E, bloom, flags, epoch = m.read_mirror(0, SLOTS)
        live = fs != 0
        assert np.array_equal(flags & 3, fs & 3)
        assert np.array_equal(E.view(np.uint32)[live], Es.view(np.uint32)[live])
        assert np.array_equal(bloom[live], bs[live]) and np.array_equal(epoch[live], es[live])
        st2 = m.sync()
        assert st2.rows_uploaded == 0 and st2.meta_updated == 0   # idle store costs no upload
        q = np.random.default_rng(1).standard_normal(DIM).astype(np.float32)
        for mask in (0, 1 << 9):
            out, cnt = m.search(q, 20, bloom_mask=mask)
            ref_hits, _ = support.oracle_search_image(img, q, 20, bloom_mask=mask)
            support.assert_topk_matches(out[0][:cnt[0]], ref_hits)
        m.close()
    finally:
        ref.splinter_close()
        if os.path.exists(path):
            os.unlink(path)
