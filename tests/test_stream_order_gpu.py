"""Mirror updates vs caller-stream searches: the seqlock promise carried onto the device.

The reference's reader never returns a half-written slot (splinter.c:598-605, contract splinter.h:368-396) and its
stress tool counts integrity failures WHILE the writers run (splinter_stress.c:123-139,170-175).  Here writer
processes mutate a live store through the reference's own API, one thread loops spl_gpu_sync (scatter kernels on the
context's stream) and another loops spl_gpu_search_device on ITS OWN stream -- the multi-GPU path's call pattern.
Every vector the writers produce is a unit vector (or the zero row a retrain / fresh set leaves), and so is the
query, so a hit computed from a row that really existed obeys   distance^2 = 2 - 2 * similarity   (zero row:
similarity 0, distance 1).  A row stitched together from two versions has another norm and breaks the identity.

Also here: a query the device-buffer API cannot answer comes back as SPL_GPU_COUNT_UNANSWERED, never as "0 hits",
and the marker survives the shard merge.
"""
import os
import subprocess
import threading

import numpy as np
import pytest

import libsplinter_b200 as sg
import support

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

DIM, SLOTS, NKEYS = 768, 4096, 3000


def _unit(rng, n):
    v = rng.standard_normal((n, DIM))
    return (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(np.float32)


def _violations(hits):
    """hits of unit queries over unit-or-zero rows: how many break distance^2 = 2 - 2 similarity"""
    sim = hits["similarity"].astype(np.float64)
    dist = hits["distance"].astype(np.float64)
    ok_unit = np.abs(dist * dist - (2.0 - 2.0 * sim)) <= 2e-4
    ok_zero = (sim == 0.0) & (np.abs(dist - 1.0) <= 1e-5)
    ok_meta = (hits["epoch"] % 2 == 0) & ((hits["bloom"] & ~np.uint64(0xFF00)) == 0)
    return int(np.count_nonzero(~((ok_unit | ok_zero) & ok_meta)))


@pytest.mark.skipif(not support.have_ref(), reason="needs oracle/_ref")
def test_device_searches_on_their_own_stream_never_see_a_torn_slot():
    ref = support.ref_store()
    name = f"/splb200_order_{os.getpid()}"
    path = "/dev/shm" + name
    if os.path.exists(path):
        os.unlink(path)
    assert ref.splinter_create(name.encode(), SLOTS, 64) == 0
    rng = np.random.default_rng(11)
    rows = _unit(rng, NKEYS)
    for k in range(NKEYS):
        key = f"doc{k:08d}".encode()
        assert ref.splinter_set(key, b"v", 1) == 0
        assert ref.splinter_set_embedding(key, support._p(rows[k])) == 0
    try:
        m = sg.Mirror.attach(name, sync_threads=4, flags=sg.OPT_BF16_MIRROR)
        writer = str(support.REF_DIR / "ref_writer")
        env = {**os.environ, "REF_WRITER_UNIT": "1"}
        procs = [subprocess.Popen([writer, name, str(lo), str(lo + NKEYS // 4), str(2000 + lo), "4000", "85", "10", "5"],
                                  stdout=subprocess.PIPE, text=True, env=env) for lo in range(0, NKEYS, NKEYS // 4)]
        stop = threading.Event()
        stats = dict(syncs=0, rows=0, searches=0, hits=0, bad=0, unanswered=0, errors=[])

        def syncer():
            try:
                while not stop.is_set():
                    st = m.sync()
                    stats["syncs"] += 1
                    stats["rows"] += st.rows_uploaded
            except Exception as e:  # noqa: BLE001
                stats["errors"].append(repr(e))

        def searcher():
            try:
                torch.cuda.set_device(0)
                stream = torch.cuda.Stream()
                q1 = _unit(np.random.default_rng(5), 1)
                qb = _unit(np.random.default_rng(6), 24)
                k = 100
                plans = []
                for q, mode in ((q1, sg.MODE_FAST), (qb, sg.MODE_BATCH_BF16), (qb[:4], sg.MODE_FAST)):
                    nq = q.shape[0]
                    plans.append((m.params(nq, k, mode=mode), torch.from_numpy(q).cuda(),
                                  torch.zeros((nq, k, 32), dtype=torch.uint8, device="cuda"),
                                  torch.zeros(nq, dtype=torch.int32, device="cuda"), nq))
                torch.cuda.synchronize()
                while not stop.is_set():
                    for p, dq, out, cnt, nq in plans:
                        with torch.cuda.stream(stream):
                            m.search_device(p, dq.data_ptr(), out.data_ptr(), k, cnt.data_ptr(), stream.cuda_stream)
                            h_out = out.cpu()
                            h_cnt = cnt.cpu()
                        stream.synchronize()
                        counts = h_cnt.numpy().astype(np.uint32)
                        recs = h_out.numpy().view(sg.RESULT_DTYPE).reshape(nq, k)
                        for i in range(nq):
                            if counts[i] == sg.COUNT_UNANSWERED:
                                stats["unanswered"] += 1
                                continue
                            assert counts[i] == k
                            stats["bad"] += _violations(recs[i][:k])
                            stats["hits"] += k
                        stats["searches"] += nq
            except Exception as e:  # noqa: BLE001
                stats["errors"].append(repr(e))

        threads = [threading.Thread(target=syncer), threading.Thread(target=searcher)]
        for t in threads:
            t.start()
        for p in procs:
            p.wait()
        stop.set()
        for t in threads:
            t.join()
        total_ops = sum(int(p.communicate()[0].split()[0]) for p in procs)
        assert not stats["errors"], stats
        assert total_ops > 1000 and stats["syncs"] >= 10 and stats["rows"] > 1000, stats
        assert stats["searches"] >= 10_000, stats
        assert stats["bad"] == 0, stats
        m.close()
    finally:
        ref.splinter_close()
        if os.path.exists(path):
            os.unlink(path)


def _dev_search(m, q, k, stream, **kw):
    nq = q.shape[0]
    dq = torch.from_numpy(q).cuda()
    out = torch.zeros((nq, k, 32), dtype=torch.uint8, device="cuda")
    cnt = torch.zeros(nq, dtype=torch.int32, device="cuda")
    p = m.params(nq, k, **kw)
    with torch.cuda.stream(stream):
        m.search_device(p, dq.data_ptr(), out.data_ptr(), k, cnt.data_ptr(), stream.cuda_stream)
    stream.synchronize()
    return out, cnt


def test_device_api_marks_what_it_cannot_answer_and_the_merge_hands_the_marker_on():
    """> 4096 exact copies of one row overflow the batched path's per-query candidate list: through the device
    API that query must come back UNANSWERED (not as 0 hits), a FAST re-issue answers it, and a merge of shard
    lists containing the marker is itself UNANSWERED."""
    n, dim, k = 16_384, 768, 10
    rng = np.random.default_rng(3)
    E = rng.standard_normal((n, dim)).astype(np.float32)
    E /= np.linalg.norm(E, axis=1, keepdims=True)
    dup = rng.choice(n, 6000, replace=False)
    E[dup] = E[dup[0]]
    q = np.stack([E[dup[0]] + 0.01 * rng.standard_normal(dim).astype(np.float32),      # hits the duplicates
                  rng.standard_normal(dim).astype(np.float32)]).astype(np.float32)
    stream = torch.cuda.Stream()
    with sg.Mirror.open_dense(n, dim, flags=sg.OPT_BF16_MIRROR) as m:
        m.dense_upload(0, E)
        out, cnt = _dev_search(m, q, k, stream, mode=sg.MODE_BATCH_BF16)
        c = cnt.cpu().numpy().astype(np.uint32)
        assert c[0] == sg.COUNT_UNANSWERED, c
        ref, rc = support.oracle_search_dense(q, E, None, None, k=k)
        if c[1] != sg.COUNT_UNANSWERED:
            got = out.cpu().numpy().view(sg.RESULT_DTYPE).reshape(2, k)
            support.assert_topk_matches(got[1][:c[1]], ref[1][:rc[1]])
        # the re-issue the header prescribes
        out1, cnt1 = _dev_search(m, q[:1], k, stream, mode=sg.MODE_FAST)
        got1 = out1.cpu().numpy().view(sg.RESULT_DTYPE).reshape(1, k)
        assert int(cnt1.cpu()[0]) == k
        support.assert_topk_matches(got1[0], ref[0][:rc[0]])
        # the host-buffer call falls back by itself and counts it
        before = m.counters().batch_unproven
        host, hc = m.search(q, k, mode=sg.MODE_BATCH_BF16)
        assert hc[0] == k and m.counters().batch_unproven > before
        support.assert_topk_matches(host[0][:k], ref[0][:rc[0]])
        # merge: two "shards", the first one unanswered for query 0
        lists = torch.cat([out.reshape(2 * k, 32), out.reshape(2 * k, 32)]).contiguous()
        counts = torch.cat([cnt, cnt]).contiguous()
        fin = torch.zeros((2, k, 32), dtype=torch.uint8, device="cuda")
        fcnt = torch.zeros(2, dtype=torch.int32, device="cuda")
        with torch.cuda.stream(stream):
            m.merge_topk_batch_device(lists.data_ptr(), counts.data_ptr(), 2, 2, k, sg.ORDER_COSINE, k, fin.data_ptr(), k,
                                      fcnt.data_ptr(), stream.cuda_stream)
        stream.synchronize()
        assert fcnt.cpu().numpy().astype(np.uint32)[0] == sg.COUNT_UNANSWERED
