"""The device-pointer entry points (spl_gpu_search_device, spl_gpu_merge_topk_device) used by the multi-GPU path:
same results, byte for byte, as the host-buffer call on the same mirror, and shard lists that merge to the
single-store answer.  Device buffers and the stream come from torch (plumbing only)."""
import numpy as np
import pytest

import libsplinter_b200 as sg
import support

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _dev_search(m, q, k, stream, **kw):
    nq = q.shape[0]
    dq = torch.from_numpy(q).cuda()
    out = torch.zeros((nq, k, 32), dtype=torch.uint8, device="cuda")
    cnt = torch.zeros(nq, dtype=torch.int32, device="cuda")
    p = m.params(nq, k, **kw)
    with torch.cuda.stream(stream):
        m.search_device(p, dq.data_ptr(), out.data_ptr(), k, cnt.data_ptr(), stream.cuda_stream)
    stream.synchronize()
    return out.cpu().numpy().view(sg.RESULT_DTYPE).reshape(nq, k), cnt.cpu().numpy().astype(np.uint32), out, cnt


@pytest.mark.parametrize("mode", [sg.MODE_FAST, sg.MODE_EXACT, sg.MODE_AUTO])
@pytest.mark.parametrize("nq,k", [(1, 10), (5, 10), (4, 32), (3, 100)])
def test_device_buffers_give_the_host_call_s_bytes(mode, nq, k):
    n, dim = 30_000, 768
    E, bloom, flags = support.synth_dense(1234, 99, 0, n, dim)
    q = support.synth_queries(55, nq, dim)
    stream = torch.cuda.Stream()
    with sg.Mirror.open_dense(n, dim) as m:
        m.dense_upload(0, E, bloom, flags)
        for kw in (dict(bloom_mask=0x1), dict(order=sg.ORDER_L2, bloom_mask=0x2, exclude_slot=11)):
            host, hc = m.search(q, k, mode=mode, **kw)
            dev, dc, _, _ = _dev_search(m, q, k, stream, mode=mode, **kw)
            assert np.array_equal(hc, dc)
            for i in range(nq):
                assert host[i][:hc[i]].tobytes() == dev[i][:dc[i]].tobytes()


def test_shard_lists_merge_to_the_single_store_answer():
    n, dim, k, shards = 45_000, 768, 20, 3
    E, bloom, flags = support.synth_dense(77, 99, 0, n, dim)
    q = support.synth_queries(56, 1, dim)
    stream = torch.cuda.Stream()
    with sg.Mirror.open_dense(n, dim) as whole:
        whole.dense_upload(0, E, bloom, flags)
        want, wc = whole.search(q, k, bloom_mask=0x1)
    per = n // shards
    lists, counts, mirrors = [], [], []
    try:
        for s in range(shards):
            lo, hi = s * per, (n if s == shards - 1 else (s + 1) * per)
            m = sg.Mirror.open_dense(hi - lo, dim, global_row0=lo)
            m.dense_upload(0, E[lo:hi], bloom[lo:hi], flags[lo:hi])
            mirrors.append(m)
            _, _, d_out, d_cnt = _dev_search(m, q, k, stream, bloom_mask=0x1)
            lists.append(d_out.reshape(k, 32)); counts.append(d_cnt)
        gathered = torch.cat(lists).contiguous()          # what all_gather_into_tensor delivers on every rank
        gcnt = torch.cat(counts).contiguous()
        fin = torch.zeros((k, 32), dtype=torch.uint8, device="cuda")
        fcnt = torch.zeros(1, dtype=torch.int32, device="cuda")
        with torch.cuda.stream(stream):
            mirrors[0].merge_topk_device(gathered.data_ptr(), gcnt.data_ptr(), shards, k, sg.ORDER_COSINE, k,
                                         fin.data_ptr(), fcnt.data_ptr(), stream.cuda_stream)
        stream.synchronize()
        got = fin.cpu().numpy().view(sg.RESULT_DTYPE).reshape(k)
        assert int(fcnt.item()) == int(wc[0]) == k
        assert got.tobytes() == want[0][:k].tobytes()     # global slot indices, same order, same scores
        # ragged: one shard contributes nothing, another only a few
        gcnt2 = gcnt.clone(); gcnt2[1] = 0; gcnt2[2] = 3
        with torch.cuda.stream(stream):
            mirrors[0].merge_topk_device(gathered.data_ptr(), gcnt2.data_ptr(), shards, k, sg.ORDER_COSINE, k,
                                         fin.data_ptr(), fcnt.data_ptr(), stream.cuda_stream)
        stream.synchronize()
        got = fin.cpu().numpy().view(sg.RESULT_DTYPE).reshape(k)[:int(fcnt.item())]
        pool = np.concatenate([lists[0].cpu().numpy().view(sg.RESULT_DTYPE).reshape(k)[:int(gcnt[0])],
                               lists[2].cpu().numpy().view(sg.RESULT_DTYPE).reshape(k)[:3]])
        order = np.lexsort((pool["slot_idx"], pool["distance"], -pool["similarity"]))
        assert got.tobytes() == pool[order][:k].tobytes()
    finally:
        for m in mirrors:
            m.close()


def test_batched_mode_through_device_buffers():
    n, dim, nq, k = 20_000, 1024, 40, 16
    E, bloom, flags = support.synth_dense(5, 99, 0, n, dim)
    q = support.synth_queries(57, nq, dim)
    stream = torch.cuda.Stream()
    with sg.Mirror.open_dense(n, dim, flags=sg.OPT_BF16_MIRROR) as m:
        m.dense_upload(0, E, bloom, flags)
        host, hc = m.search(q, k, mode=sg.MODE_BATCH_BF16, order=sg.ORDER_L2)
        dev, dc, _, _ = _dev_search(m, q, k, stream, mode=sg.MODE_BATCH_BF16, order=sg.ORDER_L2)
        assert np.array_equal(hc, dc)
        for i in range(nq):
            assert host[i][:hc[i]].tobytes() == dev[i][:dc[i]].tobytes()


def test_batched_merge_of_gathered_shard_blocks():
    """[shard][query][k] blocks as an all-gather delivers them -> one launch merges every query; equals the
    single-store answer per query, and equals the one-query merge entry point."""
    n, dim, k, shards, nq = 36_000, 768, 10, 3, 7
    E, bloom, flags = support.synth_dense(91, 99, 0, n, dim)
    q = support.synth_queries(58, nq, dim)
    stream = torch.cuda.Stream()
    with sg.Mirror.open_dense(n, dim) as whole:
        whole.dense_upload(0, E, bloom, flags)
        want, wc = whole.search(q, k, bloom_mask=0x3, mode=sg.MODE_FAST)
    per = n // shards
    blocks, counts, mirrors = [], [], []
    try:
        for s in range(shards):
            lo, hi = s * per, (n if s == shards - 1 else (s + 1) * per)
            m = sg.Mirror.open_dense(hi - lo, dim, global_row0=lo)
            m.dense_upload(0, E[lo:hi], bloom[lo:hi], flags[lo:hi])
            mirrors.append(m)
            _, _, d_out, d_cnt = _dev_search(m, q, k, stream, bloom_mask=0x3, mode=sg.MODE_FAST)
            blocks.append(d_out.reshape(nq * k, 32)); counts.append(d_cnt)
        gathered = torch.cat(blocks).contiguous()
        gcnt = torch.cat(counts).contiguous()
        fin = torch.zeros((nq, k, 32), dtype=torch.uint8, device="cuda")
        fcnt = torch.zeros(nq, dtype=torch.int32, device="cuda")
        with torch.cuda.stream(stream):
            mirrors[0].merge_topk_batch_device(gathered.data_ptr(), gcnt.data_ptr(), shards, nq, k, sg.ORDER_COSINE, k,
                                               fin.data_ptr(), k, fcnt.data_ptr(), stream.cuda_stream)
        stream.synchronize()
        got = fin.cpu().numpy().view(sg.RESULT_DTYPE).reshape(nq, k)
        gc = fcnt.cpu().numpy()
        for i in range(nq):
            assert gc[i] == wc[i]
            assert got[i][:gc[i]].tobytes() == want[i][:wc[i]].tobytes()
    finally:
        for m in mirrors:
            m.close()


def test_pipelined_query_loop_returns_what_one_at_a_time_returns():
    """SPL_SEARCH_PIPELINED lets a scan start during the previous scan's merge tail (alternating scratch sets, results
    written in stream order).  200 different queries back to back, each into its own output area, then 50 more all into
    ONE area: the bytes must be those of the unflagged loop, and the shared area must hold the LAST query's answer."""
    n, dim, k, nq = 300_000, 768, 10, 200
    stream = torch.cuda.Stream()
    rng = np.random.default_rng(8)
    q = rng.standard_normal((nq, dim)).astype(np.float32)
    with sg.Mirror.open_dense(n, dim) as m:
        m.fill_synthetic(1234, 99)
        dq = torch.from_numpy(q).cuda()
        outs = {}
        for flags in (0, sg.SEARCH_PIPELINED):
            out = torch.zeros((nq, k, 32), dtype=torch.uint8, device="cuda")
            cnt = torch.zeros(nq, dtype=torch.int32, device="cuda")
            torch.cuda.synchronize()
            for mask in (0x3, 0x7, 0):
                p = m.params(1, k, bloom_mask=mask, mode=sg.MODE_FAST, flags=flags)
                with torch.cuda.stream(stream):
                    for i in range(nq):
                        m.search_device(p, dq[i].data_ptr(), out[i].data_ptr(), k, cnt[i:].data_ptr(), stream.cuda_stream)
                stream.synchronize()
                outs[(flags, mask)] = (out.cpu().numpy().copy(), cnt.cpu().numpy().copy())
        for mask in (0x3, 0x7, 0):
            a, b = outs[(0, mask)], outs[(sg.SEARCH_PIPELINED, mask)]
            assert np.array_equal(a[1], b[1]) and np.all(a[1] == k)
            assert a[0].tobytes() == b[0].tobytes()
        # one shared output area: stream order decides who writes last
        one = torch.zeros((k, 32), dtype=torch.uint8, device="cuda")
        c1 = torch.zeros(1, dtype=torch.int32, device="cuda")
        p = m.params(1, k, bloom_mask=0x7, mode=sg.MODE_FAST, flags=sg.SEARCH_PIPELINED)
        with torch.cuda.stream(stream):
            for i in range(50):
                m.search_device(p, dq[i].data_ptr(), one.data_ptr(), k, c1.data_ptr(), stream.cuda_stream)
        stream.synchronize()
        assert one.cpu().numpy().tobytes() == outs[(0, 0x7)][0][49].tobytes()
