"""spl_gpu_scan_noalloc -- the scan entry point that allocates no device memory.

Some users of this library are not allowed to allocate on the GPU at query time: a host that has
given every byte of device memory to its own arena, a latency-critical loop that cannot take the
device-wide stall of cudaMalloc/cudaFree, a process whose device memory is full on purpose.  For
them the scan's scratch lives in a fixed array inside the library image.  This test drives that
entry point through the C ABI with caller-owned device buffers (torch tensors stand in for the
caller's arena) and checks it against the ordinary spl_gpu_search on the same data, twice in a
row -- the second call reuses the scratch the first one left behind.

Collected last on purpose (file name sorts after every other test file), so a run that stops at
the first failure has already run everything else.  It also passes alone in a fresh process.
"""
import ctypes as C

import numpy as np
import pytest

import libsplinter_b200 as sg
import support

pytestmark = pytest.mark.gpu


def _dev_ptr(t):
    return C.c_void_p(t.data_ptr())


def _call(lib, cols, n, row0, p, d_q, d_out, stride, d_cnt, stream=0):
    dE, dbloom, dflags, depoch = cols
    return lib.spl_gpu_scan_noalloc(_dev_ptr(dE), _dev_ptr(dbloom), _dev_ptr(dflags), _dev_ptr(depoch), n, row0,
                                    C.byref(p), _dev_ptr(d_q), _dev_ptr(d_out), stride, _dev_ptr(d_cnt),
                                    C.c_void_p(stream))


# k = 10 ends in the register / shuffle epilogue; k = 64 in the block-wide sort with the anchor selection and uses the
# whole fixed scratch (64 candidates per CTA); 20 rows leave all but one CTA without any work
@pytest.mark.parametrize("n,k", [(40_000, 10), (40_000, 64), (20, 10)])
def test_noalloc_scan_matches_the_allocating_entry_point_twice_in_a_row(n, k):
    import torch
    lib = sg.load()
    dim, row0 = 768, 1000
    E, bloom, flags = support.synth_dense(1234, 99, row0, n, dim)
    epoch = np.where(flags != 0, 4, 0).astype(np.uint64)
    q = support.synth_queries(4321, 3, dim)

    # reference results from the existing (allocating) entry point
    with sg.Mirror.open_dense(n, dim, global_row0=row0) as m:
        m.dense_upload(0, E, bloom, flags, epoch)
        want = {mask: m.search(q, k, bloom_mask=mask, mode=sg.MODE_FAST) for mask in (0, 0x3)}

    # the caller's own device buffers: nothing below is allocated by the library
    dev = torch.device("cuda", 0)
    cols = (torch.from_numpy(E).to(dev), torch.from_numpy(bloom.view(np.int64)).to(dev),
            torch.from_numpy(flags).to(dev), torch.from_numpy(epoch.view(np.int64)).to(dev))
    d_q = torch.from_numpy(q).to(dev)
    d_out = torch.zeros((3, k, 32), dtype=torch.uint8, device=dev)
    d_cnt = torch.zeros(3, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()

    for call in range(2):          # second call: scratch still holds the first call's contents
        for mask in (0, 0x3):
            d_out.zero_()
            d_cnt.zero_()
            torch.cuda.synchronize()
            p = sg.SearchParams(C.sizeof(sg.SearchParams), 3, dim, k, sg.ORDER_COSINE, sg.MODE_FAST, mask, 0.0, 0.0,
                                sg.NO_SLOT, 0)
            free_before = torch.cuda.mem_get_info()[0]
            rc = _call(lib, cols, n, row0, p, d_q, d_out, k, d_cnt)
            assert rc == 0, lib.spl_gpu_last_error().decode()
            torch.cuda.synchronize()
            if call == 1:  # past the one-time set-up: device memory in use must not move at all
                assert torch.cuda.mem_get_info()[0] == free_before
            got = d_out.cpu().numpy().view(sg.RESULT_DTYPE).reshape(3, k)
            cnt = d_cnt.cpu().numpy()
            ref_out, ref_cnt = want[mask]
            assert np.array_equal(cnt, ref_cnt.astype(np.int32))
            for i in range(3):
                g, r = got[i][:cnt[i]], ref_out[i][:ref_cnt[i]]
                # same kernel, same order of operations: same hits; scores within the suite's fp32 tolerance
                assert np.array_equal(g["slot_idx"], r["slot_idx"])
                assert np.all(np.abs(g["similarity"] - r["similarity"]) <= support.TOL_SIM)
                assert np.all(np.abs(g["distance"] - r["distance"]) <= support.TOL_DIST_REL * r["distance"] + 1e-30)
                assert np.array_equal(g["bloom"], r["bloom"]) and np.array_equal(g["epoch"], r["epoch"])
            # and against the CPU oracle, like every other scan test
            oref, orc = support.oracle_search_dense(q, E, bloom, flags, k=k, bloom_mask=mask)
            for i in range(3):
                g = got[i][:cnt[i]].copy()
                g["slot_idx"] -= row0
                support.assert_topk_matches(g, oref[i][:orc[i]])


def test_noalloc_scan_refuses_what_its_fixed_scratch_cannot_hold():
    import torch
    lib = sg.load()
    dev = torch.device("cuda", 0)
    n, dim = 256, 768
    cols = (torch.zeros((n, dim), device=dev), torch.zeros(n, dtype=torch.int64, device=dev),
            torch.zeros(n, dtype=torch.uint8, device=dev), torch.zeros(n, dtype=torch.int64, device=dev))
    d_q = torch.zeros(dim, device=dev)
    d_out = torch.zeros((128, 32), dtype=torch.uint8, device=dev)
    d_cnt = torch.zeros(1, dtype=torch.int32, device=dev)
    for k, dim_arg, msg in ((65, dim, "outside 1..64"), (0, dim, "outside 1..64"), (10, 96, "not served")):
        p = sg.SearchParams(C.sizeof(sg.SearchParams), 1, dim_arg, k, sg.ORDER_COSINE, sg.MODE_AUTO, 0, 0.0, 0.0,
                            sg.NO_SLOT, 0)
        rc = _call(lib, cols, n, 0, p, d_q, d_out, 128, d_cnt)
        assert rc == -2
        assert msg in lib.spl_gpu_last_error().decode()
    # a sub-allocation of the caller's arena that is not aligned for the row copies is refused, not faulted on
    p = sg.SearchParams(C.sizeof(sg.SearchParams), 1, dim, 10, sg.ORDER_COSINE, sg.MODE_AUTO, 0, 0.0, 0.0, sg.NO_SLOT, 0)
    skewed = (cols[0].view(-1)[1:1 + (n - 1) * dim].view(n - 1, dim),) + cols[1:]
    rc = _call(lib, skewed, n - 1, 0, p, d_q, d_out, 128, d_cnt)
    assert rc == -2 and "misaligned" in lib.spl_gpu_last_error().decode()
    torch.cuda.synchronize()
