"""GPU parity tests of the single-query scan (through the C ABI) against the CPU oracle.

The reference has no test for this path; these follow its behaviour as restated in
oracle/spl_oracle.c, which is itself pinned to the compiled reference by test_oracle_pinned.py.
"""
import numpy as np
import pytest

import libsplinter_b200 as sg
import support

pytestmark = pytest.mark.gpu

MODES = [sg.MODE_FAST, sg.MODE_EXACT]


@pytest.fixture(scope="module")
def corpus():
    n, dim = 50_000, 768  # not a multiple of 32: ragged last group
    E, bloom, flags = support.synth_dense(1234, 99, 0, n, dim)
    q = support.synth_queries(4321, 4, dim)
    m = sg.Mirror.open_dense(n, dim)
    m.dense_upload(0, E, bloom, flags)
    yield m, E, bloom, flags, q
    m.close()


def _check(m, E, bloom, flags, q, k, mode, **kw):
    okw = dict(order=kw.get("order", sg.ORDER_COSINE), bloom_mask=kw.get("bloom_mask", 0),
               min_sim=kw.get("min_similarity", 0.0), max_dist=kw.get("max_distance", 0.0),
               exclude=kw.get("exclude_slot", sg.NO_SLOT))
    out, counts = m.search(q, k, mode=mode, **kw)
    ref, rc = support.oracle_search_dense(q, E, bloom, flags, k=k, **okw)
    for i in range(q.shape[0]):
        if k == 0:
            assert counts[i] == rc[i]
        support.assert_topk_matches(out[i][:min(counts[i], out.shape[1])], ref[i][:rc[i]],
                                    exact=(mode == sg.MODE_EXACT), order=okw["order"])
    return out, counts


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("mask", [0, 0x3, 1 << 63, 0xF, 0xFFFF])
def test_top10_cosine_masks(corpus, mode, mask):
    m, E, bloom, flags, q = corpus
    out, counts = _check(m, E, bloom, flags, q, 10, mode, bloom_mask=mask)
    # candidate sets are bit-exact: every hit carries the mask, epoch/bloom come from the mirror
    for i in range(q.shape[0]):
        hits = out[i][:counts[i]]
        assert np.all((hits["bloom"] & np.uint64(mask)) == np.uint64(mask))
        assert np.array_equal(hits["bloom"], bloom[hits["slot_idx"]])


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("k", [1, 32, 100, 256])
def test_k_sweep(corpus, mode, k):
    m, E, bloom, flags, q = corpus
    _check(m, E, bloom, flags, q[:2], k, mode, bloom_mask=0x1)


@pytest.mark.parametrize("mode", MODES)
def test_l2_order(corpus, mode):
    m, E, bloom, flags, q = corpus
    _check(m, E, bloom, flags, q[:2], 32, mode, order=sg.ORDER_L2)


@pytest.mark.parametrize("mode", MODES)
def test_thresholds_and_exclude(corpus, mode):
    m, E, bloom, flags, q = corpus
    base, _ = m.search(q[:1], 10, mode=mode)
    top = int(base[0][0]["slot_idx"])
    out, counts = _check(m, E, bloom, flags, q[:1], 10, mode, exclude_slot=top)
    assert top not in out[0]["slot_idx"][:counts[0]]
    _check(m, E, bloom, flags, q[:1], 50, mode, min_similarity=0.08)
    _check(m, E, bloom, flags, q[:1], 50, mode, max_distance=1.36)
    _check(m, E, bloom, flags, q[:1], 50, mode, min_similarity=0.05, max_distance=1.37, bloom_mask=0x1)


@pytest.mark.parametrize("mode", MODES)
def test_unlimited_is_the_reference_default(corpus, mode):
    """--limit absent prints every survivor, sorted (_search.c:415-416)."""
    m, E, bloom, flags, q = corpus
    out, counts = _check(m, E, bloom, flags, q[:1], 0, mode, bloom_mask=0x7)
    assert counts[0] == int(np.sum(((bloom & np.uint64(7)) == 7) & ((flags & 2) != 0)))
    _check(m, E, bloom, flags, q[:1], 1000, mode, bloom_mask=0x1)  # k beyond the fused limit


@pytest.mark.parametrize("mode", MODES)
def test_zero_vectors_are_results(mode):
    """Never-embedded slots score sim=0 (denominator guard, _search.c:51) and outrank negatives."""
    n, dim = 4096, 768
    rng = np.random.default_rng(7)
    E = rng.standard_normal((n, dim)).astype(np.float32)
    q = rng.standard_normal((1, dim)).astype(np.float32)
    E[::3] = 0.0
    E[1::3] = -np.abs(E[1::3]) * np.sign(q)  # strongly negative similarity
    bloom = np.zeros(n, np.uint64)
    flags = np.full(n, 7, np.uint8)
    flags[5] = 3  # live key whose embedding copy never succeeded: sim 0, dist FLT_MAX, has_embedding 0
    with sg.Mirror.open_dense(n, dim) as m:
        m.dense_upload(0, E, bloom, flags)
        out, counts = m.search(q, 0, mode=mode)
        ref, rc = support.oracle_search_dense(q, E, bloom, flags, k=0)
        assert counts[0] == rc[0] == n
        support.assert_topk_matches(out[0][:n], ref[0][:n], exact=(mode == sg.MODE_EXACT))
        row5 = out[0][out[0]["slot_idx"] == 5][0]
        assert row5["has_embedding"] == 0 and row5["similarity"] == 0 and row5["distance"] > 3e38
        # with a threshold active such slots are dropped (_search.c:393)
        out2, c2 = m.search(q, 0, mode=mode, min_similarity=1e-6)
        assert 5 not in out2[0]["slot_idx"][:c2[0]]


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("dim", [1024, 128, 96, 2048])
def test_other_dims(mode, dim):
    n = 5000
    E, bloom, flags = support.synth_dense(5, 6, 100, n, dim)
    q = support.synth_queries(8, 2, dim)
    with sg.Mirror.open_dense(n, dim, global_row0=100) as m:
        m.dense_upload(0, E, bloom, flags)
        out, counts = m.search(q, 10, mode=mode, bloom_mask=1)
        ref, rc = support.oracle_search_dense(q, E, bloom, flags, k=10, bloom_mask=1)
        for i in range(2):
            got = out[i][:counts[i]].copy()
            assert np.all(got["slot_idx"] >= 100)
            got["slot_idx"] -= 100  # results carry GLOBAL slot indices
            support.assert_topk_matches(got, ref[i][:rc[i]], exact=(mode == sg.MODE_EXACT))


def test_empty_and_tiny():
    dim = 768
    E = np.zeros((3, dim), np.float32)
    with sg.Mirror.open_dense(3, dim) as m:
        m.dense_upload(0, E, np.zeros(3, np.uint64), np.zeros(3, np.uint8))  # no live slot at all
        for mode in MODES:
            out, counts = m.search(np.ones((1, dim), np.float32), 10, mode=mode)
            assert counts[0] == 0
            out, counts = m.search(np.ones((1, dim), np.float32), 0, mode=mode, out_stride=8)
            assert counts[0] == 0


def test_device_synthetic_generator_is_bit_identical():
    n, dim, row0 = 3000, 768, 12345
    E, bloom, flags = support.synth_dense(1234, 99, row0, n, dim)
    with sg.Mirror.open_dense(n, dim, global_row0=row0) as m:
        m.fill_synthetic(1234, 99)
        dE, dbloom, dflags, _ = m.read_mirror(row0, n)
    assert np.array_equal(dbloom, bloom)
    assert np.array_equal(dflags, flags)
    assert np.array_equal(dE.view(np.uint32), E.view(np.uint32))


def test_bad_arguments_fail_loudly(corpus):
    m = corpus[0]
    with pytest.raises(sg.SplinterGpuError):
        m.search(np.zeros((1, 512), np.float32), 10)  # wrong dim
    p = m.params(1, 10)
    p.order = 9
    q = np.zeros((1, 768), np.float32)
    out = np.zeros(10, sg.RESULT_DTYPE)
    cnt = np.zeros(1, np.uint32)
    import ctypes as C
    rc = sg.load().spl_gpu_search(m.handle, C.byref(p), q.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p),
                                  10, cnt.ctypes.data_as(C.c_void_p))
    assert rc == -2


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("dim,k", [(1024, 10), (1024, 32), (768, 256), (512, 100)])
def test_rings_whose_slots_rotate_between_consumer_warps(mode, dim, k):
    """Regression: when the shared-memory ring size is not a multiple of the consumer-warp count a slot's
    successive rounds are drained by different warps; an early warp once mistook the 1-bit mbarrier phase
    of the PREVIOUS round for its own (hang / fault on dense dim=1024 scans).  Dense, repeated."""
    n = 30_000
    E, bloom, flags = support.synth_dense(77, 78, 0, n, dim)
    q = support.synth_queries(79, 6, dim)
    with sg.Mirror.open_dense(n, dim) as m:
        m.dense_upload(0, E, bloom, flags)
        ref, rc = support.oracle_search_dense(q, E, bloom, flags, k=k)
        for rep in range(4):
            out, counts = m.search(q, k, mode=mode)
            for i in range(q.shape[0]):
                support.assert_topk_matches(out[i][:counts[i]], ref[i][:rc[i]], exact=(mode == sg.MODE_EXACT))


@pytest.mark.parametrize("k", [10, 32])
def test_best_rows_clustered_in_one_chunk(k):
    """The last CTA first tries a selection bounded by the k-th best list head; when every good row sits in one
    CTA's range that bound is loose, more candidates survive than the fast path takes, and the general merge
    fold must give the same answer.  Also a batch of four (each query with its own cluster)."""
    n, dim = 60_000, 768
    E, bloom, flags = support.synth_dense(77, 99, 0, n, dim)
    q = support.synth_queries(78, 4, dim)
    rng = np.random.default_rng(79)
    for qi in range(4):      # 100 consecutive rows close to query qi: one 256-row chunk holds the whole top-k
        base = 5_000 + 10_000 * qi
        noise = rng.standard_normal((100, dim)).astype(np.float32) * np.float32(0.02)
        E[base:base + 100] = q[qi][None, :] + noise
        flags[base:base + 100] = sg.ROW_LIST | sg.ROW_ENUM | sg.ROW_EMB
    m = sg.Mirror.open_dense(n, dim)
    try:
        m.dense_upload(0, E, bloom, flags)
        for mode in MODES:
            out, counts = _check(m, E, bloom, flags, q, k, mode)
            for qi in range(4):
                base = 5_000 + 10_000 * qi
                assert np.all((out[qi][:k]["slot_idx"] >= base) & (out[qi][:k]["slot_idx"] < base + 100))
            _check(m, E, bloom, flags, q[:1], k, mode, order=sg.ORDER_L2)
    finally:
        m.close()
