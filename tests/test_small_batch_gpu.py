"""Small-batch scans: 2, 4 or 8 queries share one pass over the store (spl_scan_tma.cu, NQ > 1), so each row
fetched from HBM is scored for all of them.  Same contract as the single-query scan; the per-row
arithmetic is the same instruction sequence, so a batch must return bit-for-bit what the same queries
return one at a time, and both must match the CPU oracle."""
import numpy as np
import pytest

import libsplinter_b200 as sg
import support

pytestmark = pytest.mark.gpu


def _mirror(n, dim, seed=1234):
    E, bloom, flags = support.synth_dense(seed, 99, 0, n, dim)
    m = sg.Mirror.open_dense(n, dim)
    m.dense_upload(0, E, bloom, flags)
    return m, E, bloom, flags


@pytest.fixture(scope="module")
def corpus():
    m, E, bloom, flags = _mirror(40_003, 768)
    yield m, E, bloom, flags, support.synth_queries(777, 16, 768)
    m.close()


@pytest.mark.parametrize("nq", [2, 3, 4, 5, 7, 8, 9, 13, 16])
@pytest.mark.parametrize("k", [1, 10, 32, 33, 100, 256])
def test_batch_equals_one_at_a_time_and_the_oracle(corpus, nq, k):
    m, E, bloom, flags, q = corpus
    out, counts = m.search(q[:nq], k, mode=sg.MODE_FAST, bloom_mask=0x1)
    ref, rc = support.oracle_search_dense(q[:nq], E, bloom, flags, k=k, bloom_mask=0x1)
    totals = m.last_totals(nq)
    for i in range(nq):
        one, c1 = m.search(q[i], k, mode=sg.MODE_FAST, bloom_mask=0x1)
        assert counts[i] == c1[0]
        assert out[i][:counts[i]].tobytes() == one[0][:c1[0]].tobytes()
        support.assert_topk_matches(out[i][:counts[i]], ref[i][:rc[i]])
        assert totals[i] == m.last_totals(1)[0]


@pytest.mark.parametrize("order", [sg.ORDER_COSINE, sg.ORDER_L2])
def test_thresholds_exclude_and_orders(corpus, order):
    m, E, bloom, flags, q = corpus
    kw = dict(order=order, bloom_mask=0x2, min_similarity=0.02, max_distance=1.41, exclude_slot=17)
    out, counts = m.search(q[:6], 16, mode=sg.MODE_FAST, **kw)
    ref, rc = support.oracle_search_dense(q[:6], E, bloom, flags, k=16, order=order, bloom_mask=0x2, min_sim=0.02,
                                          max_dist=1.41, exclude=17)
    full, fc = support.oracle_search_dense(q[:6], E, bloom, flags, k=0, order=order, bloom_mask=0x2, min_sim=0.02,
                                           max_dist=1.41, exclude=17)
    totals = m.last_totals(6)
    for i in range(6):
        support.assert_topk_matches(out[i][:counts[i]], ref[i][:rc[i]], order=order)
        assert abs(int(totals[i]) - int(fc[i])) <= 2      # rows sitting within an ulp of a threshold


def test_exact_mode_rides_the_same_batched_scan(corpus):
    m, E, bloom, flags, q = corpus
    out, counts = m.search(q[:5], 10, mode=sg.MODE_EXACT, bloom_mask=0x3)
    ref, rc = support.oracle_search_dense(q[:5], E, bloom, flags, k=10, bloom_mask=0x3)
    for i in range(5):
        support.assert_topk_matches(out[i][:counts[i]], ref[i][:rc[i]], exact=True)


@pytest.mark.parametrize("dim", [1024, 512, 256, 128])
@pytest.mark.parametrize("nq", [4, 8])
def test_other_dims_and_no_candidates(dim, nq):
    """(eight per pass exists for rows of up to 768 floats; 1024-d batches of eight go four at a time)"""
    m, E, bloom, flags = _mirror(9_001, dim, seed=5)
    q = support.synth_queries(11, nq, dim)
    for mask in (0, 1 << 63, (1 << 64) - 1):
        out, counts = m.search(q, 10, mode=sg.MODE_FAST, bloom_mask=mask)
        ref, rc = support.oracle_search_dense(q, E, bloom, flags, k=10, bloom_mask=mask)
        for i in range(nq):
            assert counts[i] == rc[i]
            support.assert_topk_matches(out[i][:counts[i]], ref[i][:rc[i]])
    m.close()


def test_identical_queries_in_one_batch_give_identical_lists(corpus):
    m, E, bloom, flags, q = corpus
    qq = np.repeat(q[:1], 4, axis=0)
    out, counts = m.search(qq, 10, mode=sg.MODE_FAST)
    for i in range(1, 4):
        assert counts[i] == counts[0] and out[i].tobytes() == out[0].tobytes()


@pytest.mark.parametrize("nq", [1, 2, 4, 8])
@pytest.mark.parametrize("k", [10, 24, 32])
def test_ties_on_the_ordering_score_are_broken_by_slot_index(nq, k):
    """A third of the rows are zero vectors (similarity 0, distance |q|: the whole 64-bit key ties), another block are
    copies of one row; every other row scores below zero.  The k best are then all ties, spread over every warp's and
    every CTA's list -- the per-CTA selection by counting and the admission bounds must keep exactly the lowest slot
    indices, for each query of a batch as for a single one."""
    n, dim = 30_011, 768
    rng = np.random.default_rng(31)
    q = rng.standard_normal((8, dim)).astype(np.float32)
    q[1:] = q[0] * np.array([[0.5], [2.0], [1.0], [0.25], [4.0], [1.5], [3.0]], np.float32)   # same direction: one corpus suits all
    E = (-np.abs(rng.standard_normal((n, dim))) * np.sign(q[0])).astype(np.float32)
    E[::3] = 0.0
    bloom = np.zeros(n, np.uint64)
    flags = np.full(n, 7, np.uint8)
    with sg.Mirror.open_dense(n, dim) as m:
        m.dense_upload(0, E, bloom, flags)
        out, counts = m.search(q[:nq], k, mode=sg.MODE_FAST)
        ref, rc = support.oracle_search_dense(q[:nq], E, bloom, flags, k=k)
        for i in range(nq):
            assert counts[i] == rc[i] == k
            assert np.array_equal(out[i][:k]["slot_idx"], np.arange(0, 3 * k, 3)), (i, out[i][:k]["slot_idx"])
            support.assert_topk_matches(out[i][:k], ref[i][:k])
    # copies of the best row: identical (similarity, distance), ranked by slot index
    E2 = rng.standard_normal((n, dim)).astype(np.float32)
    best = (q[0] / np.linalg.norm(q[0])).astype(np.float32)
    where = np.sort(rng.choice(n, 3 * k, replace=False))
    E2[where] = best
    with sg.Mirror.open_dense(n, dim) as m:
        m.dense_upload(0, E2, bloom, flags)
        out, counts = m.search(q[:nq], k, mode=sg.MODE_FAST)
        for i in range(nq):
            assert counts[i] == k
            assert np.array_equal(out[i][:k]["slot_idx"], where[:k]), (i, out[i][:k]["slot_idx"], where[:k])
            assert len(set(out[i][:k]["similarity"].tolist())) == 1


def test_eight_per_pass_with_a_similarity_threshold_and_exclusion(corpus):
    """Cosine order without a distance threshold is what the 8-queries-per-pass kernel takes (11 queries = 8 + 2 + 1);
    a similarity threshold cuts inside the candidates and one slot is excluded."""
    m, E, bloom, flags, q = corpus
    kw = dict(bloom_mask=0x2, min_similarity=0.02, exclude_slot=17)
    out, counts = m.search(q[:11], 16, mode=sg.MODE_FAST, **kw)
    ref, rc = support.oracle_search_dense(q[:11], E, bloom, flags, k=16, bloom_mask=0x2, min_sim=0.02, exclude=17)
    full, fc = support.oracle_search_dense(q[:11], E, bloom, flags, k=0, bloom_mask=0x2, min_sim=0.02, exclude=17)
    totals = m.last_totals(11)
    for i in range(11):
        one, c1 = m.search(q[i], 16, mode=sg.MODE_FAST, **kw)
        assert out[i][:counts[i]].tobytes() == one[0][:c1[0]].tobytes()
        support.assert_topk_matches(out[i][:counts[i]], ref[i][:rc[i]])
        assert abs(int(totals[i]) - int(fc[i])) <= 2      # rows sitting within an ulp of the threshold
