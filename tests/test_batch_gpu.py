"""GPU parity tests of the batched tensor-core path (SPL_MODE_BATCH_BF16, config C3) against the fp32
CPU oracle.  bf16 is used only to pick which 2k rows get re-scored; the scores returned are fp32 and
must meet the same tolerance as the single-query scan.  The oracle for this config is the restated
loop over queries (D=1024 / L2 order / batching are not expressible by the shipped CLI)."""
import numpy as np
import pytest

import libsplinter_b200 as sg
import support

pytestmark = pytest.mark.gpu


def _mirror(n, dim, row0=0, seed=1234):
    E, bloom, flags = support.synth_dense(seed, 99, row0, n, dim)
    m = sg.Mirror.open_dense(n, dim, global_row0=row0, flags=sg.OPT_BF16_MIRROR)
    m.dense_upload(0, E, bloom, flags)
    return m, E, bloom, flags


def _check(m, E, bloom, flags, q, k, row0=0, **kw):
    out, counts = m.search(q, k, mode=sg.MODE_BATCH_BF16, **kw)
    ref, rc = support.oracle_search_dense(q, E, bloom, flags, k=k, order=kw.get("order", sg.ORDER_COSINE),
                                          bloom_mask=kw.get("bloom_mask", 0), min_sim=kw.get("min_similarity", 0.0),
                                          max_dist=kw.get("max_distance", 0.0))
    bad = []
    for i in range(q.shape[0]):
        got = out[i][:counts[i]].copy()
        got["slot_idx"] -= row0
        assert counts[i] == rc[i]
        support.assert_topk_matches(got, ref[i][:rc[i]], exact=False, order=kw.get("order", sg.ORDER_COSINE))
        if not np.array_equal(got["slot_idx"], ref[i][:rc[i]]["idx"]):
            bad.append(i)
    _check.last = (out, counts)
    return bad


@pytest.mark.parametrize("dim", [1024, 768])
@pytest.mark.parametrize("order", [sg.ORDER_L2, sg.ORDER_COSINE])
def test_batched_topk_matches_oracle(dim, order):
    n, nq, k = 70_000, 200, 32       # n, nq deliberately not multiples of the 256 x 128 tile
    m, E, bloom, flags = _mirror(n, dim, row0=512)
    q = support.synth_queries(4321, nq, dim)
    try:
        _check(m, E, bloom, flags, q, k, row0=512, order=order)
        _check(m, E, bloom, flags, q[:40], 10, row0=512, order=order, bloom_mask=0x3)
    finally:
        m.close()


def test_batched_small_shapes_and_auto_mode():
    dim = 1024
    m, E, bloom, flags = _mirror(300, dim)   # less than one tile of rows
    q = support.synth_queries(7, 3, dim)
    try:
        _check(m, E, bloom, flags, q, 5, order=sg.ORDER_L2)
        _check(m, E, bloom, flags, q[:1], 32, order=sg.ORDER_COSINE)
        # AUTO picks the batched path from 16 queries up when a bf16 mirror exists
        q16 = support.synth_queries(8, 20, dim)
        out, counts = m.search(q16, 10)
        ref, rc = support.oracle_search_dense(q16, E, bloom, flags, k=10)
        for i in range(20):
            support.assert_topk_matches(out[i][:counts[i]], ref[i][:rc[i]])
        # which path AUTO takes, seen through the launch counter (batched = 5 launches per call, scan = one per
        # group of up to 4 queries): unfiltered batches go to the tensor cores from 4 queries, masked ones from 16
        def launches(nq, **kw):
            before = m.launch_count()
            m.search(q16[:nq], 10, **kw)
            return m.launch_count() - before
        assert launches(4) == 5 and launches(3) == 2                     # 3 queries: a 2-query and a 1-query scan
        assert launches(4, bloom_mask=0x1) == 1 and launches(15, bloom_mask=0x1) == 5   # 4+4+4+2+1
        assert launches(16, bloom_mask=0x1) == 5
        assert launches(20, max_distance=1.0) == 5                       # cosine order + distance threshold: 5 scans of 4
    finally:
        m.close()


def test_batched_overflow_falls_back_to_the_fp32_scan():
    """Adversarial corpus: 20 000 exact duplicates tie on every ordering value, so every row beats the
    sampled threshold, the candidate list overflows, and the affected queries are answered by the fp32
    scan instead -- same results (ties broken by ascending slot index), never a CPU fallback."""
    n, dim = 20_000, 1024
    rng = np.random.default_rng(3)
    base = rng.standard_normal(dim).astype(np.float32)
    E = np.tile(base, (n, 1))
    E[5000] = base * 0.5          # a few distinct rows
    E[17] = -base
    bloom = np.zeros(n, np.uint64)
    flags = np.full(n, 7, np.uint8)
    q = (base[None, :] + rng.standard_normal((4, dim)).astype(np.float32) * 0.5).astype(np.float32)
    with sg.Mirror.open_dense(n, dim, flags=sg.OPT_BF16_MIRROR) as m:
        m.dense_upload(0, E, bloom, flags)
        for order in (sg.ORDER_L2, sg.ORDER_COSINE):
            out, counts = m.search(q, 16, mode=sg.MODE_BATCH_BF16, order=order)
            ref, rc = support.oracle_search_dense(q, E, bloom, flags, k=16, order=order)
            for i in range(4):
                support.assert_topk_matches(out[i][:counts[i]], ref[i][:rc[i]], order=order)
                assert 17 not in out[i]["slot_idx"][:counts[i]]


def test_batched_mode_needs_the_bf16_mirror():
    with sg.Mirror.open_dense(1000, 1024) as m:
        with pytest.raises(sg.SplinterGpuError, match="not available"):
            m.search(np.zeros((4, 1024), np.float32), 10, mode=sg.MODE_BATCH_BF16)


def test_batched_threshold_on_the_ordering_score():
    """--similarity with cosine order / --distance with L2 order cut the tail of the sorted result, so the
    batched path can honour them; a threshold on the other score needs every row's exact value and is refused."""
    n, dim = 40_000, 768
    m, E, bloom, flags = _mirror(n, dim)
    # the synthetic corpus' zero vectors all sit at distance |q| = 1 from every unit query: with them in play an L2
    # threshold would fall on hundreds of exact ties.  Take them out of the candidate set for this test.
    flags = flags.copy()
    flags[~E.any(axis=1)] = 0
    m.dense_upload(0, E, bloom, flags)
    q = support.synth_queries(99, 24, dim)
    try:
        # thresholds chosen so that some queries keep all k, some only a few, some none
        full, _ = support.oracle_search_dense(q, E, bloom, flags, k=16)
        cut = float(np.median(full[:, 7]["similarity"]))
        _check(m, E, bloom, flags, q, 16, order=sg.ORDER_COSINE, min_similarity=cut)
        fl2, _ = support.oracle_search_dense(q, E, bloom, flags, k=16, order=sg.ORDER_L2)
        cut_d = float(np.median(fl2[:, 7]["distance"]))
        _check(m, E, bloom, flags, q, 16, order=sg.ORDER_L2, max_distance=cut_d)
        _check(m, E, bloom, flags, q, 16, order=sg.ORDER_L2, max_distance=cut_d, bloom_mask=0x1)
        with pytest.raises(sg.SplinterGpuError, match="not available"):
            m.search(q, 16, mode=sg.MODE_BATCH_BF16, order=sg.ORDER_COSINE, max_distance=1.0)
        with pytest.raises(sg.SplinterGpuError, match="not available"):
            m.search(q, 16, mode=sg.MODE_BATCH_BF16, order=sg.ORDER_L2, min_similarity=0.1)
        # AUTO with the cross threshold quietly takes the fp32 scan instead
        out, counts = m.search(q, 16, order=sg.ORDER_COSINE, max_distance=cut_d)
        ref, rc = support.oracle_search_dense(q, E, bloom, flags, k=16, max_dist=cut_d)
        for i in range(q.shape[0]):
            assert abs(int(counts[i]) - int(rc[i])) <= 1      # a row within an ulp of the threshold
            n_cmp = min(counts[i], rc[i])
            support.assert_topk_matches(out[i][:n_cmp], ref[i][:n_cmp])
    finally:
        m.close()


def test_fewer_candidates_than_k():
    """Found by tests/test_fuzz_gpu.py: with fewer sampled tiles than k' the bound is +inf, and rows marked
    "not a candidate" (t = +inf, including the padding past the shard's end) used to tie with it and leak into
    the candidate list.  About a quarter of the 257 rows carry the mask here and k = 128: exactly those rows must come back."""
    n, dim = 257, 768
    m, E, bloom, flags = _mirror(n, dim)
    q = support.synth_queries(5, 20, dim)
    try:
        for order in (sg.ORDER_L2, sg.ORDER_COSINE):
            out, counts = m.search(q, 128, mode=sg.MODE_BATCH_BF16, order=order, bloom_mask=0x3)
            ref, rc = support.oracle_search_dense(q, E, bloom, flags, k=128, order=order, bloom_mask=0x3)
            assert rc[0] < 128
            for i in range(20):
                assert counts[i] == rc[i]
                assert set(out[i][:counts[i]]["slot_idx"].tolist()) == set(ref[i][:rc[i]]["idx"].tolist())
    finally:
        m.close()


def _near_duplicate_corpus(n, dim, ndup, noise, seed):
    """n random unit rows, ndup of them replaced by near-copies of one base row: base + noise * |base| * g / sqrt(dim).
    The realistic adversary of an embedding store (re-ingested or lightly edited documents): bf16 cannot tell them apart."""
    rng = np.random.default_rng(seed)
    E = rng.standard_normal((n, dim)).astype(np.float32)
    E /= np.linalg.norm(E, axis=1, keepdims=True)
    base = E[0].copy()
    where = rng.choice(np.arange(1, n), ndup, replace=False)
    E[where] = base[None, :] + (noise / np.sqrt(dim)) * rng.standard_normal((ndup, dim)).astype(np.float32)
    bloom = rng.integers(0, 4, n).astype(np.uint64)
    q = np.concatenate([base[None, :] + 0.2 / np.sqrt(dim) * rng.standard_normal((5, dim)),   # land inside the crowd
                        rng.standard_normal((3, dim)) / np.sqrt(dim)]).astype(np.float32)     # far from it
    return np.ascontiguousarray(E), bloom, q


@pytest.mark.parametrize("dim", [768, 1024])
@pytest.mark.parametrize("ndup,noise", [(3000, 1e-2), (10_000, 1e-3), (3000, 1e-4), (100, 1e-2)])
def test_near_duplicates_do_not_fool_the_bf16_selection(dim, ndup, noise):
    """The bf16 pass only PROPOSES rows; the selection proof in select_rerank_kernel decides whether the proposal can
    be trusted.  With thousands of rows packed inside one bf16 error bound of the k-th score it cannot, and the query
    goes to the fp32 scan; with a hundred the kernel re-scores them all itself.  Either way the answer is
    the reference's: same indices (where fp32 itself separates the rows), both orders, with and without a mask."""
    n, k = 60_000, 32
    E, bloom, q = _near_duplicate_corpus(n, dim, ndup, noise, seed=ndup + dim)
    flags = np.full(n, 7, np.uint8)
    with sg.Mirror.open_dense(n, dim, flags=sg.OPT_BF16_MIRROR) as m:
        m.dense_upload(0, E, bloom, flags)
        for order in (sg.ORDER_COSINE, sg.ORDER_L2):
            for mask in (0, 0x1):
                c0 = m.counters()
                bad = _check(m, E, bloom, flags, q, k, order=order, bloom_mask=mask)   # oracle, ties within tolerance
                c1 = m.counters()
                # the three queries far from the crowd: the reference's indices, except where two rows sit closer together
                # than fp32 summation order can tell apart (1e-6, a twentieth of the suite's tolerance)
                out, counts = _check.last
                ref, rc = support.oracle_search_dense(q, E, bloom, flags, k=k, order=order, bloom_mask=mask)
                for i in [i for i in bad if i >= 5]:
                    g, r = out[i][:counts[i]], ref[i][:rc[i]]
                    prim = r["similarity"] if order == sg.ORDER_COSINE else r["distance"]
                    pos = {int(ix): j for j, ix in enumerate(r["idx"])}
                    for j in np.nonzero(g["slot_idx"] != r["idx"])[0]:
                        other = pos.get(int(g["slot_idx"][j]), len(r) - 1)
                        assert abs(float(prim[j]) - float(prim[other])) <= 1e-6, (i, j, order, mask)
                # inside the crowd fp32 itself cannot separate every pair of rows the way the reference's sequential
                # sums do (that is SPL_MODE_EXACT's job): there the bar is the fp32 scan's own answer, index for index
                fast, fc = m.search(q, k, mode=sg.MODE_FAST, order=order, bloom_mask=mask)
                for i in range(5):
                    assert counts[i] == fc[i]
                    if ndup >= 3000:   # answered by that very scan: the same bytes
                        assert out[i][:counts[i]].tobytes() == fast[i][:fc[i]].tobytes()
                if ndup >= 3000:    # more near-ties than the kernel may re-score: those queries took the fp32 scan
                    assert c1.batch_unproven > c0.batch_unproven
                else:               # a small crowd is re-scored inside the kernel, nothing falls back
                    assert c1.batch_unproven == c0.batch_unproven
                    if mask == 0:   # (with the mask only half of the crowd are candidates: they fit into the fixed 2k)
                        assert c1.batch_rescored_extra > c0.batch_rescored_extra


def test_random_corpus_needs_no_fallback():
    """On the synthetic corpus of the benchmark the proof goes through for every query (a few re-score a handful of
    extra rows): the slow path is for adversarial data, not the common case."""
    n, dim, nq, k = 200_000, 1024, 128, 32
    m, E, bloom, flags = _mirror(n, dim)
    q = support.synth_queries(4321, nq, dim)
    try:
        for order in (sg.ORDER_L2, sg.ORDER_COSINE):
            c0 = m.counters()
            assert not _check(m, E, bloom, flags, q, k, order=order)
            c1 = m.counters()
            assert c1.batch_queries - c0.batch_queries == nq
            assert c1.batch_unproven == c0.batch_unproven
    finally:
        m.close()


@pytest.mark.parametrize("dim", [256, 768, 1024])
def test_pair_kernel_dimensions(dim):
    """256 queries and up run on CTA pairs (cta_group::2) with separate rings for the query and the row operand (3 and
    10 units of 16 KB): 4, 12 and 16 k-blocks per tile exercise every wrap-around of both rings."""
    n, nq, k = 40_000, 256, 10
    m, E, bloom, flags = _mirror(n, dim)
    q = support.synth_queries(909, nq, dim)
    try:
        # (_check asserts scores within tolerance and indices equal up to declared ties; with 256 x 10 hits a couple of
        # pairs closer than fp32 summation order can resolve are expected, so the count of such queries is only bounded)
        assert len(_check(m, E, bloom, flags, q, k, order=sg.ORDER_COSINE)) <= 4
        assert len(_check(m, E, bloom, flags, q[:256], k, order=sg.ORDER_L2, bloom_mask=0x1)) <= 4
        # 300 queries: a pair launch is not possible (3 query tiles), the 1-CTA kernel takes it -- same answers
        q300 = support.synth_queries(910, 300, dim)
        assert len(_check(m, E, bloom, flags, q300, k, order=sg.ORDER_COSINE)) <= 4
    finally:
        m.close()
