"""Seeded random sweep over the search surface (store size, dimension, k, order, mask, thresholds, exclusion,
batch size, mode), each case checked against the CPU oracle.  The reference has no test for this path
(SURVEY 4); hand-picked cases cannot cover the cross product, and an earlier sweep of this kind is what
exposed the ring-rotation bug fixed in spl_scan_tma.cu."""
import numpy as np
import pytest

import libsplinter_b200 as sg
import support

pytestmark = pytest.mark.gpu

DIMS = [128, 256, 512, 768, 1024, 96, 320]          # 96 / 320: no TMA layout, register-path kernel


def _cases(seed, count):
    rng = np.random.default_rng(seed)
    for c in range(count):
        dim = int(rng.choice(DIMS))
        n = int(rng.choice([1, 31, 33, 700, 4096, 20_011, 65_537]))
        k = int(rng.choice([1, 2, 7, 10, 16, 31, 32, 33, 64, 100, 256, 0]))
        nq = int(rng.choice([1, 1, 2, 3, 4, 5, 8, 9]))
        order = int(rng.choice([sg.ORDER_COSINE, sg.ORDER_L2]))
        mask = int(rng.choice([0, 0x1, 0x3, 0x10, 0xF0, 1 << 63, (1 << 63) | 0x2, 0xFFFFFFFFFFFFFFFF]))
        thr = int(rng.integers(0, 4))                # 0 none, 1 min_sim, 2 max_dist, 3 both
        mode = int(rng.choice([sg.MODE_FAST, sg.MODE_FAST, sg.MODE_EXACT, sg.MODE_AUTO]))
        excl = bool(rng.integers(0, 3) == 0)
        yield c, dim, n, k, nq, order, mask, thr, mode, excl, int(rng.integers(1, 1 << 30))


@pytest.mark.parametrize("seed", [20261, 20262, 20263, 20264, 20265, 20266])
def test_random_searches_match_the_oracle(seed):
    mirrors = {}
    try:
        for c, dim, n, k, nq, order, mask, thr, mode, excl, s in _cases(seed, 60):
            if k == 0 and n > 20_011:
                k = 10                                # unlimited result sets stay small here
            key = (dim, n)
            if key not in mirrors:
                E, bloom, flags = support.synth_dense(1000 + dim, 77, 0, n, dim)
                m = sg.Mirror.open_dense(n, dim)
                m.dense_upload(0, E, bloom, flags)
                mirrors[key] = (m, E, bloom, flags)
            m, E, bloom, flags = mirrors[key]
            q = support.synth_queries(s, nq, dim)
            kw, okw = dict(order=order, bloom_mask=mask), dict(order=order, bloom_mask=mask)
            # thresholds near the bulk of the score distribution so that they cut somewhere inside the candidates
            if thr & 1:
                kw["min_similarity"] = okw["min_sim"] = float(1.5 / np.sqrt(dim))
            if thr & 2:
                kw["max_distance"] = okw["max_dist"] = float(np.sqrt(2.0 - 2.0 / np.sqrt(dim)))
            if excl:
                kw["exclude_slot"] = okw["exclude"] = int(s % n)
            what = f"case {c}: dim={dim} n={n} k={k} nq={nq} order={order} mask={mask:#x} thr={thr} mode={mode} excl={excl}"
            out, counts = m.search(q, k, mode=mode, **kw)
            ref, rc = support.oracle_search_dense(q, E, bloom, flags, k=k, **okw)
            for i in range(nq):
                if thr:
                    # a row within an ulp of a threshold may sit on either side in FAST mode
                    assert abs(int(counts[i]) - int(rc[i])) <= (0 if mode == sg.MODE_EXACT else 2), what
                    ncmp = min(int(counts[i]), int(rc[i]))
                    if mode != sg.MODE_EXACT and ncmp and counts[i] != rc[i]:
                        continue                      # membership differs by the borderline row: order check not meaningful
                else:
                    assert counts[i] == rc[i], what
                    ncmp = int(counts[i])
                try:
                    support.assert_topk_matches(out[i][:ncmp], ref[i][:ncmp], exact=(mode == sg.MODE_EXACT), order=order)
                except AssertionError as e:
                    raise AssertionError(f"{what} query {i}: {e}") from e
    finally:
        for m, *_ in mirrors.values():
            m.close()


@pytest.mark.parametrize("seed", [31, 32, 33, 34])
def test_random_batched_searches_match_the_oracle(seed):
    """SPL_MODE_BATCH_BF16 over random shapes: bf16 only selects the 2k rows that are re-scored in fp32, so the
    results must meet the fp32 tolerance and name the same rows as the oracle (away from near-ties)."""
    rng = np.random.default_rng(seed)
    for c in range(12):
        dim = int(rng.choice([64, 128, 512, 768, 1024]))
        n = int(rng.choice([257, 5_000, 33_333, 70_001]))
        k = int(rng.choice([1, 5, 10, 32, 100, 128]))
        nq = int(rng.choice([1, 17, 128, 129, 300]))
        order = int(rng.choice([sg.ORDER_COSINE, sg.ORDER_L2]))
        mask = int(rng.choice([0, 0x1, 0x3, 1 << 63]))
        E, bloom, flags = support.synth_dense(500 + c, 78, 0, n, dim)
        flags = flags.copy()
        flags[~E.any(axis=1)] = 0          # zero vectors tie exactly at distance |q|: keep them out (see test_batch_gpu)
        m = sg.Mirror.open_dense(n, dim, flags=sg.OPT_BF16_MIRROR)
        try:
            m.dense_upload(0, E, bloom, flags)
            q = support.synth_queries(int(rng.integers(1, 1 << 30)), nq, dim)
            out, counts = m.search(q, k, mode=sg.MODE_BATCH_BF16, order=order, bloom_mask=mask)
            ref, rc = support.oracle_search_dense(q, E, bloom, flags, k=k, order=order, bloom_mask=mask)
            what = f"case {c}: dim={dim} n={n} k={k} nq={nq} order={order} mask={mask:#x}"
            for i in range(nq):
                assert counts[i] == rc[i], what
                try:
                    support.assert_topk_matches(out[i][:counts[i]], ref[i][:rc[i]], order=order)
                except AssertionError as e:
                    raise AssertionError(f"{what} query {i}: {e}") from e
        finally:
            m.close()
