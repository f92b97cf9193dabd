"""Golden vectors generated from the compiled reference (tests/golden/make_golden.py) -- usable on
machines where neither /root/reference nor oracle/_ref exists.

CPU: the restatement reproduces them bit-for-bit (store image, scorers, candidate counts, order).
GPU: the CUDA path reproduces them through the C ABI on the reference-format store image (C1).
"""
from pathlib import Path

import numpy as np
import pytest

import support

GOLD = np.load(Path(__file__).resolve().parent / "golden" / "search_golden.npz")
DIM, SLOTS, NKEYS, MAXVAL = 768, 4096, 4090, 64
SEED_C, SEED_B, SEED_Q = 1234, 99, 4321


@pytest.fixture(scope="module")
def c1_image():
    """The C1 store rebuilt with the restatement's splinter_create/_set/_set_embedding/_set_label."""
    img = support.new_image(SLOTS, MAXVAL, DIM)
    row = np.empty(DIM, np.float32)
    ora = support.oracle()
    for k in range(NKEYS):
        cls = ora.splo_synth_class(SEED_C, k)
        if cls == 0:
            continue
        key = f"doc{k:08d}"
        assert support.image_set(img, key, f"v{k:08d}".encode()) >= 0
        if cls == 2:
            ora.splo_synth_row(SEED_C, k, DIM, support._p(row))
            assert support.image_set_embedding(img, key, row) == 0
        assert support.image_set_label(img, key, ora.splo_synth_bloom(SEED_B, k)) == 0
    return img


def test_store_image_equals_the_reference_built_store(c1_image):
    assert np.bitwise_xor.reduce(c1_image.view(np.uint64)) == GOLD["c1_image_crc"][0]


def test_bulk_image_builder_equals_the_call_by_call_store(c1_image):
    """splo_image_fill_synthetic (used for the large bench / stress stores) = the same store."""
    img = support.new_image(SLOTS, MAXVAL, DIM)
    placed = support.image_fill_synthetic(img, NKEYS, SEED_C, SEED_B)
    assert placed > 4000
    assert np.array_equal(img, c1_image)


def test_scorer_known_answers():
    n = 96
    a, b = support.synth_queries(7, n, DIM), support.synth_queries(8, n, DIM)
    b[5] = 0.0
    b[6] = a[6]
    a[7] *= 1e-6
    ora = support.oracle()
    cos = np.array([ora.splo_cosine(support._p(a[i]), support._p(b[i]), DIM) for i in range(n)], np.float32)
    dist = np.array([ora.splo_euclid(support._p(a[i]), support._p(b[i]), DIM) for i in range(n)], np.float32)
    assert np.array_equal(cos.view(np.uint32), GOLD["scorer_cos_bits"])
    assert np.array_equal(dist.view(np.uint32), GOLD["scorer_dist_bits"])
    assert cos[5] == 0.0 and dist[6] == 0.0


def test_restatement_reproduces_c1(c1_image):
    q = support.synth_queries(SEED_Q, 4, DIM)
    for mi, mask in enumerate(GOLD["c1_masks"]):
        for qi in range(4):
            got, surv = support.oracle_search_image(c1_image, q[qi], 10, bloom_mask=int(mask))
            assert surv == GOLD["c1_ncand"][mi]
            n = min(10, surv)  # a selective mask (0xF) leaves fewer than 10 candidates
            assert len(got) == n
            assert np.array_equal(got["idx"], GOLD["c1_idx"][mi, qi][:n])
            assert np.array_equal(got["similarity"].view(np.uint32), GOLD["c1_sim_bits"][mi, qi][:n])
            assert np.array_equal(got["distance"].view(np.uint32), GOLD["c1_dist_bits"][mi, qi][:n])


def test_cli_output_golden(c1_image):
    """keys and order of the reference's real `search --json --limit 10 --bloom 3` output"""
    import json
    doc = json.loads(bytes(GOLD["cli_json"]).decode())
    q = support.synth_queries(SEED_Q, 1, DIM)[0]
    got, surv = support.oracle_search_image(c1_image, q, 10, bloom_mask=3)
    assert doc["store"]["total_slots"] == SLOTS and doc["store"]["active_keys"] == surv
    slot_size = 128 + 4 * DIM
    for row, g in zip(doc["results"], got):
        off = support.HEADER_SIZE + int(g["idx"]) * slot_size + 64
        assert row["key"] == bytes(c1_image[off:off + 64]).split(b"\0")[0].decode()
        assert row["similarity"] == pytest.approx(float(g["similarity"]), abs=5.1e-5)
        assert row["has_embedding"] is True


@pytest.mark.gpu
def test_gpu_reproduces_c1_golden(c1_image):
    import libsplinter_b200 as sg
    q = support.synth_queries(SEED_Q, 4, DIM)
    with sg.Mirror.attach_image(c1_image) as m:
        for mode in (sg.MODE_FAST, sg.MODE_EXACT):
            for mi, mask in enumerate(GOLD["c1_masks"]):
                try:
                    out, counts = m.search(q, 10, bloom_mask=int(mask), mode=mode)
                except sg.SplinterGpuError as e:
                    if "not available" in str(e):
                        continue
                    raise
                for qi in range(4):
                    hits = out[qi][:counts[qi]]
                    n = min(10, int(GOLD["c1_ncand"][mi]))
                    ref = np.zeros(n, support.SPLO_RESULT)
                    ref["idx"] = GOLD["c1_idx"][mi, qi][:n]
                    ref["similarity"] = GOLD["c1_sim_bits"][mi, qi].view(np.float32)[:n]
                    ref["distance"] = GOLD["c1_dist_bits"][mi, qi].view(np.float32)[:n]
                    support.assert_topk_matches(hits, ref, exact=(mode == sg.MODE_EXACT))
