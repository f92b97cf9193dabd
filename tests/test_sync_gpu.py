"""GPU tests of the boundary layer: attach (geometry gate), epoch-consistent snapshot, dirty-slot
re-upload, shards.  Store images are produced by the oracle's restatement of splinter_create /
splinter_set / ... (oracle/spl_oracle.c), which test_oracle_pinned.py pins byte-for-byte to the
reference's libsplinter.so."""
import numpy as np
import pytest

import libsplinter_b200 as sg
import support

pytestmark = pytest.mark.gpu

DIM = 768


def _build(slots=1024, nkeys=700, seed=11):
    img = support.new_image(slots, 64, DIM)
    rng = np.random.default_rng(seed)
    vecs = {}
    for i in range(nkeys):
        key = f"doc{i:08d}"
        assert support.image_set(img, key, f"value {i}".encode()) >= 0
        if i % 10 != 0:  # 10 % never embedded: zero vectors are results too
            v = rng.standard_normal(DIM).astype(np.float32)
            support.image_set_embedding(img, key, v)
            vecs[key] = v
        mask = int(rng.integers(0, 16))
        if mask:
            support.image_set_label(img, key, mask)
    return img, rng


def _mirror_equals_image(m, img):
    E, bloom, flags, epoch = support.image_to_dense(img, DIM)
    n = support.image_slots(img)
    dE, dbloom, dflags, depoch = m.read_mirror(0, n)
    live = flags != 0
    assert np.array_equal(dbloom[live], bloom[live])
    assert np.array_equal(depoch[live], epoch[live])
    assert np.array_equal(dflags & 3, flags & 3)
    assert np.array_equal(dE.view(np.uint32)[live], E.view(np.uint32)[live])


def _search_equals_oracle(m, img, q, mode=sg.MODE_FAST, **kw):
    okw = dict(bloom_mask=kw.get("bloom_mask", 0), min_sim=kw.get("min_similarity", 0.0))
    k = kw.pop("k", 10)
    out, counts = m.search(q, k, mode=mode, **kw)
    ref, surv = support.oracle_search_image(img, q, k, **okw)
    support.assert_topk_matches(out[0][:counts[0]], ref, exact=(mode == sg.MODE_EXACT))
    return out[0][:counts[0]]


def test_attach_image_snapshot_and_search():
    img, rng = _build()
    q = rng.standard_normal(DIM).astype(np.float32)
    with sg.Mirror.attach_image(img) as m:
        info = m.info()
        assert (info.dim, info.total_slots, info.slot_count, info.is_dense) == (DIM, 1024, 1024, 0)
        _mirror_equals_image(m, img)
        for mask in (0, 1, 3, 8):
            _search_equals_oracle(m, img, q, bloom_mask=mask)
        _search_equals_oracle(m, img, q, k=0)


def test_incremental_sync_uploads_only_dirty_slots():
    img, rng = _build()
    q = rng.standard_normal(DIM).astype(np.float32)
    with sg.Mirror.attach_image(img) as m:
        st = m.sync()
        assert (st.rows_uploaded, st.meta_updated, st.skipped_busy) == (0, 0, 0)
        # 1) set_embedding on 5 keys -> 5 rows
        for i in (1, 2, 3, 4, 5):
            support.image_set_embedding(img, f"doc{i:08d}", rng.standard_normal(DIM).astype(np.float32))
        st = m.sync()
        assert (st.rows_uploaded, st.meta_updated) == (5, 0)
        _mirror_equals_image(m, img)
        # 2) label changes move no slot epoch (splinter.c:843): metadata-only update, no row bytes
        support.image_set_label(img, "doc00000007", 1 << 40)
        support.image_unset_label(img, "doc00000008", 0xF)
        st = m.sync()
        assert st.rows_uploaded == 0 and st.meta_updated >= 1
        _mirror_equals_image(m, img)
        hits = _search_equals_oracle(m, img, q, bloom_mask=1 << 40)
        assert len(hits) == 1 and hits[0]["slot_idx"] == support.image_find(img, "doc00000007")
        # 3) unset moves neither H->epoch nor the bus (splinter.c:326-363): only a sweep sees it
        gone = support.image_find(img, "doc00000009")
        support.image_unset(img, "doc00000009")
        st = m.sync()
        assert st.rows_uploaded == 1
        _mirror_equals_image(m, img)
        assert gone not in _search_equals_oracle(m, img, q, k=0)["slot_idx"]
        # 4) a brand-new key lands in some empty slot with a zero vector (splinter.c:406-411)
        support.image_set(img, "fresh-key", b"hello")
        st = m.sync()
        assert st.rows_uploaded == 1
        _mirror_equals_image(m, img)
        assert support.image_find(img, "fresh-key") in _search_equals_oracle(m, img, q, k=0)["slot_idx"]


def test_retrain_aba_at_epoch_4_is_detected():
    """retrain forces the epoch to 4 (splinter.c:807-825) -- the same value a set + set_embedding
    slot already has.  The mirror must notice the scrubbed (all-zero) row anyway."""
    img, rng = _build()
    key = "doc00000001"
    slot = support.image_find(img, key)
    with sg.Mirror.attach_image(img) as m:
        _, _, _, epoch = m.read_mirror(slot, 1)
        assert epoch[0] == 4
        support.image_retrain(img, key)
        st = m.sync()
        assert st.aba_rechecked == 1 and st.rows_uploaded == 1
        E, _, _, epoch = m.read_mirror(slot, 1)
        assert epoch[0] == 4 and not E.any()
        _mirror_equals_image(m, img)


def test_busy_slot_is_skipped_not_spun_on():
    """A slot stuck at an odd epoch (dead writer) is skipped like splinter_get_embedding's EAGAIN
    (splinter.c:599); the previous mirror row stays and the slot is retried on the next sync."""
    img, rng = _build()
    key = "doc00000002"
    slot = support.image_find(img, key)
    off = support.HEADER_SIZE + slot * (support.SLOT_META + 4 * DIM)
    with sg.Mirror.attach_image(img) as m:
        before, _, _, _ = m.read_mirror(slot, 1)
        ep = img[off + 8: off + 16].view(np.uint64)
        emb = img[off + 128: off + 128 + 4 * DIM].view(np.float32)
        ep[0] = 5                      # writer active
        emb[:] = 123.0                 # half-written garbage
        st = m.sync()
        assert st.skipped_busy == 1 and st.rows_uploaded == 0
        after, _, _, _ = m.read_mirror(slot, 1)
        assert np.array_equal(before, after)
        ep[0] = 6                      # writer finished
        st = m.sync()
        assert st.skipped_busy == 0 and st.rows_uploaded == 1
        _mirror_equals_image(m, img)


def test_geometry_gate():
    img = support.new_image(64, 64, DIM)
    bad = img.copy()
    bad[0] ^= 0xFF
    with pytest.raises(sg.SplinterGpuError, match="not a splinter"):
        sg.Mirror.attach_image(bad)
    # a store built without embeddings has 128-byte slots
    no_emb = np.zeros(support.HEADER_SIZE + 64 * 128 + 64 * 64, np.uint8)
    no_emb[:16] = img[:16]
    with pytest.raises(sg.SplinterGpuError, match="without embeddings"):
        sg.Mirror.attach_image(no_emb)
    # D = 1024 header variant (4224-byte slots) is accepted and its dimension inferred
    img1024 = support.new_image(32, 64, 1024)
    with sg.Mirror.attach_image(img1024) as m:
        assert m.info().dim == 1024


def test_shards_carry_global_slot_indices_and_merge_to_the_global_topk():
    img, rng = _build(slots=2048, nkeys=1500)
    q = rng.standard_normal(DIM).astype(np.float32)
    ref, _ = support.oracle_search_image(img, q, 20, bloom_mask=1)
    parts = []
    for g in range(4):
        with sg.Mirror.attach_image(img, slot_begin=g * 512, slot_count=512) as m:
            out, counts = m.search(q, 20, bloom_mask=1, mode=sg.MODE_FAST)
            hits = out[0][:counts[0]]
            assert np.all((hits["slot_idx"] >= g * 512) & (hits["slot_idx"] < (g + 1) * 512))
            parts.append(hits)
    allhits = np.concatenate(parts)
    order = np.lexsort((allhits["slot_idx"], allhits["distance"], -allhits["similarity"]))
    support.assert_topk_matches(allhits[order][:20], ref)
