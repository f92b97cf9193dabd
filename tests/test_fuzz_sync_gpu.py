"""Random mutation sequences against the sync layer: set / set_embedding / labels / retrain / unset / re-set on a
store image (through the oracle's restatement of the reference's writers, pinned byte-for-byte by
test_oracle_pinned.py), with spl_gpu_sync at random points.  After every sync the mirror must equal the store and
a search must equal the oracle -- whatever the interleaving did to hashes and epochs (the epoch is not
monotonic: retrain forces it to 4, unset + set restart it)."""
import numpy as np
import pytest

import libsplinter_b200 as sg
import support

pytestmark = pytest.mark.gpu


def _mirror_equals_image(m, img, dim):
    E, bloom, flags, epoch = support.image_to_dense(img, dim)
    n = support.image_slots(img)
    dE, dbloom, dflags, depoch = m.read_mirror(0, n)
    live = flags != 0
    assert np.array_equal(dflags & 3, flags & 3)
    assert np.array_equal(dbloom[live], bloom[live])
    assert np.array_equal(depoch[live], epoch[live])
    assert np.array_equal(dE.view(np.uint32)[live], E.view(np.uint32)[live])


@pytest.mark.parametrize("seed,dim", [(1, 128), (2, 128), (3, 768), (4, 256), (5, 1024), (6, 64), (7, 128), (8, 768)])
def test_random_mutations_then_sync(seed, dim):
    rng = np.random.default_rng(seed)
    slots, nkeys = 512, 300
    img = support.new_image(slots, 64, dim)
    keys = [f"k{i:05d}" for i in range(nkeys)]
    for key in keys[:200]:
        support.image_set(img, key, b"v")
        if rng.integers(0, 4):
            support.image_set_embedding(img, key, rng.standard_normal(dim).astype(np.float32))
    q = rng.standard_normal((3, dim)).astype(np.float32)
    with sg.Mirror.attach_image(img, sync_threads=int(rng.integers(1, 5))) as m:
        _mirror_equals_image(m, img, dim)
        for rnd in range(40):
            for _ in range(int(rng.integers(1, 80))):
                key = keys[int(rng.integers(0, nkeys))]
                op = int(rng.integers(0, 9))
                if op <= 1:
                    support.image_set(img, key, b"value" if op else b"v2")
                elif op <= 3:
                    support.image_set_embedding(img, key, rng.standard_normal(dim).astype(np.float32))
                elif op == 4:
                    support.image_set_label(img, key, int(rng.integers(1, 256)))
                elif op == 5:
                    support.image_unset_label(img, key, int(rng.integers(1, 256)))
                elif op == 6:
                    support.image_retrain(img, key)
                elif op == 7:
                    support.image_unset(img, key)
                else:  # unset + set + same vector as some other key: every detector sees "something changed"
                    support.image_unset(img, key)
                    support.image_set(img, key, b"again")
                    support.image_set_embedding(img, key, rng.standard_normal(dim).astype(np.float32))
            st = m.sync()
            assert st.skipped_busy == 0          # nobody writes while we sweep
            _mirror_equals_image(m, img, dim)
            mask = int(rng.choice([0, 1, 3, 0x80]))
            for i in range(3):
                out, cnt = m.search(q[i], 10, bloom_mask=mask)
                ref, _ = support.oracle_search_image(img, q[i], 10, bloom_mask=mask)
                support.assert_topk_matches(out[0][:cnt[0]], ref)
        assert m.sync().rows_uploaded == 0       # quiescent: nothing to do


def test_batched_search_follows_incremental_sync():
    """Config C5's combination: dirty-slot re-upload into BOTH mirrors (fp32 rows and the tiled bf16 shadow with
    its norms) and the tensor-core path searching them.  After every sync the batched result must equal the oracle
    on the store image."""
    rng = np.random.default_rng(12)
    dim, slots, nkeys = 256, 2048, 1500
    img = support.new_image(slots, 64, dim)
    keys = [f"k{i:05d}" for i in range(nkeys)]

    def unit():
        v = rng.standard_normal(dim).astype(np.float32)
        return v / np.linalg.norm(v)

    for key in keys[:1200]:
        support.image_set(img, key, b"v")
        support.image_set_embedding(img, key, unit())
        if rng.integers(0, 2):
            support.image_set_label(img, key, 0x1)
    q = np.stack([unit() for _ in range(20)])
    with sg.Mirror.attach_image(img, flags=sg.OPT_BF16_MIRROR, sync_threads=3) as m:
        assert m.info().has_bf16_mirror == 1
        for rnd in range(12):
            for _ in range(int(rng.integers(20, 200))):
                key = keys[int(rng.integers(0, nkeys))]
                op = int(rng.integers(0, 6))
                if op <= 2:
                    support.image_set(img, key, b"v")
                    support.image_set_embedding(img, key, unit())
                elif op == 3:
                    support.image_set_label(img, key, 0x1)
                elif op == 4:
                    support.image_unset_label(img, key, 0x1)
                else:
                    support.image_unset(img, key)
            m.sync()
            for order, mask in ((sg.ORDER_COSINE, 0), (sg.ORDER_L2, 0x1)):
                out, cnt = m.search(q, 8, mode=sg.MODE_BATCH_BF16, order=order, bloom_mask=mask)
                for i in range(q.shape[0]):
                    ref, _ = support.oracle_search_image(img, q[i], 8, order=order, bloom_mask=mask)
                    assert cnt[i] == len(ref)
                    support.assert_topk_matches(out[i][:cnt[i]], ref, order=order)
