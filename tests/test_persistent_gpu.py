"""file-backed (SPLINTER_PERSISTENT) stores.  The store file is created and written by
the reference's own persistent build (oracle/_ref/libsplinter_p.so = splinter.c with -DSPLINTER_PERSISTENT,
CMakeLists.txt:108-110); SPL_OPT_PERSISTENT maps the same file, and searches must equal the CPU oracle run
on the file's bytes, before and after the reference mutates it."""
import os

import numpy as np
import pytest

import libsplinter_b200 as sg
import support

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not (support.REF_DIR / "libsplinter_p.so").exists(), reason="needs oracle/_ref")]

DIM, SLOTS, NKEYS = 768, 512, 300


def _vec(rng):
    v = rng.standard_normal(DIM).astype(np.float32)
    return v / np.linalg.norm(v)


def _check(m, path, q, **kw):
    img = np.fromfile(path, dtype=np.uint8)     # page cache == the MAP_SHARED mapping
    # the oracle keeps a base-address -> dimension registry for images it did not create itself; a fresh buffer
    # can land on the address of an earlier test's image of another dimension
    assert support.oracle().splo_image_dim(support._p(img), img.nbytes) == DIM
    okw = dict(kw)
    for mode in (sg.MODE_FAST, sg.MODE_EXACT):
        out, cnt = m.search(q, 10, mode=mode, **kw)
        for i in range(q.shape[0]):
            ref, rc = support.oracle_search_image(img, q[i], 10, **okw)
            support.assert_topk_matches(out[i][:cnt[i]], ref, exact=(mode == sg.MODE_EXACT))


def test_file_backed_store_written_by_the_reference(tmp_path):
    refp = support.ref_store_persistent()
    path = str(tmp_path / "store.spl")
    assert refp.splinter_create(path.encode(), SLOTS, 64) == 0
    rng = np.random.default_rng(5)
    try:
        for k in range(NKEYS):
            key = f"doc{k:05d}".encode()
            assert refp.splinter_set(key, b"value", 5) == 0
            assert refp.splinter_set_embedding(key, support._p(_vec(rng))) == 0
            if k % 3 == 0:
                assert refp.splinter_set_label(key, 0x5) == 0
        assert os.path.getsize(path) == 5440 + SLOTS * (128 + 4 * DIM) + SLOTS * 64
        # a shm-style attach must not find it; the persistent flag must
        with pytest.raises(sg.SplinterGpuError):
            sg.Mirror.attach(path)
        m = sg.Mirror.attach(path, flags=sg.OPT_PERSISTENT)
        info = m.info()
        assert (info.total_slots, info.dim) == (SLOTS, DIM)
        q = np.stack([_vec(rng) for _ in range(3)])
        _check(m, path, q)
        _check(m, path, q, bloom_mask=0x5)
        # the reference keeps writing the file; the mirror follows
        for k in range(0, 40):
            assert refp.splinter_set_embedding(f"doc{k:05d}".encode(), support._p(_vec(rng))) == 0
        assert refp.splinter_unset(b"doc00100") >= 0
        assert refp.splinter_set_label(b"doc00007", 0x5) == 0
        assert m.wait(1000) is True
        st = m.sync()
        assert st.rows_uploaded >= 40
        _check(m, path, q)
        _check(m, path, q, bloom_mask=0x5)
        m.close()
    finally:
        refp.splinter_close()
