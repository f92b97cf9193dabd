"""Test / bench support: ctypes views of the oracle (oracle/libspl_oracle.so), of the compiled
reference (oracle/_ref/, when present) and comparison helpers.

Nothing here is imported by the product (libsplinter_b200/).
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
ORACLE_SO = ROOT / "oracle" / "libspl_oracle.so"
REF_DIR = ROOT / "oracle" / "_ref"

ORDER_COSINE, ORDER_L2 = 0, 1
F_LIST, F_ENUM, F_EMB = 1, 2, 4
HEADER_SIZE, SLOT_META = 5440, 128

SPLO_RESULT = np.dtype([("idx", "<u4"), ("similarity", "<f4"), ("distance", "<f4"), ("has_embedding", "<u4")])

# fp32 tolerance of SPL_MODE_FAST against the reference's sequential sums:
# cosine lives on a unit scale, so its bound is absolute; distance is relative.
TOL_SIM = 1e-5
TOL_DIST_REL = 1e-5

_vp, _u32, _u64, _f32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_float
_oracle = None
_ref_search = None
_ref_store = None


def _p(a):
    return None if a is None else a.ctypes.data_as(_vp)


def oracle() -> C.CDLL:
    global _oracle
    if _oracle is None:
        lib = C.CDLL(str(ORACLE_SO))
        lib.splo_fnv1a.restype = _u64
        lib.splo_fnv1a.argtypes = [C.c_char_p]
        lib.splo_cosine.restype = _f32
        lib.splo_cosine.argtypes = [_vp, _vp, C.c_int]
        lib.splo_euclid.restype = _f32
        lib.splo_euclid.argtypes = [_vp, _vp, C.c_int]
        lib.splo_compare.restype = C.c_int
        lib.splo_compare.argtypes = [C.c_int, _f32, _f32, _u32, _f32, _f32, _u32]
        lib.splo_image_size.restype = C.c_size_t
        lib.splo_image_size.argtypes = [_u32, _u32, _u32]
        lib.splo_image_init.restype = C.c_int
        lib.splo_image_init.argtypes = [_vp, _u32, _u32, _u32]
        lib.splo_image_find.restype = C.c_long
        lib.splo_image_find.argtypes = [_vp, C.c_char_p]
        lib.splo_image_set.restype = C.c_long
        lib.splo_image_set.argtypes = [_vp, C.c_char_p, C.c_char_p, _u32]
        lib.splo_image_set_embedding.restype = C.c_int
        lib.splo_image_set_embedding.argtypes = [_vp, C.c_char_p, _vp]
        for name in ("splo_image_set_label", "splo_image_unset_label"):
            getattr(lib, name).restype = C.c_int
            getattr(lib, name).argtypes = [_vp, C.c_char_p, _u64]
        for name in ("splo_image_retrain", "splo_image_unset"):
            getattr(lib, name).restype = C.c_int
            getattr(lib, name).argtypes = [_vp, C.c_char_p]
        lib.splo_image_dim.restype = _u32
        lib.splo_image_dim.argtypes = [_vp, C.c_size_t]
        lib.splo_image_fill_synthetic.restype = C.c_long
        lib.splo_image_fill_synthetic.argtypes = [_vp, _u32, _u64, _u64, C.c_int]
        lib.splo_enumerate_matches.restype = C.c_size_t
        lib.splo_enumerate_matches.argtypes = [_vp, _u64, _vp]
        lib.splo_list.restype = C.c_size_t
        lib.splo_list.argtypes = [_vp, _vp]
        lib.splo_snapshot_row.restype = C.c_int
        lib.splo_snapshot_row.argtypes = [_vp, _u32, _vp, _vp]
        lib.splo_image_to_dense.restype = None
        lib.splo_image_to_dense.argtypes = [_vp, _vp, _vp, _vp, _vp]
        lib.splo_search_image.restype = C.c_long
        lib.splo_search_image.argtypes = [_vp, _vp, C.c_int, C.c_int, _u64, _f32, _f32, C.c_char_p, _vp, _vp]
        lib.splo_search_dense.restype = C.c_long
        lib.splo_search_dense.argtypes = [_vp, _u32, _vp, _vp, _vp, _u64, _u32, C.c_int, C.c_int, _u64, _f32, _f32,
                                          _u32, C.c_int, _vp, _vp]
        lib.splo_merge_topk.restype = _u32
        lib.splo_merge_topk.argtypes = [_vp, _vp, _u32, _u32, C.c_int, _u32, _vp]
        lib.splo_mix64.restype = _u64
        lib.splo_mix64.argtypes = [_u64]
        lib.splo_synth_class.restype = C.c_int
        lib.splo_synth_class.argtypes = [_u64, _u64]
        lib.splo_synth_row.restype = None
        lib.splo_synth_row.argtypes = [_u64, _u64, _u32, _vp]
        lib.splo_synth_bloom.restype = _u64
        lib.splo_synth_bloom.argtypes = [_u64, _u64]
        lib.splo_synth_dense.restype = None
        lib.splo_synth_dense.argtypes = [_u64, _u64, _u64, _u64, _u32, _vp, _vp, _vp, C.c_int]
        _oracle = lib
    return _oracle


def have_ref() -> bool:
    return (REF_DIR / "libref_search.so").exists() and (REF_DIR / "libsplinter.so").exists()


def _bind_store(lib):
    lib.splinter_create.restype = C.c_int
    lib.splinter_create.argtypes = [C.c_char_p, C.c_size_t, C.c_size_t]
    lib.splinter_open.restype = C.c_int
    lib.splinter_open.argtypes = [C.c_char_p]
    lib.splinter_close.restype = None
    lib.splinter_set.restype = C.c_int
    lib.splinter_set.argtypes = [C.c_char_p, C.c_char_p, C.c_size_t]
    lib.splinter_unset.restype = C.c_int
    lib.splinter_unset.argtypes = [C.c_char_p]
    lib.splinter_set_embedding.restype = C.c_int
    lib.splinter_set_embedding.argtypes = [C.c_char_p, _vp]
    lib.splinter_get_embedding.restype = C.c_int
    lib.splinter_get_embedding.argtypes = [C.c_char_p, _vp]
    lib.splinter_set_label.restype = C.c_int
    lib.splinter_set_label.argtypes = [C.c_char_p, _u64]
    lib.splinter_unset_label.restype = C.c_int
    lib.splinter_unset_label.argtypes = [C.c_char_p, _u64]
    lib.splinter_retrain_slot.restype = C.c_int
    lib.splinter_retrain_slot.argtypes = [C.c_char_p]
    lib.splinter_bump_slot.restype = C.c_int
    lib.splinter_bump_slot.argtypes = [C.c_char_p]
    lib.splinter_enumerate_matches.restype = None
    lib.splinter_list.restype = C.c_int
    lib.splinter_event_bus_init.restype = C.c_int
    return lib


def ref_store() -> C.CDLL:
    """The reference's own libsplinter.so (compiled unmodified by oracle/Makefile)."""
    global _ref_store
    if _ref_store is None:
        _ref_store = _bind_store(C.CDLL(str(REF_DIR / "libsplinter.so"), mode=C.RTLD_GLOBAL))
    return _ref_store


_ref_store_p = None


def ref_store_persistent() -> C.CDLL:
    """The same source built with -DSPLINTER_PERSISTENT: file-backed stores (splinter.c:157-168, 237-241)."""
    global _ref_store_p
    if _ref_store_p is None:
        _ref_store_p = _bind_store(C.CDLL(str(REF_DIR / "libsplinter_p.so"), mode=C.RTLD_LOCAL))
    return _ref_store_p


def ref_search() -> C.CDLL:
    """The harness TU that #includes the reference's splinter_cli_cmd_search.c."""
    global _ref_search
    if _ref_search is None:
        ref_store()
        lib = C.CDLL(str(REF_DIR / "libref_search.so"))
        lib.ref_cosine_similarity.restype = _f32
        lib.ref_cosine_similarity.argtypes = [_vp, _vp, C.c_int]
        lib.ref_euclidean_distance.restype = _f32
        lib.ref_euclidean_distance.argtypes = [_vp, _vp, C.c_int]
        lib.ref_compare.restype = C.c_int
        lib.ref_compare.argtypes = [_f32, _f32, _f32, _f32]
        lib.ref_topk_dense.restype = C.c_long
        lib.ref_topk_dense.argtypes = [_vp, _vp, C.c_size_t, C.c_int, _vp, _f32, _f32, C.c_int, C.c_int, _vp, _vp, _vp]
        lib.ref_score_dense.restype = None
        lib.ref_score_dense.argtypes = [_vp, _vp, C.c_size_t, C.c_int, _vp, _vp]
        lib.ref_sizeof_slot.restype = C.c_size_t
        lib.ref_sizeof_header.restype = C.c_size_t
        lib.ref_sizeof_result.restype = C.c_size_t
        lib.ref_embed_dim.restype = C.c_int
        for f in ("ref_offsetof_bus_owner_fd", "ref_offsetof_bus_owner_pid", "ref_offsetof_header_epoch"):
            getattr(lib, f).restype = C.c_size_t
        _ref_search = lib
    return _ref_search


# ---------------------------------------------------------------- synthetic data
def synth_dense(seed_c, seed_b, row0, n, dim, threads=None):
    E = np.empty((n, dim), np.float32)
    bloom = np.empty(n, np.uint64)
    flags = np.empty(n, np.uint8)
    oracle().splo_synth_dense(seed_c, seed_b, row0, n, dim, _p(E), _p(bloom), _p(flags), threads or os.cpu_count() or 1)
    return E, bloom, flags


def synth_queries(seed, nq, dim):
    q = np.empty((nq, dim), np.float32)
    for i in range(nq):
        oracle().splo_synth_row(seed, i, dim, _p(q[i]))
    return q


# ---------------------------------------------------------------- store images
def new_image(slots, max_val_sz=64, dim=768) -> np.ndarray:
    size = oracle().splo_image_size(slots, max_val_sz, dim)
    img = np.zeros(size, np.uint8)
    assert oracle().splo_image_init(_p(img), slots, max_val_sz, dim) == 0
    return img


def image_set(img, key, val=b"x"):
    _register(img)
    return oracle().splo_image_set(_p(img), key.encode(), val, len(val))


def image_set_embedding(img, key, vec):
    _register(img)
    vec = np.ascontiguousarray(vec, np.float32)
    return oracle().splo_image_set_embedding(_p(img), key.encode(), _p(vec))


def image_set_label(img, key, mask):
    _register(img)
    return oracle().splo_image_set_label(_p(img), key.encode(), mask)


def image_unset_label(img, key, mask):
    _register(img)
    return oracle().splo_image_unset_label(_p(img), key.encode(), mask)


def image_retrain(img, key):
    _register(img)
    return oracle().splo_image_retrain(_p(img), key.encode())


def image_unset(img, key):
    _register(img)
    return oracle().splo_image_unset(_p(img), key.encode())


def image_find(img, key):
    _register(img)
    return oracle().splo_image_find(_p(img), key.encode())


def image_fill_synthetic(img, nkeys, seed_c=1234, seed_b=99, threads=None):
    _register(img)
    return oracle().splo_image_fill_synthetic(_p(img), nkeys, seed_c, seed_b, threads or os.cpu_count() or 1)


def image_slots(img):
    return int(img[8:12].view(np.uint32)[0])


def _register(img):
    """The oracle maps an image's base address to its dimension; a buffer it did not create itself (a file read
    back, a copy) may sit where an earlier image of another dimension lived.  Re-derive it from the geometry."""
    oracle().splo_image_dim(_p(img), img.nbytes)


def image_to_dense(img, dim=768):
    _register(img)
    n = image_slots(img)
    E = np.empty((n, dim), np.float32)
    bloom = np.empty(n, np.uint64)
    flags = np.empty(n, np.uint8)
    epoch = np.empty(n, np.uint64)
    oracle().splo_image_to_dense(_p(img), _p(E), _p(bloom), _p(flags), _p(epoch))
    return E, bloom, flags, epoch


# ---------------------------------------------------------------- oracle searches
def oracle_search_dense(q, E, bloom, flags, k, order=ORDER_COSINE, bloom_mask=0, min_sim=0.0, max_dist=0.0,
                        exclude=0xFFFFFFFF, threads=None):
    q = np.ascontiguousarray(q, np.float32)
    if q.ndim == 1:
        q = q[None]
    nq, dim = q.shape
    n = E.shape[0]
    lim = k if k > 0 else max(n, 1)
    out = np.zeros((nq, lim), SPLO_RESULT)
    counts = np.zeros(nq, np.uint32)
    rc = oracle().splo_search_dense(_p(q), nq, _p(E), _p(bloom), _p(flags), n, dim, order, k, bloom_mask, min_sim,
                                    max_dist, exclude, threads or os.cpu_count() or 1, _p(out), _p(counts))
    assert rc >= 0
    return out, counts


def oracle_search_image(img, q, k, order=ORDER_COSINE, bloom_mask=0, min_sim=0.0, max_dist=0.0, exclude_key=None):
    q = np.ascontiguousarray(q, np.float32)
    _register(img)
    n = image_slots(img)
    out = np.zeros(n, SPLO_RESULT)
    surv = C.c_long(0)
    cnt = oracle().splo_search_image(_p(img), _p(q), order, k, bloom_mask, min_sim, max_dist,
                                     exclude_key.encode() if exclude_key else None, _p(out), C.byref(surv))
    assert cnt >= 0
    return out[:cnt], int(surv.value)


# ---------------------------------------------------------------- comparisons
def assert_topk_matches(got, ref, exact=False, order=ORDER_COSINE, tol_sim=TOL_SIM, tol_dist_rel=TOL_DIST_REL):
    """got: RESULT_DTYPE records (slot_idx, similarity, distance); ref: SPLO_RESULT records.

    exact=True  : indices identical and both scores bit-identical (SPL_MODE_EXACT).
    exact=False : rank-wise scores within tolerance; index differences only between results whose
                  oracle scores are tied within the tolerance (SPL_MODE_FAST / bf16 paths).
    """
    assert len(got) == len(ref), f"count {len(got)} != oracle {len(ref)}"
    gi, ri = got["slot_idx"].astype(np.int64), ref["idx"].astype(np.int64)
    gs, rs = got["similarity"], ref["similarity"]
    gd, rd = got["distance"], ref["distance"]
    if exact:
        assert np.array_equal(gi, ri), f"index mismatch at {np.nonzero(gi != ri)[0][:8]}"
        assert np.array_equal(gs.view(np.uint32), rs.view(np.uint32)), "similarity not bit-identical"
        assert np.array_equal(gd.view(np.uint32), rd.view(np.uint32)), "distance not bit-identical"
        return
    if len(got) == 0:
        return
    finite = np.isfinite(rd) & (rd < 3e38)
    same = gi == ri
    if order == ORDER_COSINE:
        # rank-wise similarities agree; distances are compared where the indices agree
        assert np.all(np.abs(gs - rs) <= tol_sim), f"max |d sim| {np.abs(gs - rs).max():.3g} > {tol_sim}"
        chk = finite & same
    else:
        chk = finite
        assert np.all(np.abs(gs[same] - rs[same]) <= tol_sim), "similarity off on matching indices"
    if chk.any():
        rel = np.abs(gd[chk] - rd[chk]) / np.maximum(np.abs(rd[chk]), 1e-30)
        assert np.all((rel <= tol_dist_rel) | (np.abs(gd[chk] - rd[chk]) <= 1e-7)), f"max rel |d dist| {rel.max():.3g}"
    primary_r = rs if order == ORDER_COSINE else rd
    tol_p = 2 * tol_sim if order == ORDER_COSINE else 2 * tol_dist_rel * (float(np.max(np.abs(rd[finite]))) if finite.any() else 0.0)
    ref_pos = {int(ix): j for j, ix in enumerate(ri)}
    for i in np.nonzero(gi != ri)[0]:
        j = ref_pos.get(int(gi[i]))
        if j is not None:
            assert abs(float(primary_r[j]) - float(primary_r[i])) <= tol_p, \
                f"rank {i}: idx {gi[i]} vs oracle {ri[i]} is not a tie (oracle scores {primary_r[j]} vs {primary_r[i]})"
        else:  # boundary tie: the extra element must score like the oracle's last one
            g_primary = float(gs[i] if order == ORDER_COSINE else gd[i])
            assert abs(g_primary - float(primary_r[-1])) <= tol_p, \
                f"rank {i}: idx {gi[i]} not in oracle top-k and not tied with its boundary"
