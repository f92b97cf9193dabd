"""Pins the CPU restatement (oracle/spl_oracle.c) to the reference's OWN code compiled unmodified
into oracle/_ref/ (oracle/Makefile).  Every check is bit-exact: scores, store images, candidate
sets, result order, and the real `search` command's output.

The reference's test-suite holds no golden vector for this path, so the compiled
reference itself is the anchor; tests/golden/ carries vectors generated from it for machines
where oracle/_ref/ is absent.
"""
import ctypes as C
import json
import os
import subprocess

import numpy as np
import pytest

import support

pytestmark = pytest.mark.skipif(not support.have_ref(), reason="oracle/_ref (compiled reference) not present")

DIM = 768


def _vecs(rng, n, dim, kind):
    if kind == "normal":
        return rng.standard_normal((n, dim)).astype(np.float32)
    if kind == "unit":
        v = rng.uniform(-1, 1, (n, dim)).astype(np.float32)
        return v / np.linalg.norm(v, axis=1, keepdims=True).astype(np.float32)
    if kind == "tiny":
        return (rng.standard_normal((n, dim)) * 1e-7).astype(np.float32)
    return (rng.standard_normal((n, dim)) * 1e4).astype(np.float32)


@pytest.mark.parametrize("dim", [768, 1024, 5, 1])
@pytest.mark.parametrize("kind", ["normal", "unit", "tiny", "huge"])
def test_scorers_bit_exact(dim, kind):
    """cosine_similarity / euclidean_distance (_search.c:43-62), incl. the denom < 1e-9f guard."""
    rng = np.random.default_rng(dim * 7 + len(kind))
    a, b = _vecs(rng, 64, dim, kind), _vecs(rng, 64, dim, kind)
    b[3] = 0.0          # zero vector -> similarity exactly 0
    b[4] = a[4]         # identical -> distance exactly 0
    ref, ora = support.ref_search(), support.oracle()
    for i in range(64):
        pa, pb = support._p(a[i]), support._p(b[i])
        rc, oc = ref.ref_cosine_similarity(pa, pb, dim), ora.splo_cosine(pa, pb, dim)
        rd, od = ref.ref_euclidean_distance(pa, pb, dim), ora.splo_euclid(pa, pb, dim)
        assert np.float32(rc).view(np.uint32) == np.float32(oc).view(np.uint32)
        assert np.float32(rd).view(np.uint32) == np.float32(od).view(np.uint32)


def test_comparator_agrees_with_compare_results():
    rng = np.random.default_rng(3)
    vals = np.round(rng.standard_normal(40), 1).astype(np.float32)  # many exact ties
    ref, ora = support.ref_search(), support.oracle()
    for _ in range(2000):
        sa, da, sb, db = (float(x) for x in rng.choice(vals, 4))
        r = ref.ref_compare(sa, da, sb, db)
        o = ora.splo_compare(0, sa, da, 7, sb, db, 7)  # equal idx: the pure reference order
        assert np.sign(r) == np.sign(o)
        if r == 0:  # the reference leaves these unspecified; the restatement breaks them by idx
            assert ora.splo_compare(0, sa, da, 1, sb, db, 2) < 0


def _shm_name(tag):
    return f"/splb200_{tag}_{os.getpid()}"


def _read_shm(name):
    with open("/dev/shm" + name, "rb") as f:
        return np.frombuffer(f.read(), np.uint8).copy()


@pytest.fixture
def ref_store_name():
    name = _shm_name("pin")
    try:
        os.unlink("/dev/shm" + name)
    except FileNotFoundError:
        pass
    yield name
    try:
        os.unlink("/dev/shm" + name)
    except FileNotFoundError:
        pass


def _apply_ops(ops, on_ref=None, img=None):
    """Run one op list through the reference API or through the restatement's image functions;
    returns the list of success flags so both sides can be compared (e.g. retrain of an unset key fails)."""
    ok = []
    for op in ops:
        kind, key = op[0], op[1]
        if on_ref is not None:
            kb = key.encode()
            if kind == "set":
                rc = on_ref.splinter_set(kb, op[2], len(op[2]))
            elif kind == "emb":
                v = np.ascontiguousarray(op[2], np.float32)
                rc = on_ref.splinter_set_embedding(kb, support._p(v))
            elif kind == "label":
                rc = on_ref.splinter_set_label(kb, op[2])
            elif kind == "unlabel":
                rc = on_ref.splinter_unset_label(kb, op[2])
            elif kind == "retrain":
                rc = on_ref.splinter_retrain_slot(kb)
            else:
                rc = on_ref.splinter_unset(kb)
        else:
            if kind == "set":
                rc = support.image_set(img, key, op[2])
            elif kind == "emb":
                rc = support.image_set_embedding(img, key, op[2])
            elif kind == "label":
                rc = support.image_set_label(img, key, op[2])
            elif kind == "unlabel":
                rc = support.image_unset_label(img, key, op[2])
            elif kind == "retrain":
                rc = support.image_retrain(img, key)
            else:
                rc = support.image_unset(img, key)
        ok.append(rc >= 0)
    return ok


def _op_list(rng, nkeys):
    ops = []
    for i in range(nkeys):
        key = f"doc{i:08d}"
        ops.append(("set", key, f"value number {i}".encode()))
        if i % 7:
            ops.append(("emb", key, rng.standard_normal(DIM).astype(np.float32)))
        m = int(rng.integers(0, 16))
        if m:
            ops.append(("label", key, m))
    for i in range(0, nkeys, 11):
        ops.append(("unset", f"doc{i:08d}"))
    for i in range(3, nkeys, 13):
        ops.append(("retrain", f"doc{i:08d}"))
    for i in range(5, nkeys, 17):
        ops.append(("unlabel", f"doc{i:08d}", 0x5))
    for i in range(0, nkeys, 22):          # re-insert into slots freed by unset (zeroed embedding)
        ops.append(("set", f"again{i:05d}", b"short"))
    for i in range(1, nkeys, 9):           # overwrite with a shorter value (hybrid scrub)
        ops.append(("set", f"doc{i:08d}", b"x"))
    return ops


def test_store_images_are_byte_identical_to_the_reference(ref_store_name):
    """splinter_create / set / set_embedding / labels / unset / retrain (splinter.c) vs splo_image_*."""
    ref = support.ref_store()
    slots, maxval, nkeys = 512, 64, 400
    rng = np.random.default_rng(5)
    ops = _op_list(rng, nkeys)
    assert ref.splinter_create(ref_store_name.encode(), slots, maxval) == 0
    fresh = _read_shm(ref_store_name)
    img = support.new_image(slots, maxval, DIM)
    assert np.array_equal(fresh, img), "splo_image_init differs from splinter_create"
    ok_ref = _apply_ops(ops, on_ref=ref)
    ref.splinter_close()
    ok_img = _apply_ops(ops, img=img)
    assert ok_ref == ok_img and sum(ok_ref) > 0.9 * len(ops)
    got = _read_shm(ref_store_name)
    diff = np.nonzero(got != img)[0]
    assert diff.size == 0, f"{diff.size} differing bytes, first at {diff[:8]}"


def test_candidate_sets_and_search_order_match_the_reference(ref_store_name):
    """enumerate_matches / list (splinter.c:466-477,977-989) + scorers + qsort(compare_results) + limit."""
    ref, rs = support.ref_store(), support.ref_search()
    slots, maxval, nkeys = 1024, 64, 900
    rng = np.random.default_rng(9)
    ops = _op_list(rng, nkeys)
    assert ref.splinter_create(ref_store_name.encode(), slots, maxval) == 0
    _apply_ops(ops, on_ref=ref)
    img = _read_shm(ref_store_name)          # the reference's own bytes
    support.oracle().splo_image_dim(support._p(img), img.nbytes)
    slot_size = 128 + 4 * DIM
    # the reference hands out POINTERS into its own mapping (splinter.c:472,986): turn them back into
    # slot indices through the mapping's base address (keys are not unique once unset + set has
    # re-used a slot earlier in a probe chain, so names cannot be used)
    base = None
    for ln in open("/proc/self/maps"):
        if ref_store_name.lstrip("/") in ln:
            base = int(ln.split("-")[0], 16)
            break
    assert base is not None

    def slot_of(addr):
        off = addr - base - support.HEADER_SIZE - 64
        assert off % slot_size == 0
        return off // slot_size

    # --- candidate enumeration through the reference's callbacks ---
    CB = C.CFUNCTYPE(None, C.c_void_p, C.c_uint64, C.c_void_p)
    for mask in (1, 3, 8, 0xF):
        seen = []
        cb = CB(lambda key, ep, user: seen.append(slot_of(key)))
        ref.splinter_enumerate_matches(C.c_uint64(mask), cb, None)
        out = np.zeros(slots, np.uint32)
        n = support.oracle().splo_enumerate_matches(support._p(img), mask, support._p(out))
        assert list(out[:n]) == seen
    arr = (C.c_void_p * slots)()
    cnt = C.c_size_t(0)
    assert ref.splinter_list(arr, C.c_size_t(slots), C.byref(cnt)) == 0
    listed = [slot_of(arr[i]) for i in range(cnt.value)]
    out = np.zeros(slots, np.uint32)
    n = support.oracle().splo_list(support._p(img), support._p(out))
    assert list(out[:n]) == listed
    ref.splinter_close()

    # --- scoring + sort + limit: reference statics over the dense copy vs restatement over the image ---
    E, bloom, flags, epoch = support.image_to_dense(img, DIM)
    for mask, k, min_sim, max_dist in ((0, 10, 0.0, 0.0), (3, 25, 0.0, 0.0), (0, 0, 0.0, 0.0), (1, 50, 0.01, 0.0),
                                       (0, 50, 0.0, 39.5)):
        q = rng.standard_normal(DIM).astype(np.float32)
        need = support.F_ENUM if mask else support.F_LIST
        cand = (((flags & need) != 0) & ((bloom & np.uint64(mask)) == np.uint64(mask))).astype(np.uint8)
        lim = k if k > 0 else slots
        oi, os_, od = np.zeros(lim, np.uint32), np.zeros(lim, np.float32), np.zeros(lim, np.float32)
        surv = rs.ref_topk_dense(support._p(q), support._p(E), slots, DIM, support._p(cand), min_sim, max_dist, k, 1,
                                 support._p(oi), support._p(os_), support._p(od))
        got, osurv = support.oracle_search_image(img, q, k, bloom_mask=mask, min_sim=min_sim, max_dist=max_dist)
        assert surv == osurv
        m = min(surv, lim)
        assert len(got) == m
        # random data carries no exact ties, so the reference's (unstable) qsort order is the total order
        assert np.array_equal(got["idx"], oi[:m])
        assert np.array_equal(got["similarity"].view(np.uint32), os_[:m].view(np.uint32))
        assert np.array_equal(got["distance"].view(np.uint32), od[:m].view(np.uint32))
        # multi-threaded split of the same reference code (the bench's all-cores baseline) agrees too
        oi2, os2, od2 = np.zeros(lim, np.uint32), np.zeros(lim, np.float32), np.zeros(lim, np.float32)
        rs.ref_topk_dense(support._p(q), support._p(E), slots, DIM, support._p(cand), min_sim, max_dist, k, 4,
                          support._p(oi2), support._p(os2), support._p(od2))
        assert np.array_equal(oi2[:m], oi[:m])


def test_real_search_command_end_to_end(ref_store_name, tmp_path):
    """The reference's actual cmd_search (_search.c:136-486), with a stub standing in for splinference,
    prints the same keys in the same order (and the same %.4f scores) as the restatement."""
    ref = support.ref_store()
    slots, maxval, nkeys = 512, 64, 300
    rng = np.random.default_rng(21)
    ops = _op_list(rng, nkeys)
    assert ref.splinter_create(ref_store_name.encode(), slots, maxval) == 0
    _apply_ops(ops, on_ref=ref)
    ref.splinter_close()
    q = rng.standard_normal(DIM).astype(np.float32)
    qfile = tmp_path / "q.f32"
    q.tofile(qfile)
    exe = str(support.REF_DIR / "ref_search_cli")
    for extra, mask, k in ((["--limit", "10"], 0, 10), (["--bloom", "3", "--limit", "15"], 3, 15),
                           (["--bloom", "1"], 1, 0)):
        r = subprocess.run([exe, ref_store_name, str(qfile), "some query text", "--json"] + extra,
                           capture_output=True, text=True, timeout=60)
        assert r.returncode == 0, r.stderr
        doc = json.loads(r.stdout)
        img = _read_shm(ref_store_name)
        support.oracle().splo_image_dim(support._p(img), img.nbytes)
        got, surv = support.oracle_search_image(img, q, k, bloom_mask=mask)
        assert doc["store"]["active_keys"] == surv
        assert len(doc["results"]) == len(got)
        slot_size = 128 + 4 * DIM
        for row, g in zip(doc["results"], got):
            off = support.HEADER_SIZE + int(g["idx"]) * slot_size + 64
            key = bytes(img[off:off + 64]).split(b"\0")[0].decode()
            assert row["key"] == key
            assert row["similarity"] == pytest.approx(float(g["similarity"]), abs=5.1e-5)
            assert row["distance"] == pytest.approx(float(g["distance"]), abs=5.1e-5)


def test_event_bus_and_epoch_offsets_match_the_reference_header():
    """spl_gpu_wait reads H->event_bus.owner_fd / owner_pid and H->epoch straight out of the mapping
    (libsplinter_b200/csrc/spl_host.cu: layout::H_BUS_OWNER_FD / H_BUS_OWNER_PID / H_EPOCH)."""
    rs = support.ref_search()
    assert rs.ref_offsetof_bus_owner_fd() == 4352
    assert rs.ref_offsetof_bus_owner_pid() == 4356
    assert rs.ref_offsetof_header_epoch() == 16
