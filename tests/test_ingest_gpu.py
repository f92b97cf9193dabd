"""SURVEY 8(f) rank 3, bulk side: spl_gpu_ingest_embeddings against splinference's one-key-at-a-time loop.

Two identical stores are filled the way the reference's `ingest` command fills one (VARTEXT chunk slots `doc.N`
labelled 0x200 and bumped, splinter_cli_cmd_ingest.c:140-172), plus bystanders the embedder must leave alone (a
non-VARTEXT key, a chunk that already has a vector, a chunk without the label).  Store A is embedded by
oracle/_ref/ref_backfill -- splinference.cpp:196-325 over the reference API with a stub model -- and store B by the
bulk call with the same stub as its batch callback and the reference's own splinter_set_embedding as the writer.
The two store files must end up byte-identical, the mirror must hold the new vectors after the call's own sync, and
a search must find a chunk by its own embedding.  keys/s of both loops are printed (plumbing only: the model is a stub;
the reference's arithmetic at that boundary is llama.cpp's and unpinned)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import libsplinter_b200 as sg
import support

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not (support.REF_DIR / "ref_backfill").exists(), reason="needs oracle/_ref")]
DIM, SLOTS, NCHUNK = 768, 16384, 6000
VARTEXT, CHUNK_LABEL = 1 << 7, 0x200


def _fill(ref, name):
    path = "/dev/shm" + name
    if os.path.exists(path):
        os.unlink(path)
    assert ref.splinter_create(name.encode(), SLOTS, 256) == 0
    ref.splinter_set_named_type.restype = C.c_int
    ref.splinter_set_named_type.argtypes = [C.c_char_p, C.c_uint16]
    rng = np.random.default_rng(4)
    for i in range(NCHUNK):
        key = f"doc.{i + 1}".encode()
        txt = (f"chunk {i} " + "lorem ipsum " * int(rng.integers(1, 15))).encode()
        assert ref.splinter_set(key, txt, len(txt)) == 0
        assert ref.splinter_set_named_type(key, VARTEXT) == 0
        assert ref.splinter_set_label(key, CHUNK_LABEL) == 0
        assert ref.splinter_bump_slot(key) == 0
    pre = rng.standard_normal(DIM).astype(np.float32)
    for key, typed, label, vec in ((b"plain_kv", False, CHUNK_LABEL, None), (b"already.1", True, CHUNK_LABEL, pre),
                                   (b"unlabelled.1", True, 0, None)):
        assert ref.splinter_set(key, b"some text", 9) == 0
        if typed:
            assert ref.splinter_set_named_type(key, VARTEXT) == 0
        if label:
            assert ref.splinter_set_label(key, label) == 0
        if vec is not None:
            assert ref.splinter_set_embedding(key, support._p(vec)) == 0
    ref.splinter_close()
    return path


def test_bulk_ingest_writes_what_the_per_key_loop_writes():
    ref = support.ref_store()
    name_a, name_b = f"/splb200_ing_a_{os.getpid()}", f"/splb200_ing_b_{os.getpid()}"
    path_a, path_b = _fill(ref, name_a), _fill(ref, name_b)
    try:
        assert open(path_a, "rb").read() == open(path_b, "rb").read()       # same starting point
        # A: the reference's loop, one key at a time
        r = subprocess.run([str(support.REF_DIR / "ref_backfill"), name_a, "200"], capture_output=True, text=True, timeout=120)
        assert r.returncode == 0, r.stderr
        emb_a, failed_a, sec_a = r.stdout.split()
        assert int(emb_a) == NCHUNK and int(failed_a) == 0
        # B: the bulk call; the writer is the reference's own splinter_set_embedding on the store this process has open
        assert ref.splinter_open(name_b.encode()) == 0
        stub = C.CDLL(str(support.REF_DIR / "libstub_embed.so"))
        seen = C.c_uint64(0)
        with sg.Mirror.attach(name_b) as m:
            st = m.ingest_embeddings(CHUNK_LABEL, 512, stub.stub_embed_batch, C.addressof(seen), ref.splinter_set_embedding)
            assert st.candidates == NCHUNK and st.embedded == NCHUNK, (st.candidates, st.embedded, st.raced, st.skipped_busy)
            assert st.skipped_busy == 0 and st.raced == 0 and st.write_failed == 0 and seen.value == NCHUNK
            assert st.batches == -(-NCHUNK // 512)
            sec_b = (st.gather_ms + st.embed_ms + st.write_ms) / 1e3
            print(f"embedding {NCHUNK} chunks: per-key loop {NCHUNK / float(sec_a):,.0f} keys/s, bulk {NCHUNK / sec_b:,.0f} keys/s "
                  f"(gather {st.gather_ms:.1f} ms, stub model {st.embed_ms:.1f} ms, write-back {st.write_ms:.1f} ms, "
                  f"one sync of {NCHUNK} rows {st.sync_ms:.1f} ms)")
            ref.splinter_close()
            img_a, img_b = open(path_a, "rb").read(), open(path_b, "rb").read()
            assert img_a == img_b                                            # byte-identical stores
            # a second call finds nothing left to do
            assert ref.splinter_open(name_b.encode()) == 0
            st2 = m.ingest_embeddings(CHUNK_LABEL, 512, stub.stub_embed_batch, C.addressof(seen), ref.splinter_set_embedding)
            ref.splinter_close()
            assert st2.candidates == 0 and seen.value == NCHUNK
            # the mirror holds the vectors (the call's closing sync) and a search finds a chunk by its own embedding
            img = np.frombuffer(img_b, np.uint8).copy()
            Es, _, _, _ = support.image_to_dense(img, DIM)
            slot = m.find_key("doc.1234")
            assert slot is not None and np.abs(Es[slot]).sum() > 0
            out, cnt = m.search(Es[slot], 5, bloom_mask=CHUNK_LABEL)
            assert cnt[0] == 5 and int(out[0][0]["slot_idx"]) == slot and abs(float(out[0][0]["similarity"]) - 1.0) < 1e-5
            ref_hits, _ = support.oracle_search_image(img, Es[slot], 5, bloom_mask=CHUNK_LABEL)
            support.assert_topk_matches(out[0][:5], ref_hits)
    finally:
        for p in (path_a, path_b):
            if os.path.exists(p):
                os.unlink(p)
