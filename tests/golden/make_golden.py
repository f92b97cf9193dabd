#!/usr/bin/env python
"""Generates tests/golden/search_golden.npz from the reference's OWN code compiled into oracle/_ref/.

Run in the build container (where /root/reference exists and `make -C oracle ref` has been run):
    python tests/golden/make_golden.py

Inputs are regenerated from seeds by the counter RNG (oracle/spl_oracle.c splo_synth_*: input
generation only); every OUTPUT stored here comes from the reference:
  * scorer_*  : ref cosine_similarity / euclidean_distance (_search.c:43-62) on 96 vector pairs
  * c1_*      : config C1 (4096 slots x 768, cosine top-10) -- a store built through the reference's
                splinter_create/_set/_set_embedding/_set_label, candidates from its
                splinter_enumerate_matches / splinter_list, scored and sorted by its statics +
                qsort(compare_results); slot indices from pointer arithmetic on its own mapping
  * cli_json  : the reference's real `search --json` output for one query (text)
"""
import ctypes as C
import json
import os
import subprocess
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parent))
import support  # noqa: E402

DIM, SLOTS, NKEYS, MAXVAL = 768, 4096, 4090, 64
SEED_C, SEED_B, SEED_Q = 1234, 99, 4321


def build_reference_store(name):
    ref = support.ref_store()
    path = "/dev/shm" + name
    if os.path.exists(path):
        os.unlink(path)
    assert ref.splinter_create(name.encode(), SLOTS, MAXVAL) == 0
    row = np.empty(DIM, np.float32)
    for k in range(NKEYS):
        cls = support.oracle().splo_synth_class(SEED_C, k)
        if cls == 0:
            continue
        key = f"doc{k:08d}".encode()
        val = f"v{k:08d}".encode()
        assert ref.splinter_set(key, val, len(val)) == 0
        if cls == 2:
            support.oracle().splo_synth_row(SEED_C, k, DIM, support._p(row))
            assert ref.splinter_set_embedding(key, support._p(row)) == 0
        assert ref.splinter_set_label(key, C.c_uint64(support.oracle().splo_synth_bloom(SEED_B, k))) == 0
    return ref


def main():
    assert support.have_ref(), "run `make -C oracle ref` first"
    rs = support.ref_search()
    out = {}
    # ---- scorer known answers ----
    n = 96
    a = np.stack([support.synth_queries(7, n, DIM)[i] for i in range(n)])
    b = np.stack([support.synth_queries(8, n, DIM)[i] for i in range(n)])
    b[5] = 0.0
    b[6] = a[6]
    a[7] *= 1e-6
    cos = np.array([rs.ref_cosine_similarity(support._p(a[i]), support._p(b[i]), DIM) for i in range(n)], np.float32)
    dist = np.array([rs.ref_euclidean_distance(support._p(a[i]), support._p(b[i]), DIM) for i in range(n)], np.float32)
    out["scorer_cos_bits"] = cos.view(np.uint32)
    out["scorer_dist_bits"] = dist.view(np.uint32)

    # ---- C1 through a real reference store ----
    name = f"/splb200_golden_{os.getpid()}"
    ref = build_reference_store(name)
    base = None
    for ln in open("/proc/self/maps"):
        if name.lstrip("/") in ln:
            base = int(ln.split("-")[0], 16)
            break
    slot_size = 128 + 4 * DIM
    slot_of = lambda addr: (addr - base - support.HEADER_SIZE - 64) // slot_size  # noqa: E731
    img = np.frombuffer(open("/dev/shm" + name, "rb").read(), np.uint8).copy()
    E, bloom, flags, epoch = support.image_to_dense(img, DIM)
    q = support.synth_queries(SEED_Q, 4, DIM)
    CB = C.CFUNCTYPE(None, C.c_void_p, C.c_uint64, C.c_void_p)
    masks = [0, 0x3, 1 << 63, 0xF]
    K = 10
    idx = np.zeros((len(masks), 4, K), np.uint32)
    sim = np.zeros((len(masks), 4, K), np.float32)
    dst = np.zeros((len(masks), 4, K), np.float32)
    ncand = np.zeros(len(masks), np.uint32)
    for mi, mask in enumerate(masks):
        cand = np.zeros(SLOTS, np.uint8)
        if mask:
            seen = []
            cb = CB(lambda key, ep, user: seen.append(slot_of(key)))
            ref.splinter_enumerate_matches(C.c_uint64(mask), cb, None)
            cand[seen] = 1
        else:
            arr = (C.c_void_p * SLOTS)()
            cnt = C.c_size_t(0)
            assert ref.splinter_list(arr, C.c_size_t(SLOTS), C.byref(cnt)) == 0
            cand[[slot_of(arr[i]) for i in range(cnt.value)]] = 1
        ncand[mi] = cand.sum()
        for qi in range(4):
            rs.ref_topk_dense(support._p(q[qi]), support._p(E), SLOTS, DIM, support._p(cand), 0.0, 0.0, K, 1,
                              support._p(idx[mi, qi]), support._p(sim[mi, qi]), support._p(dst[mi, qi]))
    ref.splinter_close()
    out.update(c1_masks=np.array(masks, np.uint64), c1_idx=idx, c1_sim_bits=sim.view(np.uint32),
               c1_dist_bits=dst.view(np.uint32), c1_ncand=ncand,
               c1_image_crc=np.array([np.bitwise_xor.reduce(img.view(np.uint64))], np.uint64))

    # ---- the real CLI ----
    qfile = f"/tmp/golden_q_{os.getpid()}.f32"
    q[0].tofile(qfile)
    r = subprocess.run([str(support.REF_DIR / "ref_search_cli"), name, qfile, "golden query", "--json", "--limit", "10",
                        "--bloom", "3"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stderr
    json.loads(r.stdout)
    out["cli_json"] = np.frombuffer(r.stdout.encode(), np.uint8)
    os.unlink(qfile)
    os.unlink("/dev/shm" + name)
    np.savez_compressed(HERE / "search_golden.npz", **out)
    print("wrote", HERE / "search_golden.npz", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
