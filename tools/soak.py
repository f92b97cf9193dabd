"""Soak: random searches for SECONDS on one mirror; every result is checked against a repeat of itself and a
sample against the oracle.  Looks for rare races (ring hand-back, cross-CTA epilogue) that a single pass can miss."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, libsplinter_b200 as sg, support
secs = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 768
n = 200_003
E, bloom, flags = support.synth_dense(1234, 99, 0, n, dim)
m = sg.Mirror.open_dense(n, dim, flags=sg.OPT_BF16_MIRROR if dim % 64 == 0 else 0); m.dense_upload(0, E, bloom, flags)
rng = np.random.default_rng(int(time.time()))
t_end = time.time() + secs
it = checked = 0
while time.time() < t_end:
    nq = int(rng.choice([1, 1, 1, 2, 3, 4, 5, 8, 20]))
    k = int(rng.choice([1, 5, 10, 32, 33, 100, 256]))
    mask = int(rng.choice([0, 1, 3, 7, 0xFF, 1 << 63]))
    order = int(rng.choice([0, 1]))
    mode = int(rng.choice([sg.MODE_FAST, sg.MODE_FAST, sg.MODE_EXACT, sg.MODE_AUTO]))
    q = support.synth_queries(int(rng.integers(1, 1 << 30)), nq, dim)
    kw = dict(order=order, bloom_mask=mask, mode=mode)
    a, ca = m.search(q, k, **kw)
    b, cb = m.search(q, k, **kw)
    assert np.array_equal(ca, cb), (it, nq, k, mask, order, mode)
    for i in range(nq):
        assert a[i][:ca[i]].tobytes() == b[i][:cb[i]].tobytes(), (it, i, nq, k, mask, order, mode)
    if it % 25 == 0:
        ref, rc = support.oracle_search_dense(q, E, bloom, flags, k=k, order=order, bloom_mask=mask)
        for i in range(nq):
            assert ca[i] == rc[i]
            support.assert_topk_matches(a[i][:ca[i]], ref[i][:rc[i]], exact=(mode == sg.MODE_EXACT), order=order)
        checked += 1
    it += 1
print(f"soak ok: {it} random search pairs in {secs:.0f} s ({checked} oracle checks), dim {dim}")
