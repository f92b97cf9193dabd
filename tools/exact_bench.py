import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch, libsplinter_b200 as sg, support
n, dim, k = 1 << 20, 768, 10
m = sg.Mirror.open_dense(n, dim); m.fill_synthetic(1234, 99)
q = support.synth_queries(4321, 16, dim)
for mode, name in ((sg.MODE_FAST, "fast"), (sg.MODE_EXACT, "exact")):
    for mask in (3, 0):
        for i in range(5): m.search(q[i % 16], k, bloom_mask=mask, mode=mode)
        t0 = time.perf_counter()
        for i in range(100): out, c = m.search(q[i % 16], k, bloom_mask=mask, mode=mode)
        dt = (time.perf_counter() - t0) / 100 * 1e6
        print(f"{name:6s} mask {mask}: e2e {dt:7.1f} us/query")
E, bloom, flags = support.synth_dense(1234, 99, 0, 1 << 17, dim)
with sg.Mirror.open_dense(1 << 17, dim) as ms:
    ms.dense_upload(0, E, bloom, flags)
    out, c = ms.search(q[:8], k, bloom_mask=3, mode=sg.MODE_EXACT)
    ref, rc = support.oracle_search_dense(q[:8], E, bloom, flags, k=k, bloom_mask=3)
    for i in range(8): support.assert_topk_matches(out[i][:c[i]], ref[i][:rc[i]], exact=True)
    print("exact parity on 2^17 slots: bit-identical")
