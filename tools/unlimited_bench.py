"""k = 0 (the reference's default: every survivor, sorted) and large-k timings."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, libsplinter_b200 as sg, support
n = 1 << 20
m = sg.Mirror.open_dense(n, 768); m.fill_synthetic(1234, 99)
q = support.synth_queries(4321, 4, 768)
for mode, mname in ((sg.MODE_FAST, "FAST"), (sg.MODE_EXACT, "EXACT")):
    for name, mask, k in (("0x7 k=0", 7, 0), ("0x3 k=0", 3, 0), ("all k=0", 1 << 63, 0), ("0x3 k=100", 3, 100), ("0x3 k=256", 3, 256), ("0x3 k=1000", 3, 1000)):
        stride = n if k == 0 else k
        for _ in range(2): out, cnt = m.search(q[0], k, bloom_mask=mask, mode=mode, out_stride=stride)
        t0 = time.perf_counter(); reps = 5
        for r in range(reps): out, cnt = m.search(q[r % 4], k, bloom_mask=mask, mode=mode, out_stride=stride)
        dt = (time.perf_counter() - t0) / reps
        print(f"{mname:5s} {name:11s}: {dt*1e3:8.3f} ms e2e, {int(cnt[0])} results")
