"""k = 0 through the C call with caller-owned buffers: pageable (touched) vs pinned output, 2^20 x 768."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch, libsplinter_b200 as sg, support
n = 1 << 20
m = sg.Mirror.open_dense(n, 768); m.fill_synthetic(1234, 99)
m.set_timing(True)
q = support.synth_queries(4321, 4, 768)
counts = np.zeros(1, np.uint32)
pageable = np.zeros(n * 32, np.uint8); pageable[:] = 1
pinned = torch.zeros(n * 32, dtype=torch.uint8).pin_memory().numpy()
for name, mask in (("0x3", 3), ("unfiltered", 0)):
    p = m.params(1, 0, mode=sg.MODE_FAST, bloom_mask=mask)
    for bname, buf in (("pageable", pageable), ("pinned", pinned)):
        for _ in range(2): m.search_into(p, q[0], buf, n, counts)
        ts = []
        for r in range(7):
            t0 = time.perf_counter(); m.search_into(p, q[r % 4], buf, n, counts); ts.append(time.perf_counter() - t0)
        print(f"mask {name:10s} out {bname:8s}: {np.median(ts)*1e3:7.3f} ms e2e (min {min(ts)*1e3:.3f}), scan {m.last_scan_ms():.3f} ms, {int(counts[0])} results = {int(counts[0])*32/1e6:.1f} MB")
