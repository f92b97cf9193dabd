"""Small-batch scan throughput: nq queries per spl_gpu_search call (device time via CUDA events + end to end)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, libsplinter_b200 as sg, support
n, dim = 1 << 20, int(sys.argv[1]) if len(sys.argv) > 1 else 768
k = int(sys.argv[2]) if len(sys.argv) > 2 else 10
m = sg.Mirror.open_dense(n, dim); m.fill_synthetic(1234, 99)
m.set_timing(True)
q = support.synth_queries(4321, 64, dim)
for name, mask in (("0x3", 3), ("dense", 1 << 63)):
    for nq in (1, 2, 4, 8, 16):
        for _ in range(5): m.search(q[:nq], k, bloom_mask=mask, mode=sg.MODE_FAST)
        dev = []
        t0 = time.perf_counter()
        reps = 30
        for r in range(reps):
            m.search(q[(r % 4) * nq:(r % 4) * nq + nq], k, bloom_mask=mask, mode=sg.MODE_FAST)
            dev.append(m.last_scan_ms() * nq)
        e2e = (time.perf_counter() - t0) / reps
        d = float(np.median(dev))
        print(f"mask {name:5s} nq {nq:2d}: device {d*1e3:8.1f} us/batch = {d*1e3/nq:7.1f} us/query ({nq/d*1e3:9.0f} q/s) | e2e {e2e*1e6:8.1f} us/batch ({nq/e2e:9.0f} q/s)")
