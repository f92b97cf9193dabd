"""Developer timing of spl_gpu_group_search over different device lists (run on a multi-GPU box)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch, libsplinter_b200 as sg, support
ndev = torch.cuda.device_count()
per = 1 << 22
q = support.synth_queries(4321, 8, 768)
for devs in ([0], [0, 0], list(range(min(2, ndev))), list(range(ndev))):
    with sg.Group.open_dense(per * len(devs), 768, devs) as g:
        g.fill_synthetic(1234, 99)
        for k in (10, 100):
            for i in range(3): g.search(q[i], k, mode=sg.MODE_FAST)
            t0 = time.perf_counter(); n = 20
            for i in range(n): g.search(q[i % 8], k, mode=sg.MODE_FAST)
            ms = (time.perf_counter() - t0) / n * 1e3
            print(f"devices {devs} k={k}: {ms:.3f} ms per query ({per * len(devs)} slots)")
