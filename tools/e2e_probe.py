import os, sys, time
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np, libsplinter_b200 as sg, support
m = sg.Mirror.open_dense(1 << 20, 768); m.fill_synthetic(1234, 99)
q = support.synth_queries(4321, 8, 768)
for mask in (3, 2**64 - 1):
    for i in range(50): m.search(q[i % 8], 10, bloom_mask=mask)
    t0 = time.perf_counter()
    for i in range(400): m.search(q[i % 8], 10, bloom_mask=mask)
    print("mask", hex(mask), "e2e us/query", (time.perf_counter() - t0) / 400 * 1e6)
