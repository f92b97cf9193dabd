"""ncu target: a few scans of nq queries (usage: ncu_small_batch.py [mask] [nq]); tools/ncu_summary.py reads the report."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import libsplinter_b200 as sg, support
m = sg.Mirror.open_dense(1 << 20, 768); m.fill_synthetic(1234, 99)
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 4
q = support.synth_queries(4321, nq, 768)
mask = int(sys.argv[1], 0) if len(sys.argv) > 1 else 1 << 63
for _ in range(3):
    m.search(q, 10, bloom_mask=mask, mode=sg.MODE_FAST)
print("ok")
