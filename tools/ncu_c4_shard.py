"""ncu target: one GPU's shard of C4 at N = 2 (2^25 rows x 768, unfiltered cosine top-100)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import libsplinter_b200 as sg, support
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 25
m = sg.Mirror.open_dense(rows, 768); m.fill_synthetic(1234, 99)
q = support.synth_queries(4321, 1, 768)
m.set_timing(True)
for _ in range(3):
    out, cnt = m.search(q, 100, bloom_mask=0, mode=sg.MODE_FAST)
print("ok", int(cnt[0]), "ms", m.last_scan_ms())
