import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, libsplinter_b200 as sg
step = sys.argv[1] if len(sys.argv) > 1 else "batch"
n, dim = 20_000, 1024
rng = np.random.default_rng(3)
base = rng.standard_normal(dim).astype(np.float32)
E = np.tile(base, (n, 1)); E[5000] = base * 0.5; E[17] = -base
q = (base[None, :] + rng.standard_normal((4, dim)).astype(np.float32) * 0.5).astype(np.float32)
with sg.Mirror.open_dense(n, dim, flags=sg.OPT_BF16_MIRROR) as m:
    m.dense_upload(0, E, np.zeros(n, np.uint64), np.full(n, 7, np.uint8))
    if step == "fast":
        out, c = m.search(q, 16, mode=sg.MODE_FAST, order=sg.ORDER_L2); print("fast ok", c, out[0]["slot_idx"][:5])
    else:
        out, c = m.search(q, 16, mode=sg.MODE_BATCH_BF16, order=sg.ORDER_L2); print("batch ok", c, out[0]["slot_idx"][:5])
