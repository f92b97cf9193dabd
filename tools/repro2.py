import os, sys, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
if len(sys.argv) == 1:
    for rep in range(2):
      for dim in (1024,):
        for data in ("random", "dup"):
            for order in (0, 1):
                for k in (10, 16, 17, 32):
                    try:
                        r = subprocess.run([sys.executable, __file__, str(dim), data, str(order), str(k)], capture_output=True, text=True, timeout=25)
                        print(dim, data, "order", order, "k", k, "->", (r.stdout.strip().splitlines() or [r.stderr.strip()[-80:]])[-1][:70], "| rc", r.returncode, flush=True)
                    except subprocess.TimeoutExpired:
                        print(dim, data, "order", order, "k", k, "-> HANG", flush=True)
    sys.exit(0)
import numpy as np, libsplinter_b200 as sg
dim, data, order, k = int(sys.argv[1]), sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
n = 20_000
rng = np.random.default_rng(3)
base = rng.standard_normal(dim).astype(np.float32)
E = np.tile(base, (n, 1)) if data == "dup" else rng.standard_normal((n, dim)).astype(np.float32)
q = (base[None, :] + rng.standard_normal((2, dim)).astype(np.float32) * 0.5).astype(np.float32)
with sg.Mirror.open_dense(n, dim) as m:
    m.dense_upload(0, E, np.zeros(n, np.uint64), np.full(n, 7, np.uint8))
    out, c = m.search(q, k, mode=sg.MODE_FAST, order=order)
    print("ok", c, out[0]["slot_idx"][:4])
