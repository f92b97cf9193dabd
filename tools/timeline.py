"""Developer aid: per-CTA timeline of one TMA scan (SPL_GPU_DEBUG_TIMELINE=1)."""
import os, sys, ctypes as C
os.environ["SPL_GPU_DEBUG_TIMELINE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, libsplinter_b200 as sg, support
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
m = sg.Mirror.open_dense(n, 768); m.fill_synthetic(1234, 99)
q = support.synth_queries(4321, 1, 768)
lib = sg.load()
for name, mask in (("none", 2**64 - 1), ("0x7", 7), ("0x3", 3), ("bit63", 1 << 63)):
    for _ in range(3): m.search(q, 10, bloom_mask=mask)
    buf = np.zeros((148, 8), np.uint64)
    lib.spl_gpu_debug_timeline(m.handle, buf.ctypes.data_as(C.c_void_p), 148)
    t = buf.astype(np.int64); t0 = t[:, 0].min()
    rel = (t - t0) / 1e3
    last = int(np.argmax(t[:, 3]))
    e = rel[last]
    print(f"   last CTA {last}: consumers done {e[2]:.1f} | cta merge done {e[4]:.1f} | ticket {e[5]:.1f} | lists folded {e[6]:.1f} | final tree {e[7]:.1f} | end {e[3]:.1f};  "
          f"median CTA: consumers->merge {np.median(rel[:,4]-rel[:,2]):.1f}, merge->ticket {np.median(rel[:,5]-rel[:,4]):.1f}")
    print(f"mask {name:6s} start max {rel[:,0].max():7.1f} | producers done: min {rel[:,1].min():7.1f} med {np.median(rel[:,1]):7.1f} max {rel[:,1].max():7.1f} | "
          f"consumers done: min {rel[:,2].min():7.1f} med {np.median(rel[:,2]):7.1f} max {rel[:,2].max():7.1f} | end: med {np.median(rel[:,3]):7.1f} max {rel[:,3].max():7.1f} us")
