import os, sys, ctypes as C
os.environ["SPL_GPU_DEBUG_TIMELINE"] = "1"
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np, libsplinter_b200 as sg, support
m = sg.Mirror.open_dense(1 << 20, 768); m.fill_synthetic(1234, 99)
q = support.synth_queries(4321, 4, 768)
lib = sg.load()
for nq in (1, 2, 4):
  for name, mask in (("0x3", 3), ("bit63", 1 << 63)):
    for _ in range(3): m.search(q[:nq], 10, bloom_mask=mask, mode=sg.MODE_FAST)
    buf = np.zeros((148, 8), np.uint64)
    lib.spl_gpu_debug_timeline(m.handle, buf.ctypes.data_as(C.c_void_p), 148)
    t = buf.astype(np.int64); t0 = t[:, 0].min(); rel = (t - t0) / 1e3
    last = int(np.argmax(np.where(t[:, 6] > t[:, 5], rel[:, 3], -1)))   # the finishing CTA that took query 0 (only it stamps 6 and 7)
    print(f"nq {nq} mask {name:6s} producers done med {np.median(rel[:,1]):7.1f} max {rel[:,1].max():7.1f} | consumers done med {np.median(rel[:,2]):7.1f} max {rel[:,2].max():7.1f} | end med {np.median(rel[:,3]):7.1f} max {rel[:,3].max():7.1f} | per CTA after the scan: lists published +{np.median(rel[:,4]-rel[:,2]):.1f} ticket +{np.median(rel[:,5]-rel[:,4]):.1f} us | last CTA: ticket at {rel[:,5].max():.1f}, selection +{(rel[last,6]-rel[last,5]):.1f}, merge +{(rel[last,7]-rel[last,6]):.1f}, records +{(rel[last,3]-rel[last,7]):.1f}")
