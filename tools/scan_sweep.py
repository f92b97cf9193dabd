"""Developer sweep: device time of one scan vs label-mask selectivity (run on the GPU box)."""
import sys, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import libsplinter_b200 as sg, support
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
k = int(sys.argv[2]) if len(sys.argv) > 2 else 10
dim = 768
dev = torch.device("cuda", 0)
m = sg.Mirror.open_dense(n, dim); m.fill_synthetic(1234, 99)
_, bloom, flags, _ = m.read_mirror(0, n, want_E=False)
q = torch.from_numpy(support.synth_queries(4321, 8, dim)).to(dev)
out = torch.zeros((k, 32), dtype=torch.uint8, device=dev); cnt = torch.zeros(1, dtype=torch.int32, device=dev)
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
for name, mask in (("none", 0xFFFFFFFFFFFFFFFF), ("0xff", 0xFF), ("0x7", 7), ("0x3", 3), ("0x1", 1), ("bit63", 1 << 63), ("zero", 0)):
    p = m.params(1, k, bloom_mask=mask)
    need = 2 if mask else 1
    npass = int((((flags & need) != 0) & ((bloom & np.uint64(mask)) == np.uint64(mask))).sum())
    for i in range(10): m.search_device(p, q[i % 8].data_ptr(), out.data_ptr(), k, cnt.data_ptr(), st.cuda_stream)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(50): m.search_device(p, q[i % 8].data_ptr(), out.data_ptr(), k, cnt.data_ptr(), st.cuda_stream)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 50 * 1e3
    b = 9 * n + 3072 * npass
    print(f"mask {name:6s} pass {npass:8d} ({100*npass/n:5.1f}%)  {us:8.1f} us  {b/us/1e3:8.1f} GB/s  rows/us {npass/us:7.1f}")
