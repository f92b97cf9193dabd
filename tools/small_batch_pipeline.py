"""4-query calls in a loop on one stream, with and without SPL_SEARCH_PIPELINED (device time per call, CUDA events)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, libsplinter_b200 as sg, support
dev = torch.device("cuda:0")
m = sg.Mirror.open_dense(1 << 20, 768); m.fill_synthetic(1234, 99)
k = 10
ts = torch.cuda.Stream(); torch.cuda.set_stream(ts); stream = ts.cuda_stream
for nq in (1, 2, 4):
    q = torch.from_numpy(support.synth_queries(4321, nq, 768)).to(dev)
    out = torch.zeros((nq * k, 32), dtype=torch.uint8, device=dev); cnt = torch.zeros(nq, dtype=torch.int32, device=dev)
    for name, mask in (("0x3", 3), ("dense", 1 << 63)):
        for flags in (0, sg.SEARCH_PIPELINED):
            p = m.params(nq, k, mode=sg.MODE_FAST, bloom_mask=mask, flags=flags)
            for _ in range(10): m.search_device(p, q.data_ptr(), out.data_ptr(), k, cnt.data_ptr(), stream)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(200): m.search_device(p, q.data_ptr(), out.data_ptr(), k, cnt.data_ptr(), stream)
            e1.record(); torch.cuda.synchronize()
            print(f"nq {nq} mask {name:5s} pipelined {int(bool(flags))}: {e0.elapsed_time(e1) / 200 * 1e3:7.1f} us per call")
