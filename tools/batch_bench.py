"""Developer timing of the batched tensor-core path at C3 (run on the GPU box)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch, libsplinter_b200 as sg, support
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
nq = int(sys.argv[3]) if len(sys.argv) > 3 else 256
k = int(sys.argv[4]) if len(sys.argv) > 4 else 32
dev = torch.device("cuda", 0)
m = sg.Mirror.open_dense(n, dim, flags=sg.OPT_BF16_MIRROR); m.fill_synthetic(1234, 99)
q_host = support.synth_queries(4321, nq, dim)
q = torch.from_numpy(q_host).to(dev)
out = torch.zeros((nq, k, 32), dtype=torch.uint8, device=dev); cnt = torch.zeros(nq, dtype=torch.int32, device=dev)
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
p = m.params(nq, k, order=sg.ORDER_L2, mode=sg.MODE_BATCH_BF16)
for i in range(3): m.search_device(p, q.data_ptr(), out.data_ptr(), k, cnt.data_ptr(), st.cuda_stream)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
iters = 10
e0.record()
for i in range(iters): m.search_device(p, q.data_ptr(), out.data_ptr(), k, cnt.data_ptr(), st.cuda_stream)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
flops = 2.0 * nq * n * dim
print(f"n={n} dim={dim} nq={nq} k={k}: {ms:.3f} ms/batch  {nq/ms*1e3:.0f} q/s  {flops/ms/1e9:.1f} TFLOP/s (algorithmic, both passes counted once)  counts min {int(cnt.min())}")
# parity spot check of 4 queries against the oracle on a 2^17 prefix is done by the tests; here: self-consistency vs FAST scan
res = out.cpu().numpy().view(sg.RESULT_DTYPE).reshape(nq, k)
o2, c2 = m.search(q_host[:4], k, order=sg.ORDER_L2, mode=sg.MODE_FAST)
for i in range(4):
    same = np.array_equal(res[i]["slot_idx"], o2[i]["slot_idx"])
    print("  query", i, "indices identical to fp32 scan:", same, " max|d dist|", float(np.abs(res[i]["distance"] - o2[i]["distance"]).max()))
