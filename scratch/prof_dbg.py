"""Timing-only probe of the solve kernel under BDILU_DBG variants (results are garbage for dbg != 0)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from block_diag_ilu_b200 import device as dev, synth
nsys, N, w, d = 64, 8192, 32, 2
data, b = synth.make_batch_device(nsys, N, w, d, torch.device("cuda:0"), seed=5)
ilu = dev.BatchedILU(data, N, w, d)
for _ in range(3):
    e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
    e1.record()
    x = ilu.solve(b)
    e2.record()
    torch.cuda.synchronize()
    print("dbg=%s solve %.3f ms" % (os.environ.get("BDILU_DBG", "0"), e1.elapsed_time(e2)))
