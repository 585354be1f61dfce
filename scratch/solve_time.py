import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from block_diag_ilu_b200 import device as dev, synth
nsys = int(sys.argv[1]); N = 8192; w = 32; d = 2
data, b = synth.make_batch_device(nsys, N, w, d, torch.device("cuda:0"), seed=5)
ilu = dev.BatchedILU(data, N, w, d)
for _ in range(3):
    e1, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e1.record(); x = ilu.solve(b); e2.record(); torch.cuda.synchronize()
print("nsys %d solve %.3f ms dbg=%s" % (nsys, e1.elapsed_time(e2), os.environ.get("BDILU_QUAD_DBG")))
