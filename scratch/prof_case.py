"""Small profiling case: cfg3 shape, few systems; one factor launch + one solve launch (+ optional repeat)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from block_diag_ilu_b200 import device as dev, synth
nsys = int(sys.argv[1]) if len(sys.argv) > 1 else 8
N = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
w = int(sys.argv[3]) if len(sys.argv) > 3 else 32
d = int(sys.argv[4]) if len(sys.argv) > 4 else 2
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 2
nrhs = int(sys.argv[6]) if len(sys.argv) > 6 else 1
data, b = synth.make_batch_device(nsys, N, w, d, torch.device("cuda:0"), seed=5, nrhs=nrhs)
for _ in range(reps):
    work = data.clone()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record()
    ilu = dev.BatchedILU(work, N, w, d)
    e1.record()
    x = ilu.solve(b)
    e2.record()
    torch.cuda.synchronize()
    print("factor %.3f ms  solve %.3f ms" % (e0.elapsed_time(e1), e1.elapsed_time(e2)))
assert int(ilu.flags.abs().max()) == 0
print("ok", float(x.abs().max()))
