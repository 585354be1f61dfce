"""cfg3: assembly M = I - gamma*J followed by the factorisation, as two launches and as the fused launch."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from block_diag_ilu_b200 import device as dev, synth
nsys, N, w, d = int(sys.argv[1]) if len(sys.argv) > 1 else 64, 8192, 32, 2
J, b = synth.make_batch_device(nsys, N, w, d, torch.device("cuda:0"), seed=5)
M = torch.empty_like(J)
def timed(fn, reps=5):
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
def two():
    dev.scale_diag_add(J, -0.01, 1.0, N, w, d, out=M)
    dev.BatchedILU(M, N, w, d, check=False)
def fused():
    dev.BatchedILU.from_scaled(J, -0.01, 1.0, N, w, d, out=M, check=False)
def plain():
    dev.BatchedILU(M, N, w, d, check=False)
print("scale_diag_add + factor: %.3f ms" % timed(two))
print("fused factor_scaled    : %.3f ms" % timed(fused))
dev.scale_diag_add(J, -0.01, 1.0, N, w, d, out=M)
print("factor alone (in place): %.3f ms" % timed(plain, reps=1))
