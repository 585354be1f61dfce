"""K1 on inputs that do / do not pivot: the optimistic path (family D) against its fallback (uniform random blocks: the
optimistic attempt is abandoned at its first vote, then the general loop runs).  usage: python scratch/factor_families.py"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from block_diag_ilu_b200 import device as dev, synth
for (nsys, N, w, d) in [(64, 8192, 32, 2), (4096, 512, 8, 2)]:
    for fam in ("D", "P"):
        data, b = synth.make_batch_device(nsys, N, w, d, torch.device("cuda:0"), seed=5, family=fam)
        best = 1e9
        for _ in range(4):
            work = data.clone()
            e1, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e1.record(); ilu = dev.BatchedILU(work, N, w, d, check=False); e2.record(); torch.cuda.synchronize()
            best = min(best, e1.elapsed_time(e2))
        nonid = float((ilu.ipiv != torch.arange(1, w + 1, device="cuda", dtype=torch.int32).repeat(N)).float().mean())
        print("nsys %d N %d w %d family %s: factor %.3f ms, non-identity pivots %.0f %%" % (nsys, N, w, fam, best, 100 * nonid), flush=True)
        del data, b, ilu, work
        torch.cuda.empty_cache()
