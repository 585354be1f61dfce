"""Repeatability stress of the wavefront group kernel: the same solve many times must give bit-identical x
(a race in the publication protocol would show up as a rare difference), and must match the chain kernel to 1e-12."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from block_diag_ilu_b200 import device as dev, synth, _lib
L = _lib.lib()
bad = 0
for (nsys, N, w, d, nrhs, fam) in [(1, 2048, 64, 3, 256, "D"), (3, 700, 64, 4, 20, "P"), (8, 4096, 32, 2, 16, "D"), (5, 1500, 32, 3, 9, "P"),
                                   (16, 1024, 16, 1, 64, "D"), (7, 900, 16, 4, 3, "P"), (300, 512, 8, 2, 1, "D"), (40, 600, 8, 1, 12, "P")]:
    data, b = synth.make_batch_device(nsys, N, w, d, torch.device("cuda:0"), seed=11, nrhs=nrhs, family=fam)
    ilu = dev.BatchedILU(data, N, w, d, check=False)
    L.bdilu_debug_force_chain_solve(1)
    ref = ilu.solve(b).clone()
    L.bdilu_debug_force_chain_solve(6)
    first = None
    for rep in range(40):
        x = ilu.solve(b)
        if first is None:
            first = x.clone()
            rel = float(((x - ref).abs().amax(dim=-1) / ref.abs().amax(dim=-1)).max())
        elif not torch.equal(x, first):
            bad += 1
    L.bdilu_debug_force_chain_solve(0)
    print("nsys %d N %d w %d d %d nrhs %d %s: vs chain %.2e, non-identical repeats %d" % (nsys, N, w, d, nrhs, fam, rel, bad), flush=True)
print("STRESS", "FAILED" if bad else "ok")
