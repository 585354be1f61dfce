"""Narrow-block solve timing: the prefetched chain kernel (path 9) against the automatic choice (path 0) and the general
chain kernel (path 1).  usage: python scratch/pf_time.py [nsys,N,w,d,nrhs ...]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from block_diag_ilu_b200 import device as dev, synth, _lib

L = _lib.lib()
CASES = [  # nsys, N, w, d, nrhs
    (1, 64, 4, 1, 1), (1, 1024, 16, 1, 1), (64, 512, 8, 2, 1), (512, 512, 8, 2, 1), (1024, 512, 8, 2, 1),
    (2048, 512, 8, 2, 1), (4096, 512, 8, 2, 1), (4096, 512, 8, 2, 8), (296, 1024, 16, 1, 1), (1200, 1024, 16, 1, 1),
    (1, 1024, 16, 1, 64), (16, 1024, 16, 1, 64), (4096, 64, 4, 1, 1), (100000, 64, 4, 1, 1), (1, 4096, 8, 2, 1),
]
if len(sys.argv) > 1:
    CASES = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]]
for (nsys, N, w, d, nrhs) in CASES:
    data, b = synth.make_batch_device(nsys, N, w, d, torch.device("cuda:0"), seed=5, nrhs=nrhs)
    ilu = dev.BatchedILU(data, N, w, d)
    res, xs = {}, {}
    for name, path in (("auto", 0), ("pf", 12), ("pf4", 11), ("chain", 1)):
        old = L.bdilu_debug_force_chain_solve(path)
        best = 1e9
        for _ in range(4):
            e1, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e1.record(); x = ilu.solve(b); e2.record(); torch.cuda.synchronize()
            best = min(best, e1.elapsed_time(e2))
        L.bdilu_debug_force_chain_solve(old)
        res[name], xs[name] = best, x
    same4 = bool(torch.equal(xs["pf"], xs["pf4"]))
    same = float(((xs["pf"] - xs["chain"]).abs().amax(dim=-1) / xs["chain"].abs().amax(dim=-1)).max())
    print("nsys %d N %d w %d d %d nrhs %d : " % (nsys, N, w, d, nrhs) + "  ".join("%s %.3f ms" % kv for kv in res.items()) +
          "  | pf vs chain max rel diff: %.2e, pf4 bit-identical to pf: %s" % (same, same4), flush=True)
    del data, b, ilu, xs, x
    torch.cuda.empty_cache()
