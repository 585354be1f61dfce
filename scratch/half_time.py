"""blockw-32 solve: two-SM kernel (path 10) against quad32 (path 0).  usage: python scratch/half_time.py [nsys,N,d,nrhs ...]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from block_diag_ilu_b200 import device as dev, synth, _lib
L = _lib.lib()
CASES = [(2, 256, 2, 1), (8, 2048, 2, 1), (64, 8192, 2, 1), (32, 8192, 2, 1), (64, 8192, 1, 1), (16, 8192, 2, 4)]
if len(sys.argv) > 1:
    CASES = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]]
for (nsys, N, d, nrhs) in CASES:
    data, b = synth.make_batch_device(nsys, N, 32, d, torch.device("cuda:0"), seed=5, nrhs=nrhs)
    ilu = dev.BatchedILU(data, N, 32, d)
    res, xs = {}, {}
    for name, path in (("quad32", 0), ("half32", 10)):
        old = L.bdilu_debug_force_chain_solve(path)
        best = 1e9
        for _ in range(4):
            e1, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e1.record(); x = ilu.solve(b); e2.record(); torch.cuda.synchronize()
            best = min(best, e1.elapsed_time(e2))
        L.bdilu_debug_force_chain_solve(old)
        res[name], xs[name] = best, x
        assert int(ilu.flags.abs().max()) == 0
    print("nsys %d N %d d %d nrhs %d : " % (nsys, N, d, nrhs) + "  ".join("%s %.3f ms" % kv for kv in res.items()) +
          "  | bit-identical: %s" % bool(torch.equal(xs["quad32"], xs["half32"])), flush=True)
