"""Solve-kernel timing on many right-hand sides: the many-vector kernel (path 4) against the automatic
one-vector choice (path 5) and the chain kernel (path 1).  usage: python scratch/mrhs_time.py"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from block_diag_ilu_b200 import device as dev, synth, _lib

L = _lib.lib()
CASES = [  # nsys, N, w, d, nrhs
    (1, 2048, 64, 3, 256),   # cfg4
    (1, 2048, 64, 3, 32),
    (1, 2048, 64, 3, 8),
    (64, 8192, 32, 2, 8),
    (8, 8192, 32, 2, 64),
    (64, 8192, 32, 2, 32),
    (4096, 512, 8, 2, 8),
    (1, 1024, 16, 1, 64),
]
if len(sys.argv) > 1:
    CASES = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]]
for (nsys, N, w, d, nrhs) in CASES:
    data, b = synth.make_batch_device(nsys, N, w, d, torch.device("cuda:0"), seed=5, nrhs=nrhs)
    ilu = dev.BatchedILU(data, N, w, d)
    res = {}
    xs = {}
    for name, path in (("mrhs", 4), ("mrhs_wave", 6), ("auto_no_mrhs", 5), ("chain", 1)):
        if name == "chain" and nsys * nrhs * N * w * w > 3e10:
            continue
        old = L.bdilu_debug_force_chain_solve(path)
        best = 1e9
        for _ in range(3):
            e1, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e1.record(); x = ilu.solve(b); e2.record(); torch.cuda.synchronize()
            best = min(best, e1.elapsed_time(e2))
        L.bdilu_debug_force_chain_solve(old)
        res[name] = best
        xs[name] = x
    ref = xs["auto_no_mrhs"]
    dev_rel = float(((xs["mrhs_wave"] - ref).abs().amax(dim=-1) / ref.abs().amax(dim=-1)).max())
    flops = (2 * w * w + 4 * d * w) * N * nrhs * nsys
    print("nsys %d N %d w %d d %d nrhs %d : " % (nsys, N, w, d, nrhs) +
          "  ".join("%s %.3f ms" % kv for kv in res.items()) +
          "  | mrhs_wave %.2f TFLOP/s, max rel diff vs one-vector path %.2e" % (flops / res["mrhs_wave"] / 1e9, dev_rel), flush=True)
    del data, b, ilu, xs, x, ref
    torch.cuda.empty_cache()
