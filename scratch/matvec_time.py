"""dot_vec / residual timing over many vectors (cfg4 shape and cfg3 shape)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from block_diag_ilu_b200 import device as dev, synth
for (nsys, N, w, d, nrhs) in [(1, 2048, 64, 3, 256), (1, 2048, 64, 3, 1), (64, 8192, 32, 2, 1), (8, 8192, 32, 2, 16)]:
    data, b = synth.make_batch_device(nsys, N, w, d, torch.device("cuda:0"), seed=5, nrhs=nrhs)
    best = 1e9
    for _ in range(4):
        e1, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e1.record(); out = dev.dot_vec(data, b, N, w, d); e2.record(); torch.cuda.synchronize()
        best = min(best, e1.elapsed_time(e2))
    byt = nsys * (8 * (N * w * w + 2 * (d * N) * w) + 16 * N * w * nrhs)
    print("nsys %d N %d w %d d %d nrhs %d : dot_vec %.3f ms = %.0f GB/s of algorithmic bytes" % (nsys, N, w, d, nrhs, best, byt / best / 1e6), flush=True)
