"""Phase profile of the narrow-block chain kernel (debug build with -DBDILU_PF_PROF, loaded through BDILU_LIB_PATH)."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from block_diag_ilu_b200 import device as dev, synth, _lib
L = _lib.lib()
L.bdilu_debug_force_chain_solve(9)
for (nsys, N, w, d) in [(1, 64, 4, 1), (64, 512, 8, 2), (1, 1024, 16, 1), (4096, 512, 8, 2)]:
    data, b = synth.make_batch_device(nsys, N, w, d, torch.device("cuda:0"), seed=5)
    ilu = dev.BatchedILU(data, N, w, d)
    ilu.solve(b); torch.cuda.synchronize()
    out = (C.c_ulonglong * 8)()
    L.bdilu_debug_pfprof(out, 1)
    ilu.solve(b); torch.cuda.synchronize()
    L.bdilu_debug_pfprof(out, 1)
    n = max(out[4], 1)
    print("nsys %d N %d w %d d %d: per forward block (cycles): wait+read %.0f  gather+coupling %.0f  rows %.0f  store+issue+prefetch %.0f  (blocks %d)"
          % (nsys, N, w, d, out[0] / n, out[1] / n, out[2] / n, out[3] / n, n))
