"""Phase profile of the wave32 solve (build with BDILU_NVCC_EXTRA=-DBDILU_W32_PROF)."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from block_diag_ilu_b200 import device as dev, synth, _lib
nsys = int(sys.argv[1]) if len(sys.argv) > 1 else 64
N, w, d = 8192, 32, 2
data, b = synth.make_batch_device(nsys, N, w, d, torch.device("cuda:0"), seed=5)
L = _lib.lib()
ilu = dev.BatchedILU(data.clone(), N, w, d)
x = ilu.solve(b); torch.cuda.synchronize()
out = (C.c_ulonglong * 8)()
L.bdilu_debug_q32prof(out, 1)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); x = ilu.solve(b); e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
L.bdilu_debug_q32prof(out, 0)
v = list(out)
nb = v[7]
names = ["wait_stage", "to_regs", "first_poll", "chain", "store", "prod_wait_empty", "prod_issue"]
print("solve %.3f ms = %.0f cycles @1.965GHz; blocks counted %d (expect %d)" % (ms, ms * 1.965e6, nb, 2 * nsys * N))
for n, c in zip(names[:5], v[:5]):
    print("  consumer %-12s %8.1f cycles/block" % (n, c / nb))
for n, c in zip(names[5:], v[5:7]):
    print("  producer %-16s %8.1f cycles/block" % (n, c / (2 * nsys * N)))
