import sys, numpy as np, torch, ctypes as C
from block_diag_ilu_b200 import device as dev, synth, _lib
N,w,d=8192,32,2
data,b=synth.make_batch(2,N,w,d,"D",seed0=1)
ilu=dev.BatchedILU(torch.from_numpy(data).cuda(),N,w,d,copy=False)
B=torch.from_numpy(b).cuda()
x=ilu.solve(B); torch.cuda.synchronize()
L=_lib.lib()
buf=np.zeros((4,512,4),dtype=np.int64)
L.bdilu_debug_skew_prof.argtypes=[C.c_void_p]; L.bdilu_debug_skew_prof(buf.ctypes.data)
# stamps hold the backward sweep of CTA 0 (last written)
t0=buf[0,100,0]
print("k | near: start, after far-wait, at step7, after full-wait | far: start, after near-wait, after full | loader: start, after empty-wait, after store, after load-issue")
for k in range(100,116):
    print(k, (buf[0,k]-t0).tolist(), (buf[1,k]-t0).tolist()[:3], (buf[2,k]-t0).tolist())
print("near group-step period", np.diff(buf[0,100:200,0]).mean())
