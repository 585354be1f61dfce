"""Solve-only timing of cfg3-shaped batches (CUDA events), correctness spot check vs the oracle."""
import sys, numpy as np, torch, ctypes as C
from block_diag_ilu_b200 import device as dev, synth, _lib
from oracle import loader
orc=loader.best()
nsys=int(sys.argv[1]) if len(sys.argv)>1 else 64
N,w,d=8192,32,2
data,b=synth.make_batch(4,N,w,d,"D",seed0=1)
D=torch.from_numpy(data).cuda().repeat((nsys+3)//4,1)[:nsys].contiguous(); B=torch.from_numpy(b).cuda().repeat((nsys+3)//4,1,1)[:nsys].contiguous()
ilu=dev.BatchedILU(D,N,w,d,copy=False)
for path in (0,7):
    _lib.lib().bdilu_debug_force_chain_solve(path)
    x=ilu.solve(B); torch.cuda.synchronize()
    ts=[]
    for it in range(5):
        e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
        e0.record(); x=ilu.solve(B); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    ref=data[1].copy(); ipiv,rbc,cbr,info=orc.factor(ref,N,w,d,0,w); xr,fl=orc.solve(ref,N,w,d,0,w,rbc,cbr,b[1,0])
    err=np.abs(x[1,0].cpu().numpy()-xr).max()/np.abs(xr).max()
    print("path",path,"nsys",nsys,"solve ms: min %.3f med %.3f"%(min(ts),sorted(ts)[2]),"err %.1e"%err, flush=True)
out=(C.c_int64*3)(); _lib.lib().bdilu_debug_skew_stats(out,0); print("stats",list(out))
