import sys, time, numpy as np, torch, ctypes as C
from block_diag_ilu_b200 import device as dev, synth, _lib
from oracle import loader
orc=loader.best()
def st(reset=False):
    out=(C.c_int64*3)(); _lib.lib().bdilu_debug_skew_stats(out,int(reset)); return list(out)
for (nsys,N,d,nrhs) in [(2,8,1,1),(2,8,2,1),(2,9,1,1),(2,33,2,1),(3,77,2,3),(2,1000,2,1),(4,8192,2,1)]:
    w=32
    data,b=synth.make_batch(nsys,N,w,d,"D",seed0=N,nrhs=nrhs)
    st(True)
    ilu=dev.BatchedILU(torch.from_numpy(data).cuda(),N,w,d,copy=True)
    bb=torch.from_numpy(b).cuda()
    x=ilu.solve(bb); torch.cuda.synchronize()
    t0=time.time(); x=ilu.solve(bb); torch.cuda.synchronize(); t1=time.time()
    errs=[]
    for s in range(nsys):
        ref=data[s].copy(); ipiv,rbc,cbr,info=orc.factor(ref,N,w,d,0,w)
        for r in range(nrhs):
            xr,fl=orc.solve(ref,N,w,d,0,w,rbc,cbr,b[s,r]); errs.append(np.abs(x[s,r].cpu().numpy()-xr).max()/np.abs(xr).max())
    print((nsys,N,d,nrhs),"stats",st(),"maxerr %.2e"%max(errs),"flags",int(ilu.flags.abs().max()),"t=%.3f ms"%((t1-t0)*1e3),flush=True)
