import numpy as np, sys, torch
sys.path.insert(0,'.')
from oracle.loader import Port, Ref
from block_diag_ilu_b200 import synth, device as dev
P,R=Port(),Ref()
N,w,d=2048,32,2
data,b=synth.make_batch(1,N,w,d,'P',seed0=77)
d1,d2=data[0].copy(),data[0].copy()
P.factor(d1,N,w,d,0,w); ip,_,_,_=R.factor(d2,N,w,d,0,w)
ilu=dev.BatchedILU(torch.from_numpy(data).cuda(),N,w,d,copy=True)
g=ilu.data[0].cpu().numpy()
nb=N*w*w
G=g[:nb].reshape(N,-1); r=d2[:nb].reshape(N,-1); p=d1[:nb].reshape(N,-1)
e=np.abs(G-r).max(axis=1)/np.abs(r).max(axis=1)
print('gpu vs ref max',e.max(),'n>1e-12',(e>1e-12).sum(),'n>1e-13',(e>1e-13).sum(), 'median',np.median(e))
ep=np.abs(G-p).max(axis=1)/np.abs(p).max(axis=1)
print('gpu vs port max',ep.max())
k=e.argmax()
E=np.abs(G[k]-r[k]).reshape(w,w).T  # [r,c]
print('worst block',k,'err by column max',E.max(axis=0))
print('err by row max',E.max(axis=1))
print('ipiv',ip[k*w:(k+1)*w])
# long double truth for that block
A=data[0][:nb].reshape(N,w,w)[k].T.astype(np.longdouble).copy()
piv=ip[k*w:(k+1)*w]-1
for i in range(w):
    j=piv[i]
    if j!=i: A[[i,j]]=A[[j,i]]
    A[i+1:,i]/=A[i,i]
    A[i+1:,i+1:]-=np.outer(A[i+1:,i],A[i,i+1:])
T=A.astype(np.float64)
print('truth vs gpu',np.abs(T-G[k].reshape(w,w).T).max(),'truth vs ref',np.abs(T-r[k].reshape(w,w).T).max(),'truth vs port',np.abs(T-p[k].reshape(w,w).T).max())
print('diagU', np.abs(np.diag(T)).min())
