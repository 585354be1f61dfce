"""CPU emulation of the index logic of csrc/bdilu_solve_skew.cu (time slots, staged layout, table, loaders, drain)
against a plain block sweep.  Run: python scratch/skew_emulate.py"""
import numpy as np
GB=8
def ch(p): return 16-4*p
def cc(p): return ch(p)+(p&1)
def bc(p): return 8*cc(p)+(0 if p&1 else 5)
off=[0]
for p in range(4): off.append(off[-1]+16*GB*bc(p))
PAN=off[4]; VEC=4*GB*64; NARR=4; SLOT=PAN+NARR*VEC; RINV=3*VEC
def emulate(N,nd,blocks,sub,sup,b,fwd,yin=None):
    """blocks[N,32,32] (row,col), sub[d][N,32], sup[d][N,32]; returns the sweep result [N,32]"""
    sm=np.zeros(4*SLOT//8)           # doubles
    out=np.full((N,32),np.nan)
    memblk=(lambda i:i) if fwd else (lambda i:N-1-i)
    def fill(tau):
        slot=(tau&3)*SLOT
        for p in range(4):
            for e in range(8):
                for f in range(8):
                    i=8*(tau-p)+e; blk=memblk(i); col=8*p+f if fwd else 31-8*p-f
                    cb=slot+off[p]+(e*bc(p)+f*cc(p))*16
                    for mm in range(ch(p)):
                        m=4*p+mm if fwd else mm
                        for r in (2*m,2*m+1):
                            ok = r>col if fwd else r<col
                            if ok:
                                v=blocks[blk,r,col] if 0<=blk<N else 0.0
                                pos=r-col-1 if fwd else col-1-r
                                sm[(cb+8*pos)//8]=v
                    t=p*64+e*8+f; row=col
                    inn=0<=blk<N
                    src=b if fwd else yin
                    sm[(slot+PAN+t*8)//8]=src[blk,row] if inn else 0.0
                    for d in range(nd):
                        cr=blk-1-d if fwd else blk
                        cin=inn and 0<=cr<=N-2-d
                        arr=sub if fwd else sup
                        sm[(slot+PAN+VEC*(1+d)+t*8)//8]=arr[d][cr,row] if cin else 0.0
                    if not fwd: sm[(slot+PAN+RINV+t*8)//8]=1.0/blocks[blk,row,row] if inn else 0.0
    def drain(tau):
        slot=(tau&3)*SLOT
        for p in range(4):
            for e in range(8):
                for f in range(8):
                    blk=memblk(8*(tau-p)+e); row=8*p+f if fwd else 31-8*p-f
                    if 0<=blk<N: out[blk,row]=sm[(slot+PAN+(p*64+e*8+f)*8)//8]
    Mtot=(N+62)//32; kfill=4*Mtot
    R=np.zeros((32,32)); v1=np.zeros(32); s1=np.zeros(32); s2=np.zeros(32)
    fill(0); fill(1)
    for T in range(32*Mtot):
        u=T&31; k=T>>3
        if T%8==0:
            if k>=2: drain(k-2)
            if k+2<kfill: fill(k+2)
        newv=np.zeros(32)
        for j in range(32):
            q=(u-j)&31; p=q>>3; f=q&7; e=j&7; a=j>>3
            slot=((a+p)&3)*SLOT
            col=slot+off[p]+(e*bc(p)+f*cc(p))*16; vec=slot+PAN+(p*64+e*8+f)*8
            bv=sm[vec//8]; cf=[sm[(vec+VEC*(1+d))//8] for d in range(nd)]
            t=bv+R[j,u]
            if nd==2: t-=cf[1]*s2[j]
            t-=cf[0]*s1[j]
            v=t if fwd else t*sm[(vec+RINV)//8]
            R[j,u]=0.0; sm[vec//8]=v; newv[j]=v
            for g in range(8):
                valid = q<=30-4*g
                for s in range(4*g,min(4*g+4,31)):
                    a_=sm[(col+8*s)//8] if valid else 0.0
                    R[j,(u+1+s)&31]-=a_*v
        s2=np.array([v1[(j+30)&31] for j in range(32)]); s1=np.array([newv[(j+31)&31] for j in range(32)]); v1=newv
    for k in range(kfill,kfill+2): drain(k-2)
    return out
def ref(N,nd,blocks,sub,sup,b):
    y=np.zeros((N,32))
    for B in range(N):
        for li in range(32):
            s=sum(blocks[B,li,c]*y[B,c] for c in range(li))
            for d in range(1,min(nd,B)+1): s+=sub[d-1][B-d,li]*y[B-d,li]
            y[B,li]=b[B,li]-s
    x=np.zeros((N,32))
    for B in range(N-1,-1,-1):
        for li in range(31,-1,-1):
            s=sum(blocks[B,li,c]*x[B,c] for c in range(li+1,32))
            for d in range(1,min(N-B-1,nd)+1): s+=sup[d-1][B,li]*x[B+d,li]
            x[B,li]=(y[B,li]-s)/blocks[B,li,li]
    return y,x
rng=np.random.default_rng(0)
for N,nd in [(8,1),(9,2),(33,2),(40,1),(71,2)]:
    blocks=rng.uniform(-1,1,(N,32,32))
    for B in range(N):
        for c in range(32): blocks[B,c,c]+=32*np.sign(blocks[B,c,c])
    sub=[rng.uniform(-1,1,(N,32)) for _ in range(nd)]; sup=[rng.uniform(-1,1,(N,32)) for _ in range(nd)]
    b=rng.uniform(-1,1,(N,32))
    y,x=ref(N,nd,blocks,sub,sup,b)
    ye=emulate(N,nd,blocks,sub,sup,b,True)
    xe=emulate(N,nd,blocks,sub,sup,b,False,yin=ye)
    print(N,nd,"fwd err %.2e bwd err %.2e"%(np.abs(ye-y).max(),np.abs(xe-x).max()))
