// Micro-benchmarks that decide the layout of the systolic solve kernel (round 2).
//  T1: cp.async.bulk of small pieces, 32 lanes issuing one copy each per instruction
//  T3: LDS.64 wavefront cost for per-lane address patterns (conflict-free / panel layout / predicated lanes)
//  T4: dependent-chain latencies (DFMA, SHFL+DFMA, STS->LDS)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1;}}while(0)
__device__ __forceinline__ unsigned s32(const void*p){return (unsigned)__cvta_generic_to_shared(p);}
__device__ __forceinline__ void mbar_init(void*b,unsigned c){asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;"::"r"(s32(b)),"r"(c));}
__device__ __forceinline__ void mbar_expect(void*b,unsigned n){asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"::"r"(s32(b)),"r"(n):"memory");}
__device__ __forceinline__ bool mbar_try(void*b,unsigned ph){unsigned ok;asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0,1,0,p;\n}\n":"=r"(ok):"r"(s32(b)),"r"(ph):"memory");return ok;}
__device__ __forceinline__ void bulk(void*dst,const void*src,unsigned n,void*bar){asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"::"r"(s32(dst)),"l"(src),"r"(n),"r"(s32(bar)):"memory");}

// T1: every iteration the warp issues `per` instructions, each = 32 copies of `bytes` bytes
__global__ void t1(const double*g, int iters, int per, int bytes, long long*out, int stride_d){
  extern __shared__ __align__(128) unsigned char sm[];
  unsigned long long*bar=(unsigned long long*)sm; unsigned char*buf=sm+128;
  int lane=threadIdx.x;
  if(lane==0){mbar_init(bar,1);} __syncwarp();
  asm volatile("fence.mbarrier_init.release.cluster;":::"memory");
  long long t0=clock64();
  for(int it=0;it<iters;++it){
    if(lane==0) mbar_expect(bar,(unsigned)per*32u*bytes);
    __syncwarp();
    for(int k=0;k<per;++k){
      const double*src=g+((size_t)((it*per+k)&1023)*32+lane)*stride_d;
      bulk(buf+(size_t)(k*32+lane)*bytes,src,bytes,bar);
    }
    while(!mbar_try(bar,it&1)){}
  }
  long long t1=clock64();
  if(lane==0) out[blockIdx.x]=t1-t0;
}

// T3: LDS.64 patterns. off[lane] = byte offset; active mask
__global__ void t3(const int*offs, unsigned mask, int iters, long long*out, double*sink){
  __shared__ __align__(16) double buf[6000];
  int lane=threadIdx.x; for(int i=lane;i<6000;i+=32) buf[i]=i; __syncwarp();
  unsigned a=s32(buf)+offs[lane]; bool act=(mask>>lane)&1; unsigned acc[16];
#pragma unroll
  for(int k=0;k<16;++k) acc[k]=k;
  long long t0=clock64();
  for(int it=0;it<iters;++it){
#pragma unroll
    for(int k=0;k<16;++k){ unsigned lo=0,hi=0; if(act) asm volatile("ld.shared.v2.u32 {%0,%1},[%2];":"=r"(lo),"=r"(hi):"r"(a+8*k)); acc[k]^=lo+hi; }
  }
  long long t1=clock64();
  unsigned x=0;
#pragma unroll
  for(int k=0;k<16;++k) x^=acc[k];
  if(lane==0) out[0]=t1-t0; sink[lane]=x;
}
__global__ void t3b(const int*offs, unsigned mask, int iters, long long*out, double*sink){  // LDS.128
  __shared__ __align__(16) double buf[6000];
  int lane=threadIdx.x; for(int i=lane;i<6000;i+=32) buf[i]=i; __syncwarp();
  unsigned a=s32(buf)+offs[lane]; bool act=(mask>>lane)&1; unsigned acc[8];
#pragma unroll
  for(int k=0;k<8;++k) acc[k]=k;
  long long t0=clock64();
  for(int it=0;it<iters;++it){
#pragma unroll
    for(int k=0;k<8;++k){ unsigned x0=0,x1=0,x2=0,x3=0; if(act) asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3},[%4];":"=r"(x0),"=r"(x1),"=r"(x2),"=r"(x3):"r"(a+16*k)); acc[k]^=x0+x1+x2+x3; }
  }
  long long t1=clock64();
  unsigned x=0;
#pragma unroll
  for(int k=0;k<8;++k) x^=acc[k];
  if(lane==0) out[0]=t1-t0; sink[lane]=x;
}
// T4 latencies
__global__ void t4(int iters,long long*out,double*sink){
  __shared__ double sb[64];
  int lane=threadIdx.x; double x=1.0+lane*1e-9, c=1e-9; sb[lane]=x; __syncwarp();
  long long t0=clock64();
  for(int i=0;i<iters;++i){ x=fma(x,c,1.0); x=fma(x,c,1.0); x=fma(x,c,1.0); x=fma(x,c,1.0);}   // 4 dependent DFMA
  long long t1=clock64();
  for(int i=0;i<iters;++i){ // shfl + dfma chain x4
#pragma unroll
    for(int k=0;k<4;++k){ int lo=__shfl_sync(~0u,__double2loint(x),(lane+31)&31), hi=__shfl_sync(~0u,__double2hiint(x),(lane+31)&31); x=fma(__hiloint2double(hi,lo),c,1.0);} }
  long long t2=clock64();
  for(int i=0;i<iters;++i){ // sts -> lds -> dfma x4
#pragma unroll
    for(int k=0;k<4;++k){ asm volatile("st.volatile.shared.f64 [%0],%1;"::"r"(s32(sb+lane)),"d"(x):"memory"); double v; asm volatile("ld.volatile.shared.f64 %0,[%1];":"=d"(v):"r"(s32(sb+((lane+1)&31)))); x=fma(v,c,1.0);} }
  long long t3=clock64();
  // throughput of independent DFMAs: 32 accumulators
  double r[32];
#pragma unroll
  for(int k=0;k<32;++k) r[k]=x+k;
  long long t4=clock64();
  for(int i=0;i<iters;++i){
#pragma unroll
    for(int k=0;k<32;++k) r[k]=fma(r[k],c,x);
  }
  long long t5=clock64();
#pragma unroll
  for(int k=0;k<32;++k) x+=r[k];
  if(lane==0){out[0]=t1-t0;out[1]=t2-t1;out[2]=t3-t2;out[3]=t5-t4;} sink[lane]=x;
}
int main(){
  long long*out; double*sink; CK(cudaMallocManaged(&out,1024*8)); CK(cudaMalloc(&sink,8192));
  double*g; size_t gel=(size_t)1024*32*64; CK(cudaMalloc(&g,gel*8)); CK(cudaMemset(g,0,gel*8));
  CK(cudaFuncSetAttribute(t1,cudaFuncAttributeMaxDynamicSharedMemorySize,200*1024));
  int iters=2000;
  for(int bytes:{256}) for(int per:{1}) for(int ctas:{1}){
    if((size_t)per*32*bytes>190*1024) continue;
    t1<<<ctas,32,128+per*32*bytes>>>(g,iters,per,bytes,out,64); CK(cudaDeviceSynchronize());
    t1<<<ctas,32,128+per*32*bytes>>>(g,iters,per,bytes,out,64); CK(cudaDeviceSynchronize());
    long long mx=0; for(int i=0;i<ctas;++i) if(out[i]>mx) mx=out[i];
    printf("T1 bulk bytes=%4d per=%d ctas=%3d : %.1f cycles per 32-copy instruction (%.2f B/cycle/SM)\n",bytes,per,ctas,(double)mx/iters/per, 32.0*bytes*per*iters/mx);
  }
  // T3 patterns
  int h_offs[32]; int*d_offs; CK(cudaMalloc(&d_offs,128));
  auto run3=[&](const char*name,unsigned mask,bool v128)->int{
    CK(cudaMemcpy(d_offs,h_offs,128,cudaMemcpyHostToDevice));
    if(v128) t3b<<<1,32>>>(d_offs,mask,4000,out,sink); else t3<<<1,32>>>(d_offs,mask,4000,out,sink);
    CK(cudaDeviceSynchronize());
    printf("T3 %-40s mask=%08x %s : %.2f cycles per %s\n",name,mask,v128?"LDS.128":"LDS.64 ",(double)out[0]/4000/(v128?8:16),v128?"LDS.128":"LDS.64");
    return 0; };
  for(int l=0;l<32;++l) h_offs[l]=l*8;            if(run3("contiguous 8B",~0u,false)) return 1;
  for(int l=0;l<32;++l) h_offs[l]=l*(1024+8);     if(run3("stride 1032B (conflict-free)",~0u,false)) return 1;
  for(int l=0;l<32;++l) h_offs[l]=l*1024;         if(run3("stride 1024B (32-way)",~0u,false)) return 1;
  for(int l=0;l<32;++l) h_offs[l]=(l&7)*8+(l>>3)*1024+ ((l>>3)&1)*64;  if(run3("8B banks: lanes l,l+16 collide (cross-half only)",~0u,false)) return 1;
  for(int l=0;l<32;++l) h_offs[l]=(l&7)*8+(l>>3)*1024;  if(run3("lanes l,l+8,l+16,l+24 collide (2-way per half)",~0u,false)) return 1;
  for(int l=0;l<32;++l) h_offs[l]=(l&3)*8+(l>>2)*1024;  if(run3("4-way per half",~0u,false)) return 1;
  // tensor-panel layout: lane j, q=(u-j)&31, p=q>>3,f=q&7 ; addr = region(p)+ e*E + f*cs + (d+f+1)*8 ; only mod 128 matters
  { int hp[4]={36,28,20,12}, dl[4]={0,1,0,1}; int po[4]={0,18432,18432+14336,18432+14336+10240};
    for(int u:{0,5}){ for(int l=0;l<32;++l){int q=(u-l)&31,p=q>>3,f=q&7,e=l&7; h_offs[l]=(po[p]+e*8*hp[p]*8+f*hp[p]*8+(dl[p]+f+1)*8)%40000;}
      if(run3(u?"panel layout u=5":"panel layout u=0",~0u,false)) return 1; if(run3("panel layout half lanes",0x0000ffffu,false)) return 1; if(run3("panel layout 8 lanes",0x000000ffu,false)) return 1;} }
  for(int l=0;l<32;++l) h_offs[l]=l*(1024+8);     if(run3("conflict-free, 16 lanes (one half)",0x0000ffffu,false)) return 1;
  for(int l=0;l<32;++l) h_offs[l]=l*(1024+8);     if(run3("conflict-free, lanes 8..23",0x00ffff00u,false)) return 1;
  for(int l=0;l<32;++l) h_offs[l]=l*(1024+8);     if(run3("conflict-free, 8 lanes",0x000000ffu,false)) return 1;
  for(int l=0;l<32;++l) h_offs[l]=l*(1024+16);    if(run3("stride 1040B LDS.128 (conflict-free)",~0u,true)) return 1;
  for(int l=0;l<32;++l) h_offs[l]=l*(1024+16);    if(run3("stride 1040B LDS.128 8 lanes",0xffu,true)) return 1;
  for(int l=0;l<32;++l) h_offs[l]=l*(1024+16);    if(run3("stride 1040B LDS.128 16 lanes",0xffffu,true)) return 1;
  t4<<<1,32>>>(10000,out,sink); CK(cudaDeviceSynchronize());
  printf("T4 DFMA dependent latency %.2f cyc; SHFL64+DFMA %.2f cyc; STS->LDS->DFMA %.2f cyc; independent DFMA issue %.2f cyc/instr\n",
         (double)out[0]/40000,(double)out[1]/40000,(double)out[2]/40000,(double)out[3]/320000);
  return 0;
}
