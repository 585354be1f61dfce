// Microbenchmark: per-SM cost (cycles per warp-instruction) of the shared-memory / shuffle / redux instructions the
// factor kernel leans on, as a function of the number of active lanes.  One CTA of NW warps on every SM.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

template <int MODE>
__global__ void __launch_bounds__(512, 1) k(long long *out, int iters, int nact)
{
    extern __shared__ __align__(16) unsigned char sm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned base = smem_u32(sm) + warp * 1024;
    double a = lane * 1.5, b = warp * 0.25;
    unsigned u = lane;
    const bool act = lane < nact;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            if (MODE == 0) {  // STS.128, active lanes store 16 B each at lane*16
                asm volatile("{ .reg .pred p; setp.ne.u32 p, %3, 0; @p st.shared.v2.f64 [%0], {%1, %2}; }" ::"r"(base + lane * 16), "d"(a), "d"(b), "r"((unsigned)act) : "memory");
            } else if (MODE == 1) {  // STS.64
                asm volatile("{ .reg .pred p; setp.ne.u32 p, %2, 0; @p st.shared.f64 [%0], %1; }" ::"r"(base + lane * 8), "d"(a), "r"((unsigned)act) : "memory");
            } else if (MODE == 2) {  // STS.32
                asm volatile("{ .reg .pred p; setp.ne.u32 p, %2, 0; @p st.shared.u32 [%0], %1; }" ::"r"(base + lane * 4), "r"(u), "r"((unsigned)act) : "memory");
            } else if (MODE == 3) {  // LDS.128 broadcast (all lanes, same address)
                double x, y;
                asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x), "=d"(y) : "r"(base + j * 16) : "memory");
                a += x; b += y;
            } else if (MODE == 4) {  // LDS.64 broadcast
                double x;
                asm volatile("ld.shared.f64 %0, [%1];" : "=d"(x) : "r"(base + j * 8) : "memory");
                a += x;
            } else if (MODE == 5) {  // LDS.128 distinct addresses
                double x, y;
                asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x), "=d"(y) : "r"(base + lane * 16 + (j & 1) * 512) : "memory");
                a += x; b += y;
            } else if (MODE == 6) {  // SHFL.IDX 32-bit
                u = __shfl_sync(0xffffffffu, u, (j + nact) & 31) + 1;
            } else if (MODE == 7) {  // REDUX max
                u = __reduce_max_sync(0xffffffffu, u + j) + lane;
            } else if (MODE == 8) {  // LDS.128, 2 distinct addresses by lane parity (16x2 layout read)
                double x, y;
                asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x), "=d"(y) : "r"(base + (lane & 1) * 16 + j * 32) : "memory");
                a += x; b += y;
            } else if (MODE == 9) {  // branch-guarded STS.128 (only active lanes reach the instruction)
                if (act) asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(base + lane * 16), "d"(a), "d"(b) : "memory");
            } else if (MODE == 10) {  // DFMA dependent chains x2 as a control
                a = fma(a, 1.0000001, b); b = fma(b, 0.9999999, a);
            }
        }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
    if (a + b + u == 12345.678) out[0] = 0;
}

template <int MODE>
double run(int nw, int nact, int iters = 2000)
{
    int nsm;
    CK(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0));
    long long *d;
    CK(cudaMalloc(&d, nsm * sizeof(long long)));
    CK(cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768));
    k<MODE><<<nsm, nw * 32, 32768>>>(d, 10, nact);
    k<MODE><<<nsm, nw * 32, 32768>>>(d, iters, nact);
    CK(cudaDeviceSynchronize());
    long long h[256];
    CK(cudaMemcpy(h, d, nsm * sizeof(long long), cudaMemcpyDeviceToHost));
    double s = 0;
    for (int i = 0; i < nsm; ++i) s += (double)h[i];
    CK(cudaFree(d));
    return s / nsm / ((double)iters * 16 * nw);  // SM cycles per warp-instruction
}

int main()
{
    const char *names[] = {"STS.128 pred", "STS.64 pred", "STS.32 pred", "LDS.128 bcast", "LDS.64 bcast", "LDS.128 distinct",
                           "SHFL.IDX", "REDUX.MAX", "LDS.128 2addr", "STS.128 branch", "DFMA x2"};
    for (int nw : {4, 16}) {
        printf("--- %d warps per SM: SM cycles per warp-instruction ---\n", nw);
        for (int nact : {1, 2, 4, 8, 16, 32}) {
            printf("active %2d: %s %.2f | %s %.2f | %s %.2f | %s %.2f\n", nact, names[0], run<0>(nw, nact), names[1],
                   run<1>(nw, nact), names[2], run<2>(nw, nact), names[9], run<9>(nw, nact));
        }
        printf("%s %.2f | %s %.2f | %s %.2f | %s %.2f\n", names[3], run<3>(nw, 32), names[4], run<4>(nw, 32), names[5],
               run<5>(nw, 32), names[8], run<8>(nw, 32));
        printf("%s %.2f | %s %.2f | %s (2 instr) %.2f\n", names[6], run<6>(nw, 32), names[7], run<7>(nw, 32), names[10],
               run<10>(nw, 32));
    }
    return 0;
}
