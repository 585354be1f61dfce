// Per-SM streaming bandwidth from HBM with few CTAs: LDG vs cp.async vs TMA bulk.  nvcc -arch=sm_100a
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
// (a) plain LDG.128, NT threads, each CTA streams its own contiguous slab
__global__ void k_ldg(const double2 *__restrict__ src, size_t n_per_cta, double *out)
{
    const double2 *p = src + (size_t)blockIdx.x * n_per_cta;
    double acc = 0;
    for (size_t i = threadIdx.x; i < n_per_cta; i += blockDim.x * 4) {
        double2 a = __ldcs(p + i), b = (i + blockDim.x < n_per_cta) ? __ldcs(p + i + blockDim.x) : make_double2(0, 0);
        double2 c = (i + 2 * blockDim.x < n_per_cta) ? __ldcs(p + i + 2 * blockDim.x) : make_double2(0, 0);
        double2 d = (i + 3 * blockDim.x < n_per_cta) ? __ldcs(p + i + 3 * blockDim.x) : make_double2(0, 0);
        acc += a.x + b.y + c.x + d.y;
    }
    if (acc == 1.2345) out[0] = acc;
}
// (b) TMA bulk copies of CH bytes, NSTG stages in flight, one thread issues, everyone waits
template <int CH, int NSTG>
__global__ void k_bulk(const char *__restrict__ src, size_t bytes_per_cta, double *out)
{
    extern __shared__ __align__(128) unsigned char sm[];
    unsigned long long *bar = reinterpret_cast<unsigned long long *>(sm + (size_t)CH * NSTG);
    const char *p = src + (size_t)blockIdx.x * bytes_per_cta;
    const int nchunk = (int)(bytes_per_cta / CH);
    if (threadIdx.x == 0) {
        for (int s = 0; s < NSTG; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[s])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    double acc = 0;
    if (threadIdx.x == 0) {
        for (int c = 0; c < NSTG && c < nchunk; ++c) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar[c])), "r"(CH) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(sm + (size_t)c * CH)), "l"(p + (size_t)c * CH), "r"(CH), "r"(smem_u32(&bar[c])) : "memory");
        }
        for (int c = 0; c < nchunk; ++c) {
            const int s = c % NSTG;
            const unsigned ph = (c / NSTG) & 1;
            unsigned ok = 0;
            while (!ok)
                asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(smem_u32(&bar[s])), "r"(ph) : "memory");
            acc += reinterpret_cast<double *>(sm + (size_t)s * CH)[0];
            const int n = c + NSTG;
            if (n < nchunk) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar[s])), "r"(CH) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(sm + (size_t)s * CH)), "l"(p + (size_t)n * CH), "r"(CH), "r"(smem_u32(&bar[s])) : "memory");
            }
        }
    }
    if (acc == 1.2345) out[0] = acc;
}
template <class F>
float timeit(F f)
{
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) printf("  CUDA error: %s\n", cudaGetErrorString(e));
    return ms;
}
int main()
{
    const size_t per_cta = 64ull << 20;  // 64 MiB per CTA
    const int max_cta = 148;
    char *src; double *out;
    cudaMalloc(&src, per_cta * max_cta); cudaMalloc(&out, 8);
    cudaMemset(src, 0, per_cta * max_cta);
    for (int ncta : {1, 16, 64, 148}) {
        for (int nt : {256, 1024}) {
            float ms = timeit([&] { k_ldg<<<ncta, nt>>>((const double2 *)src, per_cta / 16, out); });
            printf("LDG.128  ctas %3d threads %4d : %7.1f GB/s per SM, %8.1f GB/s total\n", ncta, nt, per_cta / ms / 1e6, per_cta * ncta / ms / 1e6);
        }
        {
            cudaFuncSetAttribute(k_bulk<8192, 24>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192 * 24 + 256);
            float ms = timeit([&] { k_bulk<8192, 24><<<ncta, 32, 8192 * 24 + 256>>>(src, per_cta, out); });
            printf("bulk 8K x24  ctas %3d : %7.1f GB/s per SM, %8.1f GB/s total\n", ncta, per_cta / ms / 1e6, per_cta * ncta / ms / 1e6);
            cudaFuncSetAttribute(k_bulk<32768, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768 * 6 + 256);
            ms = timeit([&] { k_bulk<32768, 6><<<ncta, 32, 32768 * 6 + 256>>>(src, per_cta, out); });
            printf("bulk 32K x6  ctas %3d : %7.1f GB/s per SM, %8.1f GB/s total\n", ncta, per_cta / ms / 1e6, per_cta * ncta / ms / 1e6);
            cudaFuncSetAttribute(k_bulk<2048, 96>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2048 * 96 + 1024);
            ms = timeit([&] { k_bulk<2048, 96><<<ncta, 32, 2048 * 96 + 1024>>>(src, per_cta, out); });
            printf("bulk 2K x96  ctas %3d : %7.1f GB/s per SM, %8.1f GB/s total\n", ncta, per_cta / ms / 1e6, per_cta * ncta / ms / 1e6);
        }
    }
    return 0;
}
