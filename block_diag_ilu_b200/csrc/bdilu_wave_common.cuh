// Shared pieces of the wavefront solve kernels: mbarrier / bulk-copy / volatile shared-memory helpers, the
// stage layout, and the single-lane TMA producer of the generic kernel.
#pragma once
#include "bdilu_common.cuh"
#include "bdilu_kernels.h"

namespace bdilu {

namespace wave {

constexpr int NW = 16;       // consumer warps
constexpr int MAXD = 4;      // sub/super-diagonals handled by the fast path
constexpr int MAX_STAGES = 32;
constexpr unsigned SENT_HI = 0xFFF4DEADu;  // high word of the "not yet written" pattern (an sNaN)
constexpr unsigned long long SENT = 0xFFF4DEAD00000000ull;

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(void *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(void *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(void *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try(void *bar, unsigned parity)
{
    unsigned ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// No wait in these kernels can hang: every one is on an event that the protocol orders strictly earlier, so a wait
// that is retried SPIN_LIMIT times (seconds; legitimate waits last microseconds) is a protocol failure or corrupt
// input, and the kernel traps -> cudaErrorLaunchFailure at the caller's next synchronisation, which the C ABI
// reports as BDILU_ERR_CUDA.  The counters live on the retry paths only (a per-poll clock read or a guard object per
// wait site cost the blockw-32 kernel 16 %: measured).
constexpr unsigned SPIN_LIMIT = 1u << 26;
__device__ __forceinline__ void mbar_wait(void *bar, unsigned parity)
{
    unsigned n = 0;
    while (!mbar_try(bar, parity)) {
        if (++n > SPIN_LIMIT) __trap();
    }
}
// for the consumer warps of the blockw-32 kernel, whose instruction stream does not tolerate even a retry counter
// (measured: +6..10 % each); their waits are bounded through the producer warps' watchdog (bdilu_solve_quad32.cu)
__device__ __forceinline__ void mbar_wait_unbounded(void *bar, unsigned parity)
{
    while (!mbar_try(bar, parity)) {
    }
}
// clock-bounded wait for warps off the critical path (producers): traps after WAIT_CYCLES of waiting
constexpr long long WAIT_CYCLES = 1ll << 33;  // ~4 s
__device__ __forceinline__ void mbar_wait_timed(void *bar, unsigned parity)
{
    if (mbar_try(bar, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try(bar, parity)) {
        if (clock64() - t0 > WAIT_CYCLES) __trap();
    }
}
// global -> shared bulk copy, completion counted in bytes on `bar`
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, void *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ double ld_volatile_s(const double *p)
{
    double v;
    asm volatile("ld.volatile.shared.f64 %0, [%1];" : "=d"(v) : "r"(smem_u32(p)));
    return v;
}
__device__ __forceinline__ int ld_volatile_s(const int *p)
{
    int v;
    asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)));
    return v;
}
__device__ __forceinline__ bool is_sent(double v) { return (unsigned)__double2hiint(v) == SENT_HI; }

// byte layout of one stage (all pieces 16-byte aligned)
struct StageLayout {
    int o_rbc, o_cbr, o_vec, o_diag, bytes;  // block at offset 0
    int blk_bytes, idx_bytes, vec_bytes, diag_bytes;
};
__host__ __device__ inline StageLayout stage_layout(int w, int ld, int ndiag)
{
    StageLayout s;
    s.blk_bytes = w * ld * 8;
    s.idx_bytes = w * 4;
    s.vec_bytes = w * 8;
    s.diag_bytes = w * 8;
    auto up = [](int x) { return (x + 15) & ~15; };
    s.o_rbc = up(s.blk_bytes);
    s.o_cbr = s.o_rbc + up(s.idx_bytes);
    s.o_vec = s.o_cbr + up(s.idx_bytes);
    s.o_diag = s.o_vec + up(s.vec_bytes);
    s.bytes = s.o_diag + ndiag * up(s.diag_bytes);
    return s;
}


__device__ __forceinline__ void st_volatile_s(double *p, double v)
{
    asm volatile("st.volatile.shared.f64 [%0], %1;" ::"r"(smem_u32(p)), "d"(v) : "memory");
}
__device__ __forceinline__ void st_volatile_s(int *p, int v)
{
    asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}

struct Smem {
    unsigned char *stages;
    unsigned long long *full, *empty;
    double *ring;  // [2*NW][32]
    int *done;     // [NW]
    int *flagw;    // [2]: nan_first, zero_last
};

struct Args {
    const double *sd;             // this system's packed data
    const int *rbc, *cbr;         // this system's permutation vectors
    const double *b;              // this vector
    double *x;
    Dims D;
    int NS;
    StageLayout SL;
    int dbg;
};

// ---- producer: one lane issues the bulk copies of one sweep --------------------------------------
template <bool FWD>
__device__ void produce(const Args &A, const Smem &S, int it0)
{
    const int N = A.D.nblocks, w = A.D.w;
    for (int i = 0; i < N; ++i) {
        const int b = FWD ? i : N - 1 - i;
        const int it = it0 + i, s = it % A.NS, use = it / A.NS;
        if (use >= 1) mbar_wait(&S.empty[s], (use - 1) & 1);
        unsigned char *st = S.stages + (size_t)s * A.SL.bytes;
        const int nd = min(A.D.ndiag, i);
        unsigned tx = A.SL.blk_bytes + A.SL.idx_bytes + nd * A.SL.diag_bytes;
        if (FWD) tx += A.SL.idx_bytes + A.SL.vec_bytes;
        mbar_expect_tx(&S.full[s], tx);
        bulk_g2s(st, A.sd + A.D.block(b), A.SL.blk_bytes, &S.full[s]);
        bulk_g2s(st + A.SL.o_cbr, A.cbr + (int64_t)b * w, A.SL.idx_bytes, &S.full[s]);
        if (FWD) {
            bulk_g2s(st + A.SL.o_rbc, A.rbc + (int64_t)b * w, A.SL.idx_bytes, &S.full[s]);
            bulk_g2s(st + A.SL.o_vec, A.b + (int64_t)b * w, A.SL.vec_bytes, &S.full[s]);
        }
        for (int d = 0; d < nd; ++d) {
            // forward: sub(d, b-1-d, :) couples block b to y_{b-1-d}; backward: sup(d, b, :) to x_{b+1+d}
            const double *src = FWD ? A.sd + A.D.off_sub + A.D.diag_chunk(d, b - 1 - d)
                                    : A.sd + A.D.off_sup + A.D.diag_chunk(d, b);
            bulk_g2s(st + A.SL.o_diag + d * ((A.SL.diag_bytes + 15) & ~15), src, A.SL.diag_bytes, &S.full[s]);
        }
    }
}

}  // namespace wave

}  // namespace bdilu
