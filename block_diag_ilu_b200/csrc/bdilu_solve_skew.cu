// Kernel 2, blockw == 32, fourth generation ("skew32"): ILU_inplace<double>::solve (block_diag_ilu.hpp:486-538)
// as a SYSTOLIC sweep: one warp per (system, right-hand side), lane = block.
//
// The chain over the blocks is sequential (:502, :518), but row li of block B needs from its predecessors only
// y_{B-1-d}[colbyrow[li]] (:508-511, :524-527).  When the permutations are the identity (no row interchange was
// needed -- the diagonally dominant matrices this library is made for, README.rst:9-15) that is row li itself, so
// block B can run exactly ONE ROW behind block B-1.  A warp therefore holds 32 consecutive blocks, lane j on the
// blocks j, j+32, ..., lane j one row behind lane j-1; every lane keeps the 32 partial sums of its block in
// registers, and one step of the warp finishes one row of each of the 32 blocks in flight:
//     y   = (b[q] + R[q]) - sub1[q]*y'' - sub0[q]*y'        y', y'': the values lanes j-1 / j-2 finished one and two
//     R[r] -= L[r][q] * y   for r > q                        steps ago -- two shuffles, no shared memory, no polling
// The register that holds "the current row" must have the same NAME in all lanes although the lanes sit on different
// rows: row r of lane j lives in R[(r + j) mod 32], so at step u the current row is R[u mod 32] everywhere and the
// 32 steps of the unrolled loop body are straight-line code with static register indices.
// Per block and sweep this is ~90 warp instructions (31 DFMA + 31 LDS + a dozen others) instead of ~650 in the
// wavefront-across-warps kernel (bdilu_solve_quad32.cu), and the dependent path of a step is SHFL + 2 DFMA.
//
// Staging.  A lane reads column q of ITS block, so the blocks in flight have to sit in shared memory; 32 whole blocks
// do not fit, and the lanes need only the triangle.  Time is cut into group steps of 8 rows.  Four 8-column PANELS
// per block (rows 8p.. of columns 8p..8p+7 in the forward sweep, rows ..31-8p of columns 24-8p..31-8p backward) are
// needed for 8 steps each; panel p of the 8 blocks of group G is first needed in group step G + p, and all panels
// first needed in group step t share TIME SLOT t: 4 TMA tensor copies (one per panel height, 8 blocks each) plus the
// matching 8-row pieces of b / y and of the sub / sup rows.  A slot is in use for two group steps; 4 slots = 2 in
// use + 2 in flight (92 KB per SM).  With 4 slots the ring period equals the 32 steps of the loop body, so every
// shared-memory address a lane touches is a function of (step mod 32, lane) alone: a 1024-entry table, one LDS.64 per
// step, no address arithmetic.  Out-of-range blocks (before the first, after the last, partial groups) are TMA
// out-of-bounds boxes = zeros, which makes them harmless without a single special case in the loop; the 4 zero rows
// the boxes carry beyond the block (also out-of-bounds, no HBM traffic) let the updates be predicated per group of 4.
// HBM traffic: 62.5 % of every block per sweep instead of 100 %.
//
// Scope: identity permutations, finite b, U diagonals inside the exact range of the fast reciprocal.  Everything else
// (a system that pivoted, NaN in b, a zero diagonal) is handed back through `status` and solved by the general
// wavefront kernel in a second launch, which also produces the reference's flag values for those cases.
#include <cuda.h>

#include "bdilu_wave_common.cuh"

namespace bdilu {

namespace skew {

using namespace wave;  // mbarrier helpers

constexpr int GB = 8;      // blocks per group
constexpr int NSLOT = 4;   // time slots
constexpr int PADR = 4;    // zero rows a staged column carries beyond the block
constexpr int MAXND = 2;

__device__ unsigned long long g_stats[3];  // vectors solved here / handed back at the end / skipped at the start

__host__ __device__ constexpr int pan_hp(int p) { return 32 - 8 * p + PADR; }  // staged rows of panel p
__host__ __device__ constexpr int pan_cs(int p) { return pan_hp(p) * 8; }      // bytes per staged column
__host__ __device__ constexpr int pan_eb(int p) { return 8 * pan_cs(p); }      // bytes per block (8 columns)
__host__ __device__ constexpr int pan_off(int p) { return 64 * (288 * p - 32 * p * (p - 1)); }  // sum of 8*pan_eb
constexpr int PAN_BYTES = pan_off(4);    // 49152
constexpr int VEC_BYTES = 4 * GB * 64;   // 4 panels x 8 blocks x 8 rows of doubles
__host__ __device__ constexpr int slot_bytes(int nd) { return PAN_BYTES + VEC_BYTES * (1 + nd); }
constexpr int TAB_BYTES = 32 * 32 * 8;

struct Maps {
    CUtensorMap pan[4];      // {row, col, block, system}, box {hp_p, 8, 8, 1}
    CUtensorMap b, x;        // {row in panel, panel, block, rhs, system}, box {8, 1, 8, 1, 1}
    CUtensorMap sub[MAXND];  // {row in panel, panel, chunk row, system}, box {8, 1, 8, 1}
    CUtensorMap sup[MAXND];
};

__device__ __forceinline__ void tma4(unsigned dst, const CUtensorMap *m, int c0, int c1, int c2, int c3, unsigned bar)
{
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];" ::"r"(dst),
        "l"(m), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(bar)
        : "memory");
}
__device__ __forceinline__ void tma5(unsigned dst, const CUtensorMap *m, int c0, int c1, int c2, int c3, int c4, unsigned bar)
{
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];" ::"r"(dst),
        "l"(m), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4), "r"(bar)
        : "memory");
}

__device__ __forceinline__ double lds64(unsigned a)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
    return v;
}
template <int OFF>
__device__ __forceinline__ double lds64o(unsigned a)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1+%2];" : "=d"(v) : "r"(a), "n"(OFF));
    return v;
}
__device__ __forceinline__ void sts64(unsigned a, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v) : "memory"); }
__device__ __forceinline__ uint2 lds_u2(unsigned a)
{
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
    return v;
}

// A wait that cannot hang: a protocol failure (or a lost copy) ends in a trap = cudaErrorLaunchFailure at the next
// synchronisation, which the C ABI reports as BDILU_ERR_CUDA.
__device__ __forceinline__ void mbar_wait_bounded(unsigned bar, unsigned parity)
{
    for (int spin = 0;; ++spin) {
        unsigned ok;
        asm volatile(
            "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
        if (ok) return;
        if (spin > (1 << 24)) __trap();
    }
}
__device__ __forceinline__ unsigned mbar_test(unsigned bar, unsigned parity)
{
    unsigned ok;
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(ok)
                 : "r"(bar), "r"(parity)
                 : "memory");
    return ok;
}
__device__ __forceinline__ void mbar_arrive_a(unsigned bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void mbar_expect_a(unsigned bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}

// lanes whose row position q = (u - lane) mod 32 is <= qmax
__host__ __device__ constexpr unsigned lanes_q_le(int u, int qmax)
{
    unsigned m = 0;
    for (int j = 0; j < 32; ++j)
        if (((u - j) & 31) <= qmax) m |= 1u << j;
    return m;
}

// ---- consumer ----------------------------------------------------------------------------------------------------
struct Cons {
    double R[32];     // partial sums: row r of my block in R[(r + lane) & 31]
    double v1, v2;    // the values I finished one and two steps ago
    double s1, s2;    // what lanes j-1 / j-2 finished one / two steps ago (shuffled at the end of the previous step)
    double rinv;      // backward: 1/U[li][li] of the current step
    unsigned col, vec;  // shared-memory addresses of this step (from the table)
    unsigned zcol;    // a column of zeros (forward: its first entry, backward: its last)
    unsigned tab;     // my table column
    unsigned lanebit;
    int src1, src2;
    int blkbase;      // backward: lane + 32*M - 32 (block index in sweep order if I have not wrapped yet in this round)
    int N;
    unsigned punt;
    unsigned okfull;
    bool last;        // last round of the sweep
    unsigned bar_full, bar_empty;  // addresses of full[0], empty[0]
};

__device__ __forceinline__ double shfl_d(double v, int src)
{
    const int lo = __shfl_sync(FULL, __double2loint(v), src);
    const int hi = __shfl_sync(FULL, __double2hiint(v), src);
    return __hiloint2double(hi, lo);
}

// coefficients of the rows 4g .. 4g+3 after (forward) / before (backward) the current one.  Lanes whose block has no
// such rows left read a zero area instead (one select per group, no predicated arithmetic): their updates add 0.
template <bool FWD, int U, int G>
__device__ __forceinline__ void group_load(const Cons &C, double (&a)[32])
{
    constexpr unsigned MASK = lanes_q_le(U, 30 - 4 * G);
    constexpr int S0 = 4 * G;
    constexpr int STEP = FWD ? 8 : -8;
    const unsigned base = (C.lanebit & MASK) ? C.col : C.zcol;
    a[S0] = lds64o<STEP * S0>(base);
    a[S0 + 1] = lds64o<STEP * (S0 + 1)>(base);
    a[S0 + 2] = lds64o<STEP * (S0 + 2)>(base);
    if constexpr (G < 7) a[S0 + 3] = lds64o<STEP * (S0 + 3)>(base);
}

template <bool FWD, int ND, int U>
__device__ __forceinline__ void step(Cons &C, int M)
{
    // ---- group-step boundary (every 8 steps): the slot last used 1..8 steps ago... no: the slot of the group step
    // before the previous one is free; the slot of the NEXT group step must have landed one step early, because the
    // backward sweep reads the next step's diagonal at the end of this step
    if constexpr ((U & 7) == 0) {
        __syncwarp();
        if (C.lanebit == 1u) mbar_arrive_a(C.bar_empty + 8 * (((U >> 3) + 2) & 3));
    }
    if constexpr ((U & 7) == 3 || (U & 7) == 7) {
        constexpr int SLN = ((U >> 3) + 1) & 3;
        const unsigned par = ((U >> 3) == 3) ? ((unsigned)(M + 1) & 1u) : ((unsigned)M & 1u);
        const bool more = ((U >> 3) != 3) || !C.last;  // nothing is staged beyond the last round
        if (more) {
            if constexpr ((U & 7) == 3) C.okfull = mbar_test(C.bar_full + 8 * SLN, par);  // early, non-blocking
            else if (!C.okfull) mbar_wait_bounded(C.bar_full + 8 * SLN, par);
        }
    }

    // ---- addresses of the next step
    const uint2 nxt = lds_u2(C.tab + ((U + 1) & 31) * 256);

    // ---- coefficients of my column (they do not depend on this step's value: issued ahead of the chain)
    double a[32];
    group_load<FWD, U, 0>(C, a);
    group_load<FWD, U, 1>(C, a);
    group_load<FWD, U, 2>(C, a);
    group_load<FWD, U, 3>(C, a);
    group_load<FWD, U, 4>(C, a);
    group_load<FWD, U, 5>(C, a);
    group_load<FWD, U, 6>(C, a);
    group_load<FWD, U, 7>(C, a);

    // ---- this step's row: right-hand side and coupling coefficients
    const double bv = lds64(C.vec);
    double cf[ND];
#pragma unroll
    for (int d = 0; d < ND; ++d) cf[d] = lds64(C.vec + VEC_BYTES * (1 + d));

    double t = bv + C.R[U];
    if constexpr (ND == 2) t = fma(-cf[1], C.s2, t);
    t = fma(-cf[0], C.s1, t);
    double v;
    if constexpr (FWD) {
        v = t;
        C.punt |= (bv != bv) ? 1u : 0u;  // NaN in b: the general kernel produces the reference's flag
    } else {
        v = t * C.rinv;
    }
    C.R[U] = 0.0;
    sts64(C.vec, v);
    // hand my value to the lanes behind me (used in their next step)
    C.s1 = shfl_d(v, C.src1);
    if constexpr (ND == 2) C.s2 = shfl_d(C.v1, C.src2);
    C.v1 = v;
    const double nv = -v;

    // ---- rank-1 update of the rows still open in my block
#pragma unroll
    for (int sidx = 0; sidx < 31; ++sidx) C.R[(U + 1 + sidx) & 31] = fma(a[sidx], nv, C.R[(U + 1 + sidx) & 31]);

    if constexpr (!FWD) {
        // reciprocal of the next step's diagonal, one step ahead of its use.  Lanes on blocks outside the system get
        // 0 (their values stay 0 and never reach a real block with a nonzero coefficient anyway).
        constexpr unsigned WRAPPED = lanes_q_le((U + 1) & 31, (U + 1) & 31);  // lanes j <= U+1: already in this round's block
        const double dg = lds64(nxt.x + 8);
        const int blk = C.blkbase + ((C.lanebit & WRAPPED) ? 32 : 0) + ((U == 31) ? 32 : 0);
        const bool valid = (unsigned)blk < (unsigned)C.N;
        const bool okr = rcp_fast_ok(dg);
        C.punt |= (valid && !okr) ? 1u : 0u;  // zero / denormal / huge / NaN diagonal: general kernel
        const double r = rcp_fast(dg);
        C.rinv = valid ? r : 0.0;
    }
    C.col = nxt.x;
    C.vec = nxt.y;
}

template <bool FWD, int ND, int U>
struct Steps {
    static __device__ __forceinline__ void run(Cons &C, int M)
    {
        step<FWD, ND, U>(C, M);
        Steps<FWD, ND, U + 1>::run(C, M);
    }
};
template <bool FWD, int ND>
struct Steps<FWD, ND, 32> {
    static __device__ __forceinline__ void run(Cons &, int) {}
};

template <bool FWD, int ND>
__device__ __noinline__ unsigned consume(unsigned zero_area, unsigned tabbase, unsigned bar_full, unsigned bar_empty, int N, int Mtot,
                                         int lane)
{
    Cons C;
#pragma unroll
    for (int r = 0; r < 32; ++r) C.R[r] = 0.0;
    C.v1 = C.v2 = C.s1 = C.s2 = 0.0;
    C.rinv = 0.0;
    C.tab = tabbase + lane * 8;
    C.lanebit = 1u << lane;
    C.src1 = (lane + 31) & 31;
    C.src2 = (lane + 30) & 31;
    C.N = N;
    C.punt = 0;
    C.okfull = 0;
    C.bar_full = bar_full;
    C.bar_empty = bar_empty;
    const uint2 e0 = lds_u2(C.tab);
    C.col = e0.x;
    C.vec = e0.y;
    C.zcol = FWD ? zero_area : zero_area + 31 * 8;
    mbar_wait_bounded(bar_full, 0);  // slot of group step 0
    if constexpr (!FWD) {
        // step 0 of round 0: only lane 0 sits on a real block
        const double dg = lds64(C.col + 8);
        const bool valid = (lane == 0);
        C.punt |= (valid && !rcp_fast_ok(dg)) ? 1u : 0u;
        const double r = rcp_fast(dg);
        C.rinv = valid ? r : 0.0;
    }
    for (int M = 0; M < Mtot; ++M) {
        C.blkbase = lane + 32 * M - 32;
        C.last = (M == Mtot - 1);
        Steps<FWD, ND, 0>::run(C, M);
    }
    // the two slots still held
    __syncwarp();
    if (lane == 0) {
        mbar_arrive_a(bar_empty + 8 * 2);
        mbar_arrive_a(bar_empty + 8 * 3);
    }
    return C.punt;
}

// ---- producer ----------------------------------------------------------------------------------------------------
struct Prod {
    const Maps *maps;
    unsigned sbase, bar_full, bar_empty;
    int N, sys, rhs, slot_b;
    double *x;  // this vector's solution
};

template <bool FWD, int ND>
__device__ __forceinline__ void fill_slot(const Prod &P, int tau)
{
    const int sl = tau & 3;
    const unsigned base = P.sbase + sl * P.slot_b;
    const unsigned bar = P.bar_full + 8 * sl;
    mbar_expect_a(bar, (unsigned)P.slot_b);
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int G = tau - p;
        const int cb = FWD ? 8 * G : P.N - 8 - 8 * G;  // first block of the box (may lie outside: zeros)
        const int prow = FWD ? p : 3 - p;              // which 8 rows of the vectors
        tma4(base + pan_off(p), &P.maps->pan[p], FWD ? 8 * p : -PADR, FWD ? 8 * p : 24 - 8 * p, cb, P.sys, bar);
        tma5(base + PAN_BYTES + p * 512, FWD ? &P.maps->b : &P.maps->x, 0, prow, cb, P.rhs, P.sys, bar);
#pragma unroll
        for (int d = 0; d < ND; ++d)
            tma4(base + PAN_BYTES + VEC_BYTES * (1 + d) + p * 512, FWD ? &P.maps->sub[d] : &P.maps->sup[d], 0, prow,
                 FWD ? cb - 1 - d : cb, P.sys, bar);
    }
}

// results of time slot tau (8 rows of 8 blocks per panel) -> x, by the whole warp
template <bool FWD>
__device__ __forceinline__ void drain_slot(const Prod &P, int tau, int lane)
{
    const unsigned base = P.sbase + (tau & 3) * P.slot_b + PAN_BYTES;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        const int idx = it * 32 + lane;
        const int p = idx >> 6, e = (idx >> 3) & 7, f = idx & 7;
        const int G = tau - p;
        const int blk = (FWD ? 8 * G : P.N - 8 - 8 * G) + e;
        const int row = (FWD ? 8 * p : 24 - 8 * p) + f;
        const double v = lds64(base + idx * 8);
        if ((unsigned)blk < (unsigned)P.N) P.x[(int64_t)blk * 32 + row] = v;
    }
}

template <bool FWD, int ND>
__device__ __noinline__ void produce(const Prod &P, int Mtot, int lane)
{
    const int kfill = 4 * Mtot;
    if (lane == 0) {
        fill_slot<FWD, ND>(P, 0);
        fill_slot<FWD, ND>(P, 1);
    }
    for (int k = 0; k <= kfill + 1; ++k) {
        mbar_wait_bounded(P.bar_empty + 8 * ((k + 2) & 3), (unsigned)(k >> 2) & 1u);
        if (k >= 2) drain_slot<FWD>(P, k - 2, lane);
        if (k + 2 < kfill) {
            // my generic-proxy reads of the slot are ordered before the async-proxy writes that refill it
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) fill_slot<FWD, ND>(P, k + 2);
        }
    }
}

__device__ __forceinline__ void build_table(unsigned char *tabp, unsigned sbase, int slot_b, bool fwd, int tid, int nthreads)
{
    uint2 *tab = reinterpret_cast<uint2 *>(tabp);
    for (int idx = tid; idx < 1024; idx += nthreads) {
        const int u = idx >> 5, j = idx & 31;
        const int q = (u - j) & 31, p = q >> 3, f = q & 7, e = j & 7, a = j >> 3;
        const unsigned slot = sbase + ((a + p) & 3) * slot_b;
        unsigned col, vec;
        if (fwd) {
            col = slot + pan_off(p) + e * pan_eb(p) + f * pan_cs(p) + (f + 1) * 8;  // row q+1 of column q
            vec = slot + PAN_BYTES + p * 512 + e * 64 + f * 8;
        } else {
            const int li = 31 - q;
            col = slot + pan_off(p) + (7 - e) * pan_eb(p) + (7 - f) * pan_cs(p) + (PADR + li - 1) * 8;  // row li-1 of column li
            vec = slot + PAN_BYTES + p * 512 + (7 - e) * 64 + (7 - f) * 8;
        }
        tab[idx] = make_uint2(col, vec);
    }
}

template <int ND>
__global__ void __launch_bounds__(64, 1)
solve_skew32_kernel(const __grid_constant__ Maps maps, int N, int nrhs, double *xvec, int64_t rhs_stride, int64_t sys_stride_v,
                    int *__restrict__ flags, int *__restrict__ status)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned char *smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);  // TMA tensor copies: 128-byte aligned
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int vecid = blockIdx.x;
    if (status[vecid] != 0) {  // permutations not the identity: left to the general kernel
        if (tid == 0) atomicAdd(&g_stats[2], 1ull);
        return;
    }
    const int sys = vecid / nrhs, rhs = vecid % nrhs;
    constexpr int SLOT = slot_bytes(ND);
    unsigned char *tabp = smem + NSLOT * SLOT;
    unsigned long long *bars = reinterpret_cast<unsigned long long *>(tabp + TAB_BYTES);  // full[4], empty[4]
    double *zeros = reinterpret_cast<double *>(bars + 8);  // 32 doubles
    const unsigned zero_area = smem_u32(zeros);
    const unsigned sbase = smem_u32(smem);
    const unsigned tabbase = smem_u32(tabp);
    const unsigned bar_full = smem_u32(bars), bar_empty = smem_u32(bars + 4);
    const int Mtot = (N + 31 + 31) / 32;

    Prod P;
    P.maps = &maps;
    P.sbase = sbase; P.bar_full = bar_full; P.bar_empty = bar_empty;
    P.N = N; P.sys = sys; P.rhs = rhs; P.slot_b = SLOT;
    P.x = xvec + (int64_t)sys * sys_stride_v + (int64_t)rhs * rhs_stride;

    unsigned punt = 0;
#pragma unroll 1
    for (int sweep = 0; sweep < 2; ++sweep) {
        // slot 3 stands for "the group before the first one" during the first group step: zeros
        for (int t = tid; t < SLOT / 16; t += 64) reinterpret_cast<uint4 *>(smem + 3 * SLOT)[t] = make_uint4(0, 0, 0, 0);
        build_table(tabp, sbase, SLOT, sweep == 0, tid, 64);
        if (tid < 32) zeros[tid] = 0.0;
        if (tid == 0) {
            for (int s = 0; s < 8; ++s) {
                if (sweep) asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(bar_full + 8 * s) : "memory");
                mbar_init(bars + s, 1);
            }
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        // generic-proxy writes (zeros; in the second sweep also the y values stored to global memory) before the
        // async-proxy copies that follow
        __threadfence();
        asm volatile("fence.proxy.async;" ::: "memory");
        __syncthreads();
        if (warp == 1) {
            if (sweep == 0) produce<true, ND>(P, Mtot, lane);
            else produce<false, ND>(P, Mtot, lane);
        } else {
            punt |= (sweep == 0) ? consume<true, ND>(zero_area, tabbase, bar_full, bar_empty, N, Mtot, lane)
                                 : consume<false, ND>(zero_area, tabbase, bar_full, bar_empty, N, Mtot, lane);
        }
        __syncthreads();
    }
    if (warp == 0) {
        const bool any = __any_sync(FULL, punt != 0);
        if (lane == 0) {
            if (any) status[vecid] = 2;       // redo on the general kernel (it also writes the flag)
            else if (flags) flags[vecid] = 0;
            atomicAdd(&g_stats[any ? 1 : 0], 1ull);
        }
    }
}

// status[v] = 1 for every vector of a system whose colbyrow is not the identity (rowbycol is its inverse)
__global__ void perm_scan_kernel(const int *__restrict__ colbyrow, int64_t n_per_sys, int nrhs, int *__restrict__ status)
{
    const int sys = blockIdx.y;
    const int4 *p = reinterpret_cast<const int4 *>(colbyrow + (int64_t)sys * n_per_sys);
    const int64_t n4 = n_per_sys >> 2;
    bool bad = false;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const int4 c = __ldg(p + i);
        const int e0 = (int)((i * 4) & 31);
        bad |= (c.x != e0) | (c.y != e0 + 1) | (c.z != e0 + 2) | (c.w != e0 + 3);
    }
    if (__any_sync(FULL, bad) && (threadIdx.x & 31) == 0)
        for (int r = 0; r < nrhs; ++r) status[(int64_t)sys * nrhs + r] = 1;
}

// ---- host: tensor maps -------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn()
{
    static EncodeTiledFn fn = []() -> EncodeTiledFn {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            return nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

static bool encode(CUtensorMap *m, const void *base, int rank, const cuuint64_t *dims, const cuuint64_t *strides_bytes,
                   const cuuint32_t *box)
{
    EncodeTiledFn fn = encode_fn();
    if (!fn) return false;
    cuuint32_t es[5] = {1, 1, 1, 1, 1};
    return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, (cuuint32_t)rank, const_cast<void *>(base), dims, strides_bytes, box, es,
              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static bool make_maps(Maps &M, const double *data, int64_t sys_stride, int nsys, const Dims &D, const double *b, double *x,
                      const VecLayout &V)
{
    const cuuint64_t N = (cuuint64_t)D.nblocks, ld8 = (cuuint64_t)D.ld * 8;
    for (int p = 0; p < 4; ++p) {
        const cuuint64_t dims[4] = {32, 32, N, (cuuint64_t)nsys};
        const cuuint64_t str[3] = {ld8, 32 * ld8, (cuuint64_t)sys_stride * 8};
        const cuuint32_t box[4] = {(cuuint32_t)pan_hp(p), 8, GB, 1};
        if (!encode(&M.pan[p], data, 4, dims, str, box)) return false;
    }
    {
        // a single vector per system: the rhs stride is never used, but it has to be a legal stride
        const cuuint64_t rs = V.nrhs > 1 ? (cuuint64_t)V.rhs_stride * 8 : 256 * N;
        const cuuint64_t ss = nsys > 1 ? (cuuint64_t)V.sys_stride * 8 : rs * (cuuint64_t)V.nrhs;
        const cuuint64_t dims[5] = {8, 4, N, (cuuint64_t)V.nrhs, (cuuint64_t)nsys};
        const cuuint64_t str[4] = {64, 256, rs, ss};
        const cuuint32_t box[5] = {8, 1, GB, 1, 1};
        if (!encode(&M.b, b, 5, dims, str, box)) return false;
        if (!encode(&M.x, x, 5, dims, str, box)) return false;
    }
    for (int d = 0; d < MAXND; ++d) {
        const int dd = d < D.ndiag ? d : 0;  // unused maps: any legal one
        const cuuint64_t rows = (cuuint64_t)(D.nblocks - 1 - dd);
        const cuuint64_t dims[4] = {8, 4, rows, (cuuint64_t)nsys};
        const cuuint64_t str[3] = {64, ld8, (cuuint64_t)sys_stride * 8};
        const cuuint32_t box[4] = {8, 1, GB, 1};
        if (!encode(&M.sub[d], data + D.off_sub + D.diag_chunk(dd, 0), 4, dims, str, box)) return false;
        if (!encode(&M.sup[d], data + D.off_sup + D.diag_chunk(dd, 0), 4, dims, str, box)) return false;
    }
    return true;
}

}  // namespace skew

cudaError_t skew_stats(unsigned long long out[3], bool reset)
{
    cudaError_t e = cudaMemcpyFromSymbol(out, skew::g_stats, sizeof(unsigned long long) * 3);
    if (e == cudaSuccess && reset) {
        const unsigned long long z[3] = {0, 0, 0};
        e = cudaMemcpyToSymbol(skew::g_stats, z, sizeof z);
    }
    return e;
}

bool try_launch_solve_skew32(unsigned nvec, const double *data, int64_t sys_stride, const Dims &D, const int *rowbycol,
                             const int *colbyrow, const double *b, double *x, const VecLayout &V, int *flags, int *status,
                             cudaStream_t st, cudaError_t *err)
{
    using namespace skew;
    (void)rowbycol;
    *err = cudaSuccess;
    if (D.w != 32 || D.nsat != 0 || D.ndiag < 1 || D.ndiag > MAXND || (D.ld & 1) || D.nblocks < 8) return false;
    if ((sys_stride & 1) || (V.rhs_stride & 1) || (V.sys_stride & 1)) return false;
    if ((reinterpret_cast<uintptr_t>(data) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(x) |
         reinterpret_cast<uintptr_t>(colbyrow)) & 15)
        return false;
    const int nsys = (int)(nvec / V.nrhs);
    if (nsys > 65535) return false;
    Maps M;
    if (!make_maps(M, data, sys_stride, nsys, D, b, x, V)) return false;

    // which systems pivoted
    *err = cudaMemsetAsync(status, 0, sizeof(int) * nvec, st);
    if (*err != cudaSuccess) return true;
    const int64_t n_per_sys = (int64_t)D.nblocks * 32;
    const unsigned gx = (unsigned)std::min<int64_t>((n_per_sys / 4 + 255) / 256, 64);
    perm_scan_kernel<<<dim3(gx, nsys), 256, 0, st>>>(colbyrow, n_per_sys, V.nrhs, status);
    const size_t smem = (size_t)NSLOT * slot_bytes(D.ndiag) + TAB_BYTES + 64 + 256 + 128;
    if (D.ndiag == 1) {
        *err = cudaFuncSetAttribute(solve_skew32_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (*err != cudaSuccess) return true;
        solve_skew32_kernel<1><<<nvec, 64, smem, st>>>(M, D.nblocks, V.nrhs, x, V.rhs_stride, V.sys_stride, flags, status);
    } else {
        *err = cudaFuncSetAttribute(solve_skew32_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (*err != cudaSuccess) return true;
        solve_skew32_kernel<2><<<nvec, 64, smem, st>>>(M, D.nblocks, V.nrhs, x, V.rhs_stride, V.sys_stride, flags, status);
    }
    count_launch(2);
    *err = cudaGetLastError();
    return true;
}

}  // namespace bdilu
