// Kernel 2, blockw == 32 fast path, third generation ("quad32"): ILU_inplace<double>::solve
// (block_diag_ilu.hpp:486-538) as a wavefront over consecutive blocks, one CTA per (system, right-hand side).
//
// What bounds this sweep is not arithmetic but the length of the instruction sequence between "my predecessor
// published a value" and "I publish the value that needed it" (in-order issue: every dependent instruction costs
// 4-10 cycles, a shuffle ~25, a shared-memory round trip ~60).  Measured on B200 (profiles/), the one-row-per-step
// kernel spent ~230 cycles per row and its single producer warp ~560 cycles per block.  Hence:
//
//   * FOUR rows per step.  The four accumulators of a quad are broadcast by shuffle together with the six entries
//     of the 4x4 triangle that couples them; every lane solves that triangle redundantly (three dependent DFMAs)
//     and applies the four finished values to its own accumulator.  Each row still sees exactly the operations
//     of the one-row sweep, in the same order: a - l0*y0 - l1*y1 - ..., so the result is unchanged.
//   * The part of a row that belongs to the other triangle (U entries in the forward sweep, L entries in the
//     backward sweep) is zeroed when the row is copied into registers.  A finished lane then keeps its value
//     through all later steps (it multiplies zeros), nothing has to be captured or selected.
//   * Coupling terms (sub/sup diagonals times the predecessors' values, :508-511, :524-527) are merged per quad,
//     one step ahead, from a shared-memory ring whose "not yet written" state is an in-band signalling-NaN
//     pattern; the poll is a warp-uniform vote on values loaded before the shuffles of the previous quad.
//   * Stages hold GROUPS of G consecutive blocks: blocks, colbyrow, rowbycol, b and the sub/sup rows of G
//     consecutive blocks are each contiguous in memory, so one TMA bulk copy per piece feeds G blocks and the
//     producer's issue cost per block drops G-fold; two producer warps alternate groups.
//   * U rows are pre-scaled by 1/U[r][r], so the backward chain has no division on it.
//
// When b holds a NaN or U a zero diagonal the return flags are exact (:500-501, :533-534); the polluted x then
// holds NaN/inf as the reference's does, but not necessarily in the same elements (0 * NaN in the zeroed part).
#include <cstdio>
#include <cstdlib>

#include "bdilu_wave_common.cuh"

namespace bdilu {

namespace quad {

using namespace wave;  // mbarrier / bulk-copy / volatile shared-memory helpers, MAXD, SENT

#ifndef Q32_ARRIVE_ROW
#define Q32_ARRIVE_ROW 0  // lean chain: the successor is woken once this row is published.  Measured on cfg3: after row 0
                          // 2.60 ms, after row PG-1 = 1 (its first poll then never fails) 2.69 ms, before the chain (-1) 2.75 ms
#endif
constexpr int W = 32;
constexpr int G = 4;    // blocks per stage group
#ifndef Q32_NP
#define Q32_NP 2
#endif
#ifndef Q32_QW
#define Q32_QW 13  // measured on cfg3 (solve ms): 10: 2.73, 11: 2.63, 12: 2.57, 13: 2.56, 14: 2.59, 15 with one producer: 2.94
#endif
constexpr int NP = Q32_NP;   // producer warps
constexpr int QW = Q32_QW;  // consumer warps (QW + NP <= 16 warps, so each thread may hold 128 registers)
constexpr int RS = 32;  // ring slots (power of two, >= QW + MAXD + 1: see the reset protocol below)
static_assert(QW + MAXD + 1 <= RS, "ring too small for the reset protocol");

// Phase profile (debug builds only, -DBDILU_W32_PROF): cycles summed over consumer warps / producer warps.
//   [0] wait for the stage  [1] stage -> registers  [2] first coupling poll  [3] the 8-step chain  [4] store
//   [5] producer: wait for an empty stage  [6] producer: issue  [7] blocks counted
#ifdef BDILU_W32_PROF
__device__ unsigned long long g_prof[8];
#define Q32_T(var) const long long var = clock64()
#define Q32_ADD(slot, a, b) if (r == 0) atomicAdd(&g_prof[slot], (unsigned long long)((b) - (a)))
#else
#define Q32_T(var)
#define Q32_ADD(slot, a, b)
#endif

// Production builds: no wait can hang unnoticed: the producer warps' waits are clock-bounded and act as the watchdog of the
// consumers (see produce()); a protocol failure ends in a trap = cudaErrorLaunchFailure, reported as BDILU_ERR_CUDA.
// Debug builds (-DBDILU_SPIN_GUARD): a wait that spins for too long also prints where it is stuck.
#ifdef BDILU_SPIN_GUARD
#define Q32_GUARD_DECL long long guard_n = 0
#define Q32_GUARD(what, a, b2, c)                                                                          \
    if (++guard_n > (1ll << 22)) {                                                                         \
        printf("quad32 stuck: %s cta %d warp %d lane %d  i=%d a=%d b=%d\n", what, (int)blockIdx.x,          \
               (int)(threadIdx.x >> 5), (int)(threadIdx.x & 31), (int)(a), (int)(b2), (int)(c));           \
        __trap();                                                                                          \
    }
__device__ __forceinline__ void mbar_wait_guard(void *bar, unsigned parity, const char *what, int a, int b2, int c)
{
    Q32_GUARD_DECL;
    while (!mbar_try(bar, parity)) { Q32_GUARD(what, a, b2, c) }
}
#else
#define Q32_GUARD_DECL
#define Q32_GUARD(what, a, b2, c)
#define mbar_wait_guard(bar, parity, what, a, b2, c) mbar_wait_unbounded(bar, parity)
#endif

// byte layout of one stage = one group of G blocks (every piece 16-byte aligned since ld is even)
struct GroupLayout {
    int blk_bytes, ld8;                   // one block; one sub/sup row
    int o_cbr, o_rbc, o_vec, o_diag, bytes;  // blocks at offset 0
};
__host__ __device__ inline GroupLayout group_layout(int ld, int ndiag)
{
    GroupLayout g;
    g.blk_bytes = W * ld * 8;
    g.ld8 = ld * 8;
    g.o_cbr = G * g.blk_bytes;
    g.o_rbc = g.o_cbr + G * W * 4;
    g.o_vec = g.o_rbc + G * W * 4;
    g.o_diag = g.o_vec + G * W * 8;
    g.bytes = g.o_diag + ndiag * G * g.ld8;
    return g;
}

struct Ctx {
    const double *sd;      // this system's packed data
    const int *rbc, *cbr;  // this system's permutation vectors
    const double *b;       // this vector
    double *x;
    Dims D;
    GroupLayout GL;
    int NS, NG;            // stages; groups per sweep
    int dbg;
    unsigned char *stages;
    unsigned long long *full, *empty;
    unsigned long long *started;  // [RS]: block of that ring slot has published its first quad
    unsigned long long *sweep_done;  // one arrival per consumer warp at the end of each sweep (the producers' watchdog)
    double *ring;          // [RS][32]
    int *flagw;            // [2]: nan_first, zero_last
};

__device__ __forceinline__ void ring_reset(const Ctx &C, int tid, int nthreads, int nd)
{
    for (int t = tid; t < RS * 32; t += nthreads)
        C.ring[t] = ((t >> 5) >= RS - nd) ? 0.0 : __longlong_as_double((long long)SENT);
}

// ---- producer warps: lane = piece, one bulk copy per piece per GROUP --------------------------------------------
// Group g of a sweep holds the blocks of sweep orders gG .. gG+nb-1.  Forward: blocks b_lo = gG upwards, block b in
// slot b - b_lo.  Backward: blocks b_hi = N-1-gG downwards; they are contiguous in memory from b_lo = b_hi-nb+1 and
// are placed so that block b sits in slot b - (b_hi-G+1), i.e. sweep order i -> slot G-1 - i%G, partial group or not.
// gseq (0 .. 2NG-1) numbers the groups of both sweeps; stage = gseq % NS, barrier phase = (gseq / NS) & 1.
template <bool FWD, int ND>
__device__ void produce(const Ctx &C, int p, int lane)
{
    const int N = C.D.nblocks, ld = C.D.ld;
    const GroupLayout &GL = C.GL;
    const int np = (C.dbg & 2) ? 1 : NP;
    if (p >= np) return;
    for (int g = p; g < C.NG; g += np) {
#ifdef BDILU_W32_PROF
        const int r = lane;
#endif
        Q32_T(p0);
        const int gseq = (FWD ? 0 : C.NG) + g;
        const int s = gseq % C.NS, use = gseq / C.NS;
        if (use >= 1) mbar_wait_timed(&C.empty[s], (unsigned)(use - 1) & 1u);  // watchdog: see the header
        Q32_T(p1);
        const int nb = min(G, N - g * G);
        const int b_lo = FWD ? g * G : N - g * G - nb;
        const int mo = FWD ? 0 : G - nb;  // slot of block b_lo
        const char *src = nullptr;
        unsigned bytes = 0, dst = 0;
        if (lane == 0) {
            src = reinterpret_cast<const char *>(C.sd + C.D.block(b_lo));
            bytes = ((C.dbg & 4) ? 1 : nb) * GL.blk_bytes;
            dst = mo * GL.blk_bytes;
        } else if ((C.dbg & 4) && lane >= 8 && lane < 8 + nb - 1) {
            const int k = lane - 7;  // blocks 1..nb-1 of the group, one copy each
            src = reinterpret_cast<const char *>(C.sd + C.D.block(b_lo + k));
            bytes = GL.blk_bytes;
            dst = (mo + k) * GL.blk_bytes;
        } else if (lane == 1) {
            src = reinterpret_cast<const char *>(C.cbr + (int64_t)b_lo * W);
            bytes = nb * W * 4;
            dst = GL.o_cbr + mo * W * 4;
        } else if (lane == 2 && FWD) {
            src = reinterpret_cast<const char *>(C.rbc + (int64_t)b_lo * W);
            bytes = nb * W * 4;
            dst = GL.o_rbc + mo * W * 4;
        } else if (lane == 3 && FWD) {
            src = reinterpret_cast<const char *>(C.b + (int64_t)b_lo * W);
            bytes = nb * W * 8;
            dst = GL.o_vec + mo * W * 8;
        } else if (lane >= 4 && lane < 4 + ND) {
            const int d = lane - 4;
            // forward: block b needs sub(d, b-1-d, :) (exists for b >= d+1); backward: sup(d, b, :) (b <= N-2-d)
            const int bs = FWD ? max(b_lo, d + 1) : b_lo;
            const int be = FWD ? b_lo + nb - 1 : min(b_lo + nb - 1, N - 2 - d);
            if (be >= bs) {
                const double *q = FWD ? C.sd + C.D.off_sub + C.D.diag_chunk(d, bs - 1 - d)
                                      : C.sd + C.D.off_sup + C.D.diag_chunk(d, bs);
                src = reinterpret_cast<const char *>(q);
                bytes = (be - bs + 1) * GL.ld8;
                dst = GL.o_diag + d * G * GL.ld8 + (mo + bs - b_lo) * GL.ld8;
            }
        }
        const unsigned total = __reduce_add_sync(FULL, bytes);
        if (lane == 0) mbar_expect_tx(&C.full[s], total);
        __syncwarp();
        if (bytes != 0) bulk_g2s(C.stages + (size_t)s * GL.bytes + dst, src, bytes, &C.full[s]);
        Q32_T(p2);
        Q32_ADD(5, p0, p1);
        Q32_ADD(6, p1, p2);
        (void)ld;
    }
    // Watchdog.  The consumers' waits and polls are unbounded (their instruction stream is too tight for a counter), but
    // none of them can hang unnoticed: a stuck consumer stops releasing stages, so a producer's wait above times out; and
    // after the last group the first producer warp waits, bounded, for every consumer warp to finish the sweep.
    if (p == 0) mbar_wait_timed(C.sweep_done, FWD ? 0u : 1u);
}

// ---- consumers: lane = row, warp j owns sweep orders j, j+QW, ... -------------------------------------------------
// Ring protocol (RS = 32 slots of 32 doubles, slot of sweep order i = i & 31):
//   * block i publishes row values into slot i and polls slots i-1 .. i-ND ("not yet written" = SENT);
//   * after its LAST coupling read, and BEFORE it publishes its last quad, block i resets slot i-ND to SENT.  When
//     block i has read everything it needs, blocks i-1 .. i-ND+1 have published everything, hence finished reading
//     slot i-ND too (every row is read before the quad that needs it is published), so block i is its last reader.
//     The next writer of that slot, block i-ND+32, is started by a warp that first finished block i-ND+32-QW >= i+1
//     (QW + ND + 1 <= 32), which cannot complete before block i published its last quad -> the reset is ordered
//     before the reuse, with no counters and no extra waiting.
// MODE 0: quads by shuffle (every lane solves the 4x4 triangle);  MODE 4 / 2 / 1: lean chain, one row per step
// through shared memory, coupling merged every 4 / 2 / 1 rows.
template <bool FWD, int ND, int MODE, bool LD32>
__device__ void consume(const Ctx &C, int j, int r, int &nan_first, int &zero_last)
{
    const int N = C.D.nblocks, ld = LD32 ? W : C.D.ld;  // LD32: tight blocks, every stage offset is an immediate
    const GroupLayout &GL = C.GL;
    const int q = FWD ? r : W - 1 - r;  // position of my row in sweep order
    constexpr int PG = (MODE == 0) ? 4 : MODE;  // rows per coupling group
    constexpr int NGRP = W / PG;
    const int myquad = q / PG;              // coupling group of my row
    // phase numbering of the "started" barriers: forward blocks count from 0, backward blocks from 32*ceil(N/32)
    const int ibase = FWD ? 0 : ((N + RS - 1) & ~(RS - 1));
    double ynext = 0.0;
    if (!FWD && j < N) ynext = C.x[(int64_t)(N - 1 - j) * W + r];
    // stage / barrier phase of the group my block sits in, advanced incrementally (no division per block)
    int gcur = (FWD ? 0 : C.NG) + j / G;
    int s = gcur % C.NS;
    unsigned ph = (unsigned)(gcur / C.NS) & 1u;

    for (int i = j; i < N; i += QW) {
        const int b = FWD ? i : N - 1 - i;
        const double y_cur = ynext;
        if (!FWD && i + QW < N) ynext = C.x[(int64_t)(N - 1 - (i + QW)) * W + r];

        Q32_T(t0);
        const int mo = i % G;
        mbar_wait_guard(&C.full[s], ph, FWD ? "cons full fwd" : "cons full bwd", i, s, gcur);
        Q32_T(t1);
        const int m = FWD ? mo : G - 1 - mo;  // slot of my block inside the group
        const unsigned char *st = C.stages + (size_t)s * GL.bytes;
        const double *Ls = reinterpret_cast<const double *>(st + (size_t)m * GL.blk_bytes);
        // my row.  The shuffle variant needs the other triangle (and the diagonal) to read as zero; the lean chain
        // does not care: a finished lane's accumulator is dead, its value is read back from the ring at the end.
        double l[W];
#pragma unroll
        for (int k = 0; k < W; ++k) {
            const bool mine = (MODE != 0) || (FWD ? (k < r) : (k > r));
            l[k] = mine ? Ls[k * ld + r] : 0.0;
        }
        const int ci = reinterpret_cast<const int *>(st + GL.o_cbr)[m * W + r];
        const int nd = min(ND, i);
        double cv[ND];
#pragma unroll
        for (int d = 0; d < ND; ++d)
            cv[d] = (d < nd) ? reinterpret_cast<const double *>(st + GL.o_diag + (d * G + m) * GL.ld8)[ci] : 0.0;
        double acc;
        if (FWD) {
            const int src = reinterpret_cast<const int *>(st + GL.o_rbc)[m * W + r];
            const double bv = reinterpret_cast<const double *>(st + GL.o_vec)[m * W + src];
            if (isnan(bv)) nan_first = min(nan_first, b * W + src + 1);
            // x + (-0.0) == x for every x, but quiets a signalling NaN: the sentinel can never be published
            asm volatile("add.rn.f64 %0, %1, %2;" : "=d"(acc) : "d"(bv), "d"(-0.0));
        } else {
            const double dg = Ls[r * ld + r];
            if (dg == 0.0) zero_last = max(zero_last, b * W + r);
            const double rinv = rcp_fast_ok(dg) ? rcp_fast(dg) : 1.0 / dg;
            acc = y_cur * rinv;
#pragma unroll
            for (int k = 0; k < W; ++k) l[k] *= rinv;
#pragma unroll
            for (int d = 0; d < ND; ++d) cv[d] *= rinv;
        }
        __syncwarp();
        if (r == 0) {
            // everything this block needs is in registers; the last block of a partial group stands in for the
            // missing ones so that the stage's "empty" barrier (G arrivals) still completes
            const int extra = (i == N - 1) ? (G - 1 - mo) : 0;
            for (int e = 0; e <= extra; ++e) mbar_arrive(&C.empty[s]);
        }
        {
            const int gn = (FWD ? 0 : C.NG) + (i + QW) / G;
            s += gn - gcur;
            gcur = gn;
            while (s >= C.NS) {
                s -= C.NS;
                ph ^= 1u;
            }
        }
        Q32_T(t2);

        double *myslot = C.ring + (i & (RS - 1)) * 32 + r;
        double *resetslot = C.ring + ((i - ND) & (RS - 1)) * 32 + r;
        const double *pd[ND];
#pragma unroll
        for (int d = 0; d < ND; ++d) pd[d] = C.ring + ((i - 1 - d) & (RS - 1)) * 32 + ci;

        double v[ND];
#pragma unroll
        for (int d = 0; d < ND; ++d) v[d] = 0.0;
        // only the lanes of quad `qd` read the ring: a spinning warp costs one shared-memory wavefront per load
        auto load_pred = [&](int qd) {
            if (myquad == qd) {
#pragma unroll
                for (int d = 0; d < ND; ++d) v[d] = ld_volatile_s(pd[d]);
            }
        };
        // Warp-uniform wait until the lanes of quad `qd` hold real predecessor values in v[] (no lane ever spins on
        // its own, so the warp never splits); the first test uses values loaded before the previous quad's shuffles.
        auto merge_quad = [&](int qd) {
            const bool mine = (myquad == qd);
            double t;
            Q32_GUARD_DECL;
            for (;;) {
                Q32_GUARD(FWD ? "merge fwd" : "merge bwd", i, qd, ci)
                // speculative: the DFMAs run while the vote below decides whether v[] was real
                t = acc;
#pragma unroll
                for (int d = ND - 1; d >= 0; --d) t = fma(-cv[d], v[d], t);
                bool bad = false;
#pragma unroll
                for (int d = 0; d < ND; ++d) bad |= is_sent(v[d]);
                if (!__any_sync(FULL, mine && bad)) break;
                load_pred(qd);
            }
            if (mine) acc = t;
        };
        // sweep positions 4*jq .. 4*jq+3 live in lanes LN(jq, 0..3); CF(jq, n) is my coefficient for position 4*jq+n
#define LN(jq, n) (FWD ? 4 * (jq) + (n) : W - 1 - (4 * (jq) + (n)))
#define CF(jq, n) l[LN(jq, n)]
        // the six entries of quad jq's 4x4 triangle, broadcast from the lanes that own those rows
        auto triangle = [&](int jq, double (&t)[6]) {
            t[0] = group_bcast<32>(CF(jq, 0), LN(jq, 1));
            t[1] = group_bcast<32>(CF(jq, 0), LN(jq, 2));
            t[2] = group_bcast<32>(CF(jq, 1), LN(jq, 2));
            t[3] = group_bcast<32>(CF(jq, 0), LN(jq, 3));
            t[4] = group_bcast<32>(CF(jq, 1), LN(jq, 3));
            t[5] = group_bcast<32>(CF(jq, 2), LN(jq, 3));
        };

        // Sleep (mbarrier, no polling traffic) until my predecessor has published its first group; from then on it
        // stays about one group ahead of me and the in-band polls below almost always succeed at once.
        mbar_wait_guard(&C.started[(i - 1) & (RS - 1)], (unsigned)((ibase + i - 1) >> 5) & 1u,
                        FWD ? "started fwd" : "started bwd", i, 0, 0);
        load_pred(0);
        if constexpr (MODE == 0) {
            double tri[6];
            triangle(0, tri);
            merge_quad(0);
            Q32_T(t3);
#pragma unroll
            for (int j4 = 0; j4 < W / 4; ++j4) {
                if (j4 + 1 < W / 4) load_pred(j4 + 1);  // for the merge at the end of this step
                const double a0 = group_bcast<32>(acc, LN(j4, 0));
                const double a1 = group_bcast<32>(acc, LN(j4, 1));
                const double a2 = group_bcast<32>(acc, LN(j4, 2));
                const double a3 = group_bcast<32>(acc, LN(j4, 3));
                double trin[6];
                if (j4 + 1 < W / 4) triangle(j4 + 1, trin);  // next step's triangle: fills the shuffle latency
                const double y0 = a0;
                const double y1 = fma(-tri[0], y0, a1);
                const double y2 = fma(-tri[2], y1, fma(-tri[1], y0, a2));
                const double y3 = fma(-tri[5], y2, fma(-tri[4], y1, fma(-tri[3], y0, a3)));
                acc = fma(-CF(j4, 0), y0, acc);
                acc = fma(-CF(j4, 1), y1, acc);
                acc = fma(-CF(j4, 2), y2, acc);
                acc = fma(-CF(j4, 3), y3, acc);
                if (myquad == j4) st_volatile_s(myslot, acc);
                if (j4 == 0) {
                    __syncwarp();
                    if (r == 0) mbar_arrive(&C.started[i & (RS - 1)]);
                }
                if (j4 + 1 < W / 4) {
                    merge_quad(j4 + 1);
                    if (j4 + 2 == W / 4)  // all my coupling reads are done: recycle the slot of block i-ND
                        st_volatile_s(resetslot, __longlong_as_double((long long)SENT));
#pragma unroll
                    for (int e = 0; e < 6; ++e) tri[e] = trin[e];
                }
            }
        } else {
            // Lean chain: the lane whose row is finished stores it into the ring (that store IS the publication) and
            // everybody reads it back with one broadcast load: three instructions per row, no shuffles, no selects.
            merge_quad(0);
            Q32_T(t3);
            const double *slotbase = C.ring + (i & (RS - 1)) * 32;
            if (Q32_ARRIVE_ROW < 0) {
                __syncwarp();
                if (r == 0) mbar_arrive(&C.started[i & (RS - 1)]);
            }
#pragma unroll
            for (int kk = 0; kk < W; ++kk) {
                const int k = FWD ? kk : W - 1 - kk;
                const int g = kk / PG;
                if (kk % PG == 0 && g + 1 < NGRP) load_pred(g + 1);  // for the merge at the end of this group
                if (r == k) st_volatile_s(myslot, acc);
                if (kk + 1 < W) {
                    const double vk = ld_volatile_s(slotbase + k);
                    acc = fma(-l[k], vk, acc);
                }
                if (kk == Q32_ARRIVE_ROW) {
                    __syncwarp();
                    if (r == 0) mbar_arrive(&C.started[i & (RS - 1)]);
                }
                if (kk % PG == PG - 1 && g + 1 < NGRP) {
                    merge_quad(g + 1);
                    if (g + 2 == NGRP)  // all my coupling reads are done: recycle the slot of block i-ND
                        st_volatile_s(resetslot, __longlong_as_double((long long)SENT));
                }
            }
        }
#undef LN
#undef CF
        Q32_T(t4);
        C.x[(int64_t)b * W + r] = (MODE == 0) ? acc : ld_volatile_s(myslot);
        Q32_T(t5);
        Q32_ADD(0, t0, t1);
        Q32_ADD(1, t1, t2);
        Q32_ADD(2, t2, t3);
        Q32_ADD(3, t3, t4);
        Q32_ADD(4, t4, t5);
        Q32_ADD(7, 0, 1);
    }
    __syncwarp();
    if (r == 0) mbar_arrive(C.sweep_done);
}

template <int ND, int MODE, bool LD32>
__global__ void __launch_bounds__((QW + NP) * 32, 1)
solve_quad32_kernel(const double *__restrict__ data, int64_t sys_stride, Dims D, const int *__restrict__ rowbycol,
                    const int *__restrict__ colbyrow, const double *__restrict__ bvec, double *xvec, VecLayout V,
                    int *__restrict__ flags, int NS, int dbg, const int *__restrict__ status)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int vecid = blockIdx.x;
    // second pass behind the systolic kernel (bdilu_solve_skew.cu): only the vectors it handed back (status != 0)
    if (status != nullptr && status[vecid] == 0) return;
    const int sys = vecid / V.nrhs, rhs = vecid % V.nrhs;
    const int N = D.nblocks;

    Ctx C;
    C.dbg = dbg;
    C.sd = data + (int64_t)sys * sys_stride;
    C.rbc = rowbycol + (int64_t)sys * N * W;
    C.cbr = colbyrow + (int64_t)sys * N * W;
    C.b = bvec + (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride;
    C.x = xvec + (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride;
    C.D = D;
    C.GL = group_layout(D.ld, ND);
    C.NS = NS;
    C.NG = (N + G - 1) / G;
    C.stages = smem_raw;
    size_t off = (size_t)NS * C.GL.bytes;
    C.full = reinterpret_cast<unsigned long long *>(smem_raw + off);
    C.empty = C.full + NS;
    off += (size_t)2 * NS * 8;
    C.started = C.empty + NS;
    C.sweep_done = C.started + RS;
    off += (size_t)(RS + 1) * 8;
    C.ring = reinterpret_cast<double *>(smem_raw + off);
    off += (size_t)RS * 32 * 8;
    C.flagw = reinterpret_cast<int *>(smem_raw + off);

    if (tid == 0) {
        for (int s = 0; s < NS; ++s) {
            mbar_init(&C.full[s], 1);
            mbar_init(&C.empty[s], G);
        }
        for (int t = 0; t < RS; ++t) mbar_init(&C.started[t], 1);
        mbar_init(C.sweep_done, QW);
        C.flagw[0] = 0x7fffffff;
        C.flagw[1] = -1;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    ring_reset(C, tid, blockDim.x, ND);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();

    int nan_first = 0x7fffffff, zero_last = -1;
    if (warp >= QW) produce<true, ND>(C, warp - QW, lane);
    else consume<true, ND, MODE, LD32>(C, warp, lane, nan_first, zero_last);
    __syncthreads();  // y (in the x buffer) is complete and visible to the whole CTA
    ring_reset(C, tid, blockDim.x, ND);
    // "started" barriers keep counting phases across the sweeps: slots that saw one block fewer in the forward sweep
    // take a ghost arrival, so every slot has completed ceil(N/32) phases and the backward sweep numbers its blocks
    // from 32*ceil(N/32) (see consume(): ibase)
    if (tid < RS && (N & (RS - 1)) != 0 && tid >= (N & (RS - 1))) mbar_arrive(&C.started[tid]);
    __syncthreads();
    if (warp >= QW) produce<false, ND>(C, warp - QW, lane);
    else consume<false, ND, MODE, LD32>(C, warp, lane, nan_first, zero_last);
    if (warp < QW) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            nan_first = min(nan_first, __shfl_xor_sync(FULL, nan_first, o));
            zero_last = max(zero_last, __shfl_xor_sync(FULL, zero_last, o));
        }
        if (lane == 0) {
            atomicMin(&C.flagw[0], nan_first);
            atomicMax(&C.flagw[1], zero_last);
        }
    }
    __syncthreads();
    if (tid == 0 && flags) {
        const int nf = C.flagw[0], zl = C.flagw[1];
        flags[vecid] = (nf != 0x7fffffff) ? nf : (zl >= 0 ? N * W + zl : 0);
    }
}

// Timing experiments (builds with -DBDILU_QUAD_VARIANTS=<mask> only; every variant gives correct results):
//   2 = one producer warp, 4 = one bulk copy per block instead of per group, 32 / 64 / 128 = chain variant: shuffle
//   quads (32), lean chain merging the coupling every 4 rows (64) or every row (128).  Product builds: lean chain,
//   coupling merged every 2 rows (the fastest on B200); nothing is read from the environment.
static int debug_switches()
{
#ifdef BDILU_QUAD_VARIANTS
    return BDILU_QUAD_VARIANTS;
#else
    return 0;
#endif
}

template <int ND, int MODE, bool LD32>
static cudaError_t launch_one(unsigned nvec, size_t smem, int NS, cudaStream_t st, const double *data, int64_t sys_stride,
                              const Dims &D, const int *rowbycol, const int *colbyrow, const double *b, double *x,
                              const VecLayout &V, int *flags, const int *status)
{
    cudaError_t e = cudaFuncSetAttribute(solve_quad32_kernel<ND, MODE, LD32>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         227 * 1024);
    if (e != cudaSuccess) return e;
    solve_quad32_kernel<ND, MODE, LD32><<<nvec, (QW + NP) * 32, smem, st>>>(data, sys_stride, D, rowbycol, colbyrow, b, x,
                                                                            V, flags, NS, debug_switches(), status);
    return cudaGetLastError();
}

template <int ND>
static cudaError_t launch_nd(unsigned nvec, size_t smem, int NS, cudaStream_t st, const double *data, int64_t sys_stride,
                             const Dims &D, const int *rowbycol, const int *colbyrow, const double *b, double *x,
                             const VecLayout &V, int *flags, const int *status)
{
#define Q32_ARGS nvec, smem, NS, st, data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, status
#ifdef BDILU_QUAD_VARIANTS  // the other chain variants, for timing comparisons
    const int dbg = debug_switches();
    if (dbg & 32) return launch_one<ND, 0, false>(Q32_ARGS);
    if (dbg & 64) return launch_one<ND, 4, false>(Q32_ARGS);
    if (dbg & 128) return launch_one<ND, 1, false>(Q32_ARGS);
#endif
    if (D.ld == W) return launch_one<ND, 2, true>(Q32_ARGS);
    return launch_one<ND, 2, false>(Q32_ARGS);
#undef Q32_ARGS
}

}  // namespace quad

#ifdef BDILU_W32_PROF
extern "C" __attribute__((visibility("default"))) int bdilu_debug_q32prof(unsigned long long *out8, int reset)
{
    if (out8) cudaMemcpyFromSymbol(out8, quad::g_prof, sizeof(unsigned long long) * 8);
    if (reset) {
        unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        cudaMemcpyToSymbol(quad::g_prof, z, sizeof z);
    }
    return 0;
}
#endif

// blockw == 32, 1 <= ndiag <= 4, nsat == 0, even ld, 16-byte aligned arrays (checked by try_launch_solve_wave).
// Returns false when the shape does not fit (too few stages): the caller falls back to the older kernels.
bool try_launch_solve_quad32(unsigned nvec, const double *data, int64_t sys_stride, const Dims &D, const int *rowbycol,
                             const int *colbyrow, const double *b, double *x, const VecLayout &V, int *flags,
                             cudaStream_t st, cudaError_t *err, const int *status)
{
    using namespace quad;
    const GroupLayout GL = group_layout(D.ld, D.ndiag);
    const size_t fixed = (size_t)RS * 32 * 8 + (size_t)(RS + 1) * 8 + 64;
    const size_t budget = 227 * 1024;
    int NS = (int)((budget - fixed) / (GL.bytes + 16));
    if (NS > 16) NS = 16;
    if (NS * G < QW + 2 * G) return false;  // the stages must cover every consumer warp plus groups in flight
    const size_t smem = (size_t)NS * (GL.bytes + 16) + fixed;
    switch (D.ndiag) {
    case 1: *err = launch_nd<1>(nvec, smem, NS, st, data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, status); break;
    case 2: *err = launch_nd<2>(nvec, smem, NS, st, data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, status); break;
    case 3: *err = launch_nd<3>(nvec, smem, NS, st, data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, status); break;
    default: *err = launch_nd<4>(nvec, smem, NS, st, data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, status); break;
    }
    return true;
}

}  // namespace bdilu
