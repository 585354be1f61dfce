// Kernel 2, narrow blocks (blockw <= 16; also blockw <= 32 when satellites rule out the wavefront kernels): ILU_inplace<double>::solve (block_diag_ilu.hpp:486-538) with one lane group per
// (system, right-hand side) and NOTHING but shuffles and FMAs on the dependent path of the chain.
//
// The general chain kernel (bdilu_solve.cu) spends ~2.3 us per block at blockw 8: every block waits for a string of
// dependent global-memory round trips (rowbycol -> b[src]; colbyrow -> sub[ci]; y of the previous blocks read back from
// the x buffer).  Here every global load has an address that depends on the block index and the lane only:
//   * lane r copies b[bi*w + r], rowbycol[r], colbyrow[r], its row of the block and sub(di, bi-1-di, r) / sup(di, bi, r),
//     PF blocks ahead of their use, into its own slots of a shared-memory ring (cp.async groups);
//   * the permuted reads become shuffles inside the group: b[rowbycol[r]] = shfl(b_r, rowbycol[r]),
//     sub(di, ., colbyrow[r]) = shfl(sub_r, colbyrow[r]), and y_{bi-1-di}[colbyrow[r]] = shfl(y kept in registers, ...);
//   * the triangular solve broadcasts each finished value by shuffle, as the general kernel does.
// The forward sweep performs the operations of bdilu_solve.cu (and of the reference's loop, :502-517) in the same order;
// the backward sweep scales each row by the reciprocal of its diagonal ahead of the chain instead of dividing on it.
// 1 <= ndiag <= 4; satellite couplings (nsat > 0) touch only the nsat blocks at either end of the chain and read the x
// buffer there, as the general kernel does.
#include "bdilu_common.cuh"
#include "bdilu_kernels.h"

namespace bdilu {
namespace pf {

constexpr int WARPS = 2;
#ifndef PF_QUAD_WARPS_PER_SM
#define PF_QUAD_WARPS_PER_SM 8
#endif

// Phase profile (debug builds only, -DBDILU_PF_PROF): cycles of warp 0 of CTA 0, forward sweep:
//   [0] wait_group + ring -> registers  [1] b gather + couplings  [2] the rows  [3] store, copies of the next block,
//   far prefetch, pointer updates  [4] blocks counted
#ifdef BDILU_PF_PROF
__device__ unsigned long long g_pfprof[8];
#define PF_T(v) const long long v = clock64()
#define PF_ADD(slot, a, b) if (threadIdx.x == 0 && blockIdx.x == 0) g_pfprof[slot] += (unsigned long long)((b) - (a))
#else
#define PF_T(v)
#define PF_ADD(slot, a, b)
#endif

template <int WP, int ND>
struct Stage {
    double m[WP];    // my row of the block
    double dg;       // backward: U[r][r]
    double vec;      // forward: b[bi*w + r]; backward: y[bi*w + r]
    double cf[ND];   // forward: sub(di, bi-1-di, r); backward: sup(di, bi, r)
    int cbr, rbc;
};

// Running pointers of the prefetch front (lane r's elements of the block PF ahead of the one being processed): one add
// per pointer and block instead of the 64-bit index arithmetic of Dims per load.
template <int ND>
struct Front {
    const double *blk;     // block(bi) + r
    const double *vec;     // b or y, + bi*w + r
    const int *cbr, *rbc;  // + bi*w + r
    const double *cf[ND];  // forward: sub(d, bi-1-d, r); backward: sup(d, bi, r)
    int bi;
};

// The ring lives in shared memory and is filled by cp.async: a register ring does not pipeline, because every load of
// every stage lands on the same hardware scoreboard and the first use of the oldest stage then waits for the newest
// loads as well (measured: ~900 cycles per block whatever the depth).  cp.async groups complete in order and
// wait_group<PF-1> waits for the oldest only.  Each thread copies and later reads its own slots: no barrier needed.
// Slot of item `it` of stage `s` for thread `tid`: ((s*NITEM + it)*THREADS + tid) doubles -> conflict-free.
template <int WP, int ND>
struct Ring {
    static constexpr int NITEM = WP + 3 + ND;  // row, dg, vec, cf[ND], (cbr | rbc)
};

__device__ __forceinline__ void cp_async8(unsigned dst, const void *src, bool have)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(dst), "l"(src), "r"(have ? 8 : 0) : "memory");
}
__device__ __forceinline__ void cp_async4(unsigned dst, const void *src, bool have)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(have ? 4 : 0) : "memory");
}

// issue the copies of the front's block into stage slot `slot` (shared-memory address of item 0 for this thread)
template <int WP, int ND, bool FWD, bool TIGHT>
__device__ __forceinline__ void issue_stage(unsigned slot, const Front<ND> &F, int w, int ld, int N, int r, bool rowok,
                                            bool inrange)
{
    constexpr unsigned IS = WARPS * 32 * 8;  // bytes between items
    const int ldk = TIGHT ? WP : ld;
    const bool ok = rowok && inrange;
    if (TIGHT && WP >= 4) {
        // tight blocks: the block is copied as an image, 16 bytes per lane and instruction, so that an instruction
        // touches one whole 128-byte line per group (at blockw 8) instead of half a line twice: the copies of a warp
        // are bound by the number of distinct lines they touch (~9 cycles each, scratch/pf_prof.py).  The first WP items
        // of the stage hold the images of the CTA's groups back to back; lanes read each other's copies after the barrier
        // in read_stage.
        // The rows (= columns of the block) of the image are rotated by the group index modulo 128 bytes' worth of rows, so
        // that the groups of a warp read from different banks (images are a multiple of 128 bytes apart).
        constexpr int ROT = (WP < 16) ? 16 / WP : 1;
        const unsigned g = threadIdx.x / WP;
        const unsigned img = slot - (threadIdx.x * 8u) + g * (WP * WP * 8u);  // my group's image
        const unsigned rot = g % ROT;
        const char *src = reinterpret_cast<const char *>(F.blk - r);
#pragma unroll
        for (int t = 0; t < WP / 2; ++t) {
            const unsigned e = 2u * (unsigned)(t * WP + r);            // first of the two elements of my chunk
            const unsigned kd = ((e / WP) + rot) & (WP - 1);           // image row it goes to
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(img + (kd * WP + (e % WP)) * 8u), "l"(src + e * 8u),
                         "r"(ok ? 16 : 0)
                         : "memory");
        }
    } else {
#pragma unroll
        for (int k = 0; k < WP; ++k) cp_async8(slot + k * IS, F.blk + (int64_t)k * ldk, ok && (TIGHT || k < w));
        if (!FWD) cp_async8(slot + WP * IS, F.blk + (int64_t)r * ldk, ok);
    }
    cp_async8(slot + (WP + 1) * IS, F.vec, ok);
#pragma unroll
    for (int d = 0; d < ND; ++d) {
        const bool have = ok && (FWD ? (F.bi - 1 - d >= 0) : (F.bi + 1 + d <= N - 1));
        cp_async8(slot + (WP + 2 + d) * IS, F.cf[d], have);
    }
    cp_async4(slot + (WP + 2 + ND) * IS, F.cbr, ok);
    if (FWD) cp_async4(slot + (WP + 2 + ND) * IS + 4, F.rbc, ok);
    asm volatile("cp.async.commit_group;" ::: "memory");
}

template <int WP, int ND, bool FWD, bool TIGHT>
__device__ __forceinline__ void read_stage(Stage<WP, ND> &S, const double *slot, int r)
{
    constexpr int IS = WARPS * 32;  // doubles between items
    if (TIGHT && WP >= 4) {
        __syncwarp();  // the image was copied by all lanes of my group (each waited for its own copies)
        constexpr int ROT = (WP < 16) ? 16 / WP : 1;
        const int g = threadIdx.x / WP, rot = g % ROT;
        const double *img = slot - threadIdx.x + g * (WP * WP);
#pragma unroll
        for (int k = 0; k < WP; ++k) S.m[k] = img[((k + rot) & (WP - 1)) * WP + r];
        S.dg = FWD ? 1.0 : img[((r + rot) & (WP - 1)) * WP + r];
    } else {
#pragma unroll
        for (int k = 0; k < WP; ++k) S.m[k] = slot[k * IS];
        S.dg = FWD ? 1.0 : slot[WP * IS];
    }
    S.vec = slot[(WP + 1) * IS];
#pragma unroll
    for (int d = 0; d < ND; ++d) S.cf[d] = slot[(WP + 2 + d) * IS];
    const int2 pr = *reinterpret_cast<const int2 *>(slot + (WP + 2 + ND) * IS);
    S.cbr = pr.x;
    S.rbc = FWD ? pr.y : 0;
}

template <int ND, bool FWD>
__device__ __forceinline__ void prefetch_far(const Front<ND> &F, int w, int64_t blk_step, int ld, int N, int r, int nlines,
                                             int pd)
{
    const int bi = FWD ? F.bi + pd : F.bi - pd;
    const int64_t sgn = FWD ? pd : -pd;
    const int a = r - nlines;  // 0: colbyrow, 1: b / y, 2: rowbycol (forward), 3..: sub / sup rows
    const char *p = reinterpret_cast<const char *>(F.blk - r + sgn * blk_step) + 128 * r;
    p = (a == 0) ? reinterpret_cast<const char *>(F.cbr + sgn * w) : p;
    p = (a == 1) ? reinterpret_cast<const char *>(F.vec + sgn * w) : p;
    p = (a == 2) ? reinterpret_cast<const char *>(FWD ? (const void *)(F.rbc + sgn * w) : (const void *)(F.cf[0] + sgn * ld)) : p;
    p = (a == 3) ? reinterpret_cast<const char *>(F.cf[(FWD || ND < 2) ? 0 : 1] + sgn * ld) : p;
    if (bi >= 0 && bi < N && a <= 3) asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

template <int ND, bool FWD>
__device__ __forceinline__ void advance(Front<ND> &F, int w, int64_t blk_step, int ld)
{
    if (FWD) {
        F.blk += blk_step; F.vec += w; F.cbr += w; F.rbc += w; ++F.bi;
#pragma unroll
        for (int d = 0; d < ND; ++d) F.cf[d] += ld;
    } else {
        F.blk -= blk_step; F.vec -= w; F.cbr -= w; F.rbc -= w; --F.bi;
#pragma unroll
        for (int d = 0; d < ND; ++d) F.cf[d] -= ld;
    }
}

template <int WP, int ND, int PF, bool TIGHT, bool QUADS>
__global__ void __launch_bounds__(WARPS * 32)
solve_chain_pf_kernel(const double *__restrict__ data, int64_t sys_stride, int nsys, Dims D,
                      const int *__restrict__ rowbycol, const int *__restrict__ colbyrow,
                      const double *__restrict__ bvec, double *xvec, VecLayout V, int *__restrict__ flags)
{
    constexpr int G = WP, GPW = 32 / G;
    const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
    const int grp = lane / G, r = lane % G;
    const int w = TIGHT ? WP : D.w, ld = TIGHT ? WP : D.ld, N = D.nblocks;
    const int64_t total = (int64_t)nsys * V.nrhs;
    int64_t gid = ((int64_t)blockIdx.x * WARPS + wic) * GPW + grp;
    const bool live = gid < total;
    if (!live) gid = total - 1;  // dead groups shadow the last vector and store nothing
    const int sys = (int)(gid / V.nrhs), rhs = (int)(gid % V.nrhs);
    const double *sd = data + (int64_t)sys * sys_stride;
    const double *b = bvec + (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride;
    double *x = xvec + (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride;
    const bool rowok = TIGHT || r < w;
    const int64_t blk_step = (int64_t)w * ld;
    const int nlines = (int)((blk_step * 8 + 127) / 128);  // <= 16 at blockw 16 with ld = 16; wider ld: the first G lines only
#ifndef BDILU_PF_FAR
#define BDILU_PF_FAR 0  // far L2 prefetch (blocks ahead of the ring): measured, and off -- its up to 32 line requests per block cost more than the DRAM latency they hide (cfg5 per GPU at 2 GPUs: 0.62 -> 0.41 ms without)
#endif
    constexpr int PD = BDILU_PF_FAR;

    int nan_first = 0x7fffffff;  // 1-based index of the first NaN in b (:500-501)
    int zero_last = -1;          // the zero U diagonal the reverse loop meets last (:533-534)
    extern __shared__ __align__(16) double ring[];
    constexpr int NITEM = Ring<WP, ND>::NITEM;
    constexpr int STAGE = NITEM * WARPS * 32;  // doubles per stage
    const double *myslot = ring + threadIdx.x;
    const unsigned myslot_s = (unsigned)__cvta_generic_to_shared(ring) + threadIdx.x * 8u;
    double prev[ND];             // forward: y of the ND blocks before; backward: x of the ND blocks after (mine: row r)
#pragma unroll
    for (int d = 0; d < ND; ++d) prev[d] = 0.0;

    // ---------------- forward: L y = P b - couplings (:502-517) ----------------
    Front<ND> F;
    F.blk = sd + r; F.vec = b + r; F.bi = 0;
    F.cbr = colbyrow + (int64_t)sys * N * w + r;
    F.rbc = rowbycol + (int64_t)sys * N * w + r;
#pragma unroll
    for (int d = 0; d < ND; ++d) F.cf[d] = sd + D.off_sub + D.diag_chunk(d, -1 - d) + r;  // sub(d, bi-1-d, r) at bi = 0
#pragma unroll
    for (int s = 0; s < PF; ++s) {
        issue_stage<WP, ND, true, TIGHT>(myslot_s + s * STAGE * 8u, F, w, ld, N, r, rowok, s < N);
        advance<ND, true>(F, w, blk_step, ld);
    }
    double *xo = x + r;  // x + bi*w + r of the block being processed
    for (int bi0 = 0; bi0 < N; bi0 += PF) {
#pragma unroll
        for (int s = 0; s < PF; ++s) {
            const int bi = bi0 + s;
            if (bi < N) {
                Stage<WP, ND> S;
                PF_T(t0);
                asm volatile("cp.async.wait_group %0;" ::"n"(PF - 1) : "memory");
                read_stage<WP, ND, true, TIGHT>(S, myslot + s * STAGE, r);
                PF_T(t1);
                double acc = group_bcast<G>(S.vec, S.rbc);
                if (rowok && isnan(acc)) nan_first = min(nan_first, bi * w + S.rbc + 1);
#pragma unroll
                for (int d = 0; d < ND; ++d) {
                    const double cf = group_bcast<G>(S.cf[d], S.cbr);
                    const double yv = group_bcast<G>(prev[d], S.cbr);
                    const double t = fma(-cf, yv, acc);
                    acc = (d < bi) ? t : acc;
                }
                if (D.nsat + bi - N >= 0) {
                    // satellite rows (periodic couplings, :512-514): the last nsat blocks also see y of the first blocks,
                    // read back from the x buffer (written by this group long ago; same SM, made visible by the barrier)
                    __syncwarp();
                    if (rowok)
                        for (int bci = 0; bci <= D.nsat + bi - N; ++bci)
                            acc = fma(-__ldg(sd + D.off_bot + D.sat_chunk(N + bci - bi - 1, bci) + S.cbr), x[bci * w + S.cbr], acc);
                }
                PF_T(t2);
                if constexpr (QUADS && TIGHT && WP >= 4) {
                    // four rows per shuffle round: the four accumulators and the six entries of the 4 x 4 triangle that
                    // couples them are broadcast together, every lane resolves the triangle redundantly (the same fmas, in
                    // the same order, as the owners would do one row at a time) and applies the four finished values.
                    // One dependent shuffle per four rows instead of four: ~25 instead of ~41 cycles per row, twice the
                    // instructions -- for batches that leave the SMs mostly idle (the chain's latency is the run time).
#pragma unroll
                    for (int k = 0; k < WP; k += 4) {
                        const double y0 = group_bcast<G>(acc, k);
                        const double a1 = group_bcast<G>(acc, k + 1), a2 = group_bcast<G>(acc, k + 2), a3 = group_bcast<G>(acc, k + 3);
                        const double m10 = group_bcast<G>(S.m[k], k + 1), m20 = group_bcast<G>(S.m[k], k + 2);
                        const double m30 = group_bcast<G>(S.m[k], k + 3), m21 = group_bcast<G>(S.m[k + 1], k + 2);
                        const double m31 = group_bcast<G>(S.m[k + 1], k + 3), m32 = group_bcast<G>(S.m[k + 2], k + 3);
                        const double y1 = fma(-m10, y0, a1);
                        const double y2 = fma(-m21, y1, fma(-m20, y0, a2));
                        const double y3 = fma(-m32, y2, fma(-m31, y1, fma(-m30, y0, a3)));
                        double t = fma(-S.m[k], y0, acc);
                        acc = (r > k) ? t : acc;
                        t = fma(-S.m[k + 1], y1, acc);
                        acc = (r > k + 1) ? t : acc;
                        t = fma(-S.m[k + 2], y2, acc);
                        acc = (r > k + 2) ? t : acc;
                        t = fma(-S.m[k + 3], y3, acc);
                        acc = (r > k + 3) ? t : acc;
                    }
                } else {
#pragma unroll
                    for (int k = 0; k < WP; ++k) {
                        if (TIGHT || k < w) {
                            const double yk = group_bcast<G>(acc, k);
                            const double t = fma(-S.m[k], yk, acc);
                            acc = (r > k) ? t : acc;
                        }
                    }
                }
                PF_T(t3);
                if (live && rowok) *xo = acc;
                xo += w;
#pragma unroll
                for (int d = ND - 1; d > 0; --d) prev[d] = prev[d - 1];
                prev[0] = acc;
                issue_stage<WP, ND, true, TIGHT>(myslot_s + s * STAGE * 8u, F, w, ld, N, r, rowok, bi + PF < N);
                if (PD > 0) prefetch_far<ND, true>(F, w, blk_step, ld, N, r, nlines, PD);
                advance<ND, true>(F, w, blk_step, ld);
                PF_T(t4);
                PF_ADD(0, t0, t1); PF_ADD(1, t1, t2); PF_ADD(2, t2, t3); PF_ADD(3, t3, t4); PF_ADD(4, 0, 1);
            }
        }
    }

    // ---------------- backward: U x = y - couplings (:518-536) ----------------
    // y (written with ordinary stores above, by this thread) comes back through cp.async: all groups of the forward
    // sweep are drained and the stores are ordered before the copies issued below
    asm volatile("cp.async.wait_all;" ::: "memory");
    __threadfence();
#pragma unroll
    for (int d = 0; d < ND; ++d) prev[d] = 0.0;
    F.blk = sd + D.block(N - 1) + r; F.vec = x + (int64_t)(N - 1) * w + r; F.bi = N - 1;
    F.cbr = colbyrow + (int64_t)sys * N * w + (int64_t)(N - 1) * w + r;
    F.rbc = F.cbr;  // unused in this sweep
#pragma unroll
    for (int d = 0; d < ND; ++d) F.cf[d] = sd + D.off_sup + D.diag_chunk(d, N - 1) + r;  // sup(d, bi, r) at bi = N-1
#pragma unroll
    for (int s = 0; s < PF; ++s) {
        issue_stage<WP, ND, false, TIGHT>(myslot_s + s * STAGE * 8u, F, w, ld, N, r, rowok, s < N);
        advance<ND, false>(F, w, blk_step, ld);
    }
    xo = x + (int64_t)(N - 1) * w + r;
    for (int i0 = 0; i0 < N; i0 += PF) {
#pragma unroll
        for (int s = 0; s < PF; ++s) {
            const int i = i0 + s;  // sweep order
            if (i < N) {
                const int bi = N - 1 - i;
                Stage<WP, ND> S;
                asm volatile("cp.async.wait_group %0;" ::"n"(PF - 1) : "memory");
                read_stage<WP, ND, false, TIGHT>(S, myslot + s * STAGE, r);
                // x[r] = (y[r] - sum U[r][k] x[k]) / U[r][r] (:524-532) with the division off the chain: my row, its
                // coupling coefficients and y[r] are scaled by 1/U[r][r] while the predecessors still work (as the
                // blockw-32 kernel does), so a row costs one shuffle and one FMA.  Not the reference's rounding sequence
                // (one rounding more per term), well inside the 1e-12 bar; flags are exact.
                const double dg = S.dg;
                if (rowok && dg == 0.0) zero_last = max(zero_last, bi * w + r);
                const double rinv = rcp_fast_ok(dg) ? rcp_fast(dg) : 1.0 / dg;
                double acc = S.vec * rinv;
#pragma unroll
                for (int d = 0; d < ND; ++d) {
                    const double cf = group_bcast<G>(S.cf[d], S.cbr) * rinv;
                    const double xv = group_bcast<G>(prev[d], S.cbr);
                    const double t = fma(-cf, xv, acc);
                    acc = (d < i) ? t : acc;
                }
                if (bi < D.nsat) {
                    // satellite columns (:528-530): the first nsat blocks also see x of the last blocks
                    __syncwarp();
                    if (rowok)
                        for (int sati = D.nsat; sati > bi; --sati)
                            acc = fma(-(__ldg(sd + D.off_top + D.sat_chunk(sati - 1, bi) + S.cbr) * rinv),
                                      x[(N - sati + bi) * w + S.cbr], acc);
                }
                if constexpr (QUADS && TIGHT && WP >= 4) {
#pragma unroll
                    for (int kq = 0; kq < WP; kq += 4) {
                        const int k = WP - 1 - kq;  // rows k, k-1, k-2, k-3
                        const double s0 = S.m[k] * rinv, s1 = S.m[k - 1] * rinv, s2 = S.m[k - 2] * rinv, s3 = S.m[k - 3] * rinv;
                        const double x0 = group_bcast<G>(acc, k);
                        const double a1 = group_bcast<G>(acc, k - 1), a2 = group_bcast<G>(acc, k - 2), a3 = group_bcast<G>(acc, k - 3);
                        const double c10 = group_bcast<G>(s0, k - 1), c20 = group_bcast<G>(s0, k - 2), c30 = group_bcast<G>(s0, k - 3);
                        const double c21 = group_bcast<G>(s1, k - 2), c31 = group_bcast<G>(s1, k - 3), c32 = group_bcast<G>(s2, k - 3);
                        const double x1 = fma(-c10, x0, a1);
                        const double x2 = fma(-c21, x1, fma(-c20, x0, a2));
                        const double x3 = fma(-c32, x2, fma(-c31, x1, fma(-c30, x0, a3)));
                        double t = fma(-s0, x0, acc);
                        acc = (r < k) ? t : acc;
                        t = fma(-s1, x1, acc);
                        acc = (r < k - 1) ? t : acc;
                        t = fma(-s2, x2, acc);
                        acc = (r < k - 2) ? t : acc;
                        t = fma(-s3, x3, acc);
                        acc = (r < k - 3) ? t : acc;
                    }
                } else {
#pragma unroll
                    for (int kr = 0; kr < WP; ++kr) {
                        const int k = WP - 1 - kr;
                        if (TIGHT || k < w) {
                            const double mk = S.m[k] * rinv;
                            const double xk = group_bcast<G>(acc, k);
                            const double t = fma(-mk, xk, acc);
                            acc = (r < k) ? t : acc;
                        }
                    }
                }
                if (live && rowok) *xo = acc;
                xo -= w;
#pragma unroll
                for (int d = ND - 1; d > 0; --d) prev[d] = prev[d - 1];
                prev[0] = acc;
                issue_stage<WP, ND, false, TIGHT>(myslot_s + s * STAGE * 8u, F, w, ld, N, r, rowok, i + PF < N);
                if (PD > 0) prefetch_far<ND, false>(F, w, blk_step, ld, N, r, nlines, PD);
                advance<ND, false>(F, w, blk_step, ld);
            }
        }
    }

#pragma unroll
    for (int off = G / 2; off > 0; off >>= 1) {
        nan_first = min(nan_first, __shfl_xor_sync(FULL, nan_first, off, G));
        zero_last = max(zero_last, __shfl_xor_sync(FULL, zero_last, off, G));
    }
    if (live && r == 0 && flags)
        flags[gid] = (nan_first != 0x7fffffff) ? nan_first : (zero_last >= 0 ? N * w + zero_last : 0);
}

template <int WP, int ND, int PF, bool TIGHT, bool QUADS>
static cudaError_t launch_one(unsigned grid, const double *data, int64_t sys_stride, int nsys, const Dims &D, const int *rowbycol,
                              const int *colbyrow, const double *b, double *x, const VecLayout &V, int *flags, cudaStream_t st)
{
    const size_t smem = (size_t)PF * Ring<WP, ND>::NITEM * WARPS * 32 * sizeof(double);
    cudaError_t e = cudaFuncSetAttribute(solve_chain_pf_kernel<WP, ND, PF, TIGHT, QUADS>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);  // per (function, device)
    if (e != cudaSuccess) return e;
    solve_chain_pf_kernel<WP, ND, PF, TIGHT, QUADS><<<grid, WARPS * 32, smem, st>>>(data, sys_stride, nsys, D, rowbycol, colbyrow,
                                                                                   b, x, V, flags);
    return cudaGetLastError();
}

template <int WP, int PF>
static cudaError_t launch_w(const double *data, int64_t sys_stride, int nsys, const Dims &D, const int *rowbycol,
                            const int *colbyrow, const double *b, double *x, const VecLayout &V, int *flags, cudaStream_t st)
{
    const int64_t total = (int64_t)nsys * V.nrhs;
    const int64_t per_cta = (int64_t)WARPS * (32 / WP);
    const unsigned grid = (unsigned)((total + per_cta - 1) / per_cta);
    // tight blocks are copied as 16-byte chunks: the blocks of every system must then sit on 16-byte boundaries
    const bool tight = D.w == WP && D.ld == WP && (reinterpret_cast<uintptr_t>(data) & 15) == 0 && (sys_stride & 1) == 0;
    // four rows per shuffle round while the chains' latency is the run time (few warps per SM); one row per round once
    // the batch fills the SMs' load/store pipe (the quad form issues 20 shuffles per 4 rows instead of 8)
    const int64_t warps = (total * WP + 31) / 32;
    // (test hooks 11 / 12 of bdilu_debug_force_chain_solve force four rows / one row per round)
    const bool quads = tight && WP >= 4 &&
                       (g_solve_path == 11 || (g_solve_path != 12 && warps <= (int64_t)PF_QUAD_WARPS_PER_SM * sm_count()));
#define PF_GO(ND)                                                                                                              \
    return !tight ? launch_one<WP, ND, PF, false, false>(grid, data, sys_stride, nsys, D, rowbycol, colbyrow, b, x, V, flags, st) \
           : quads ? launch_one<WP, ND, PF, true, true>(grid, data, sys_stride, nsys, D, rowbycol, colbyrow, b, x, V, flags, st)  \
                   : launch_one<WP, ND, PF, true, false>(grid, data, sys_stride, nsys, D, rowbycol, colbyrow, b, x, V, flags, st)
    switch (D.ndiag) {
    case 1: PF_GO(1);
    case 2: PF_GO(2);
    case 3: PF_GO(3);
    default: PF_GO(4);
    }
#undef PF_GO
}

}  // namespace pf

#ifdef BDILU_PF_PROF
extern "C" __attribute__((visibility("default"))) int bdilu_debug_pfprof(unsigned long long *out8, int reset)
{
    if (out8) cudaMemcpyFromSymbol(out8, pf::g_pfprof, sizeof(unsigned long long) * 8);
    if (reset) {
        unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        cudaMemcpyToSymbol(pf::g_pfprof, z, sizeof z);
    }
    return 0;
}
#endif

// blockw <= 32, 1 <= ndiag <= 4, any nsat.  Returns false when the shape does not fit.
bool try_launch_solve_pf(const double *data, int64_t sys_stride, int nsys, const Dims &D, const int *rowbycol,
                         const int *colbyrow, const double *b, double *x, const VecLayout &V, int *flags, cudaStream_t st,
                         cudaError_t *err)
{
    if (D.ndiag < 1 || D.ndiag > 4 || D.w > 32 || D.w < 2) return false;
    const int w = D.w;
#ifndef BDILU_PF_DEPTH8
#define BDILU_PF_DEPTH8 6
#endif
    if (w <= 2) *err = pf::launch_w<2, 8>(data, sys_stride, nsys, D, rowbycol, colbyrow, b, x, V, flags, st);
    else if (w <= 4) *err = pf::launch_w<4, 8>(data, sys_stride, nsys, D, rowbycol, colbyrow, b, x, V, flags, st);
    else if (w <= 8) *err = pf::launch_w<8, BDILU_PF_DEPTH8>(data, sys_stride, nsys, D, rowbycol, colbyrow, b, x, V, flags, st);
    else if (w <= 16) *err = pf::launch_w<16, 4>(data, sys_stride, nsys, D, rowbycol, colbyrow, b, x, V, flags, st);
    else *err = pf::launch_w<32, 3>(data, sys_stride, nsys, D, rowbycol, colbyrow, b, x, V, flags, st);
    count_launch(1);
    return true;
}

}  // namespace bdilu
