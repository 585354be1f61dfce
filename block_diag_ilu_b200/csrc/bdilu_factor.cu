// Kernel 1: batched per-block LU with partial pivoting + sub/bot scaling + permutation vectors.
//
// Replaces ILU_inplace<double>::ILU_inplace (block_diag_ilu.hpp:430-485) and the dgetrf_ it calls
// (anyode_blas_lapack.hpp:55-57).  One G-lane group (G = blockw rounded up to a power of two, 32/G
// blocks per warp) owns one diagonal block; lane r keeps ROW r of the block in registers.
//
// Pivoting is implicit: rows never move between lanes.  Each lane tracks `pos`, the position its row
// currently occupies in LAPACK's swapped ordering, so that
//   * the pivot of step k is the candidate of maximum modulus with the LOWEST current position
//     (idamax's "first maximum" rule, which is what makes ipiv bit-exact),
//   * ipiv[k] = (position of the chosen row) + 1, and the row that sat at position k takes its place,
//   * at the end lane r stores its row at row `pos` of the column-major block: the stored L\U is
//     LAPACK's in-place layout, rowbycol[pos] = r, colbyrow[r] = pos (block_diag_ilu.hpp:34-52).
// The arithmetic on each element is the right-looking elimination of dgetf2 / anyode_blasless.hpp:
// l = a/piv by true division, a -= l*u.
#include "bdilu_common.cuh"
#include "bdilu_kernels.h"

namespace bdilu {

constexpr int FACTOR_WARPS = 4;
#ifndef BDILU_FACTOR_SHFL_ROW
#define BDILU_FACTOR_SHFL_ROW 0
#endif
#ifndef BDILU_FACTOR_PAIRED
#define BDILU_FACTOR_PAIRED 1
#endif
constexpr bool PAIRED = BDILU_FACTOR_PAIRED != 0;  // blockw 32: two elimination steps per pivot-row publication
#ifndef BDILU_FACTOR_OPTIMISTIC
#define BDILU_FACTOR_OPTIMISTIC 1
#endif
constexpr bool OPTIMISTIC = BDILU_FACTOR_OPTIMISTIC != 0;  // blockw 32: diagonal pivots assumed and checked first (eliminate_diag32)
#ifndef BDILU_FACTOR_GRID2D
#define BDILU_FACTOR_GRID2D 1
#endif
#ifndef BDILU_FACTOR_GRID_PAIRED
#define BDILU_FACTOR_GRID_PAIRED 0
#endif
constexpr bool GRID_PAIRED = BDILU_FACTOR_GRID_PAIRED != 0;  // ... with two steps per publication (eliminate_diag32_grid2): measured
                                                             // 2.55 ms against 2.41 ms unpaired on cfg3, so off
constexpr bool GRID2D = BDILU_FACTOR_GRID2D != 0;  // blockw == ld == 32: the optimistic path on a 16 x 2 lane grid (eliminate_diag32_grid)
constexpr bool SHFL_ROW = BDILU_FACTOR_SHFL_ROW != 0;  // blockw 32: pivot row by shuffle (1: 3.45 ms on cfg3) or mailbox (0: 3.2 ms)

// argmax of v over the group with ties broken towards the lowest `pos`; returns the winning lane
// (group-local index).  Inactive lanes pass v = -1.
template <int G>
__device__ __forceinline__ int group_argmax(double v, int pos, int r)
{
    if (G == 32) {
        // |x| >= 0 orders like its bit pattern: reduce the high word, then the low word among ties.
        const int hi = __double2hiint(v);
        const unsigned lo = (unsigned)__double2loint(v);
        const int mhi = __reduce_max_sync(FULL, hi);
        const bool c1 = (hi == mhi);
        const unsigned mlo = __reduce_max_sync(FULL, c1 ? lo : 0u);
        const bool c2 = c1 && (lo == mlo);
        unsigned bal = __ballot_sync(FULL, c2);
        if (__popc(bal) > 1) {
            const unsigned mp = __reduce_min_sync(FULL, c2 ? (unsigned)pos : 0xffffu);
            bal = __ballot_sync(FULL, c2 && (unsigned)pos == mp);
        }
        return __ffs(bal) - 1;
    } else {
        int win = r;
#pragma unroll
        for (int off = G / 2; off > 0; off >>= 1) {
            const double v2 = group_bcast<G>(v, r ^ off);
            const int p2 = __shfl_sync(FULL, pos, r ^ off, G);
            const int w2 = __shfl_sync(FULL, win, r ^ off, G);
            if (v2 > v || (v2 == v && p2 < pos)) { v = v2; pos = p2; win = w2; }
        }
        return win;
    }
}

// ---- lean path -------------------------------------------------------------------------------------
// Straight-line elimination step: no branch inside the 32-step unrolled loop, so ptxas can overlap the
// pivot search of step k+1 (it only needs column k+1) with the trailing update of step k.
//   * pivot search on integer keys: |a| as (hi & 0x7fffffff, lo); three warp reductions give
//     max hi -> max lo among those -> min position among those (idamax: first maximum);
//   * every candidate lane computes 1/a[k] speculatively while the search runs; the winner publishes
//     it next to its row, and l = a/piv is formed as q = a*rinv; q += (a - q*piv)*rinv, which is the
//     correctly rounded quotient when rinv is the correctly rounded reciprocal (Markstein) -- the same
//     value the reference's true division gives (tests/test_block_diag.cpp:139-142 check exact quotients);
//   * rows that are already pivot rows, and steps with an exactly zero pivot (dgetf2 leaves that column
//     unscaled and records info), use l = 0 instead of a branch.
// Pivots outside the range where the branch-free reciprocal is exact (|x| < 2^-990 or > 2^990, inf, NaN)
// set `weird`; such a warp redoes its block on the slow true-division path below before storing anything.

template <int G>
__device__ __forceinline__ int group_max_i(int v)
{
    if (G == 32) return __reduce_max_sync(FULL, v);
#pragma unroll
    for (int off = G / 2; off > 0; off >>= 1) v = max(v, __shfl_xor_sync(FULL, v, off, G));
    return v;
}
template <int G>
__device__ __forceinline__ unsigned group_max_u(unsigned v)
{
    if (G == 32) return __reduce_max_sync(FULL, v);
#pragma unroll
    for (int off = G / 2; off > 0; off >>= 1) v = max(v, __shfl_xor_sync(FULL, v, off, G));
    return v;
}
template <int G>
__device__ __forceinline__ unsigned group_min_u(unsigned v)
{
    if (G == 32) return __reduce_min_sync(FULL, v);
#pragma unroll
    for (int off = G / 2; off > 0; off >>= 1) v = min(v, __shfl_xor_sync(FULL, v, off, G));
    return v;
}

// 16-byte shared-memory store under a predicate (never a branch: see the paired loop of factor_kernel)
__device__ __forceinline__ void sts128_if(bool p, unsigned saddr, double x, double y)
{
    asm volatile("{ .reg .pred p; setp.ne.u32 p, %3, 0; @p st.shared.v2.f64 [%0], {%1, %2}; }" ::"r"(saddr), "d"(x),
                 "d"(y), "r"((unsigned)p)
                 : "memory");
}

__device__ __forceinline__ void sts64_if(bool p, unsigned saddr, double x)
{
    asm volatile("{ .reg .pred p; setp.ne.u32 p, %2, 0; @p st.shared.f64 [%0], %1; }" ::"r"(saddr), "d"(x), "r"((unsigned)p)
                 : "memory");
}

// Slow path (true divisions, branches): the algorithm exactly as documented at the top of this file.
// Also the only path for 32 < blockw (see factor_wide_kernel).  a[] holds the unfactored row on entry.
struct RowState {
    int pos, ipiv, info;
    double diag;
    double rinv;  // correctly rounded 1/diag when `fast` (lean path), unused otherwise
    bool fast;
};

// Where the unfactored numbers of one system come from.  In place (`scaled` false): the destination itself.  Fused
// assembly of M = scale*J + diag_add*I (BlockDiagMatrix::scale_diag_add, block_diag_ilu.hpp:276-287, the step an
// integrator runs before every factorisation): J is read from `sys`, scaled on the fly -- product and sum rounded
// separately, like the reference's two statements -- and only the factors are ever written.
struct SrcView {
    const double *sys;
    double scale, diag_add;
    bool scaled;
    __device__ __forceinline__ double at(int64_t off) const
    {
        const double v = sys[off];
        return scaled ? __dmul_rn(scale, v) : v;
    }
    __device__ __forceinline__ double block_at(int64_t off, bool on_diag) const
    {
        double v = ld_stream(sys + off);
        if (scaled) {
            v = __dmul_rn(scale, v);
            if (on_diag) v = __dadd_rn(v, diag_add);
        }
        return v;
    }
};

// Stores what one lane owns after the elimination: its row of L\U at row `pos`, the permutation
// vectors, the scaled sub/bot entries of column r, and the block's info.
template <int WP, bool ROWS = true>
__device__ __forceinline__ void store_block(const double (&a)[WP], const RowState &S, const SrcView &V, double *sysdata,
                                            double *blk, const Dims &D, int w, int ld, int sys, int bi, int r,
                                            int *__restrict__ ipiv, int *__restrict__ rowbycol,
                                            int *__restrict__ colbyrow, int *__restrict__ info)
{
    // sub(di,bi,c) /= U[c][c]  (block_diag_ilu.hpp:471-475): the loads go out first so that their latency hides
    // behind the column stores; the quotient is s*rinv corrected by one FMA pair (Markstein: correctly rounded when
    // rinv is the correctly rounded reciprocal), a true division only for a zero / out-of-range diagonal.
    constexpr int PF = 2;  // sub rows prefetched into registers
    const int nsub = min(D.ndiag, D.nblocks - 1 - bi);
    double sv[PF];
#pragma unroll
    for (int di = 0; di < PF; ++di)
        sv[di] = (di < nsub) ? V.at(D.off_sub + D.diag_chunk(di, bi) + r) : 0.0;
    if (ROWS) {
#pragma unroll
        for (int c = 0; c < WP; ++c)
            if (c < w) blk[(int64_t)c * ld + S.pos] = a[c];
    }
    const int64_t ib = ((int64_t)sys * D.nblocks + bi) * w;
    ipiv[ib + r] = S.ipiv;
    rowbycol[ib + S.pos] = r;
    colbyrow[ib + r] = S.pos;
    const bool fastq = S.fast && S.diag != 0.0;
    auto quot = [&](double sval) {
        if (!fastq) return sval / S.diag;
        double qv = sval * S.rinv;
        return fma(fma(-qv, S.diag, sval), S.rinv, qv);
    };
#pragma unroll
    for (int di = 0; di < PF; ++di)
        if (di < nsub) sysdata[D.off_sub + D.diag_chunk(di, bi) + r] = quot(sv[di]);
    for (int di = PF; di < nsub; ++di) {
        const int64_t off = D.off_sub + D.diag_chunk(di, bi) + r;
        sysdata[off] = quot(V.at(off));
    }
    // bot(sati,bi,c) /= U[c][c]  (:476-479)
    if (bi < D.nsat)
        for (int sati = bi; sati < D.nsat; ++sati) {
            const int64_t off = D.off_bot + D.sat_chunk(sati, bi) + r;
            sysdata[off] = V.at(off) / S.diag;
        }
    if (V.scaled) {
        // sup and top are not touched by the factorisation; the fused assembly still owes the destination their
        // scaled copies (same index sets as sub and bot)
        for (int di = 0; di < nsub; ++di) {
            const int64_t off = D.off_sup + D.diag_chunk(di, bi) + r;
            sysdata[off] = V.at(off);
        }
        if (bi < D.nsat)
            for (int sati = bi; sati < D.nsat; ++sati) {
                const int64_t off = D.off_top + D.sat_chunk(sati, bi) + r;
                sysdata[off] = V.at(off);
            }
    }
    if (S.info != 0 && r == 0) {
        // keep the failing block with the lowest index (serial loop order, :469-470)
        const int key = (bi << 7) | S.info;
        int old = atomicCAS(&info[sys], 0, key);
        while (old != 0 && (old >> 7) > bi) {
            const int assumed = old;
            old = atomicCAS(&info[sys], assumed, key);
            if (old == assumed) break;
        }
    }
}

// Slow path (true divisions, branches, rows in local memory): the algorithm exactly as documented at the
// top of this file, from the pristine block in global memory.  Taken only by warps that met a pivot
// outside the fast reciprocal's exact range.
template <int WP>
__device__ __noinline__ void factor_block_slow(SrcView V, double *sysdata, double *blk, Dims D, int w, int ld, int sys,
                                               int bi, int r, bool live, double *ubuf, int *__restrict__ ipiv,
                                               int *__restrict__ rowbycol, int *__restrict__ colbyrow,
                                               int *__restrict__ info)
{
    constexpr int G = WP;
    const bool rowok = r < w;
    double a[WP];
    for (int c = 0; c < WP; ++c)
        a[c] = (rowok && c < w) ? V.block_at(D.block(bi) + (int64_t)c * ld + r, c == r) : ((c == r) ? 1.0 : 0.0);
    RowState S;
    S.pos = r; S.ipiv = r + 1; S.info = 0; S.diag = 1.0; S.rinv = 1.0; S.fast = false;
    bool active = rowok;
    for (int k = 0; k < w; ++k) {
        const double v = active ? fabs(a[k]) : -1.0;
        const int p = group_argmax<G>(v, S.pos, r);
        const int q = __shfl_sync(FULL, S.pos, p, G);
        if (r == k) S.ipiv = q + 1;
        if (S.pos == k) S.pos = q;
        if (r == p) { S.pos = k; active = false; }
        __syncwarp();
        if (r == p)
            for (int c = k; c < WP; ++c) ubuf[c] = a[c];
        __syncwarp();
        const double piv = ubuf[k];
        if (r == k) S.diag = piv;
        if (piv == 0.0) {
            if (S.info == 0) S.info = k + 1;
        } else if (active) {
            const double l = a[k] / piv;
            a[k] = l;
            for (int c = k + 1; c < WP; ++c) a[c] = fma(-l, ubuf[c], a[c]);
        }
    }
    if (live && rowok) {
        for (int c = 0; c < w; ++c) blk[(int64_t)c * ld + S.pos] = a[c];
        const int64_t ib = ((int64_t)sys * D.nblocks + bi) * w;
        ipiv[ib + r] = S.ipiv;
        rowbycol[ib + S.pos] = r;
        colbyrow[ib + r] = S.pos;
        for (int di = 0; di < D.ndiag && bi + di < D.nblocks - 1; ++di) {
            const int64_t off = D.off_sub + D.diag_chunk(di, bi) + r;
            sysdata[off] = V.at(off) / S.diag;
            if (V.scaled) sysdata[off - D.off_sub + D.off_sup] = V.at(off - D.off_sub + D.off_sup);
        }
        if (bi < D.nsat)
            for (int sati = bi; sati < D.nsat; ++sati) {
                const int64_t off = D.off_bot + D.sat_chunk(sati, bi) + r;
                sysdata[off] = V.at(off) / S.diag;
                if (V.scaled) sysdata[off - D.off_bot + D.off_top] = V.at(off - D.off_bot + D.off_top);
            }
        if (S.info != 0 && r == 0) {
            const int key = (bi << 7) | S.info;
            int old = atomicCAS(&info[sys], 0, key);
            while (old != 0 && (old >> 7) > bi) {
                const int assumed = old;
                old = atomicCAS(&info[sys], assumed, key);
                if (old == assumed) break;
            }
        }
    }
}

// ---- optimistic elimination (blockw 32) ------------------------------------------------------------------------------
// The matrices this library exists for (I - gamma*J of a diffusion-reaction Jacobian, block_diag_ilu.hpp:288-290) are
// diagonally dominant: partial pivoting then picks the diagonal at every step and ipiv is the identity.  This loop
// ASSUMES that -- the pivot of step k is the row at position k, i.e. lane k, known at compile time -- and only CHECKS
// it: idamax takes the first maximum, so the assumption holds exactly when no row below has a strictly larger modulus
// in column k.  No reductions, no position bookkeeping, static pivot lanes; the arithmetic on every element is that of
// the general loop (same publication in pairs, same Markstein quotients), so a block that passes the check comes out
// bit-identical.  A block that fails it (a larger candidate, a NaN, a zero or out-of-range pivot) is reloaded and
// redone by the general loop of factor_kernel, which then decides everything itself: the results never depend on
// which path ran.  The check is voted every 8 steps so that blocks which do pivot leave early.
// Returns true (warp-uniform) when the block is done: a[] holds row r of L\U, dgs[k] = (U[k][k], 1/U[k][k]).
__device__ __forceinline__ bool eliminate_diag32(double (&a)[32], double (*urow)[34], double2 *dgs, int r)
{
    constexpr int WP = 32;
    double2 *ub0 = reinterpret_cast<double2 *>(urow[0]);
    double2 *ub1 = reinterpret_cast<double2 *>(urow[1]);
    double *nx = &urow[0][WP];
    const unsigned dgs_s = (unsigned)__cvta_generic_to_shared(dgs), nx_s = (unsigned)__cvta_generic_to_shared(nx);
    bool viol = false;
#pragma unroll
    for (int k = 0; k < WP; k += 2) {
        // ---- step k: lane k is the pivot row; the three numbers column k+1 needs ----
        const double rcp0 = rcp_fast(a[k]);
        sts128_if(r == k, dgs_s + 16u * k, a[k], rcp0);
        sts64_if(r == k, nx_s, a[k + 1]);
        __syncwarp();
        const double2 d0 = dgs[k];
        const double u0n = *nx;
        const double piv0 = d0.x, rinv0 = d0.y;
        const bool act0 = r > k;
        viol |= act0 & !(fabs(a[k]) <= fabs(piv0));  // a larger candidate below, or a NaN on either side
        double l0 = a[k] * rinv0;
        l0 = fma(fma(-l0, piv0, a[k]), rinv0, l0);
        a[k] = act0 ? l0 : a[k];
        l0 = act0 ? l0 : 0.0;
        a[k + 1] = fma(-l0, u0n, a[k + 1]);

        // ---- step k+1: lane k+1 is the pivot row ----
        const double rcp1 = rcp_fast(a[k + 1]);
        sts128_if(r == k + 1, dgs_s + 16u * (k + 1), a[k + 1], rcp1);
        {
            // lanes k and k+1 in one predicated store sequence (see the general loop); buffer by arithmetic, no select
            const unsigned rk = (unsigned)(r - k);
            if (rk < 2u) {
                double2 *dst = ub0 + rk * (WP / 2 + 1);  // urow[rk]
#pragma unroll
                for (int m = k / 2; m < WP / 2; ++m) dst[m] = make_double2(a[2 * m], a[2 * m + 1]);
            }
        }
        __syncwarp();
        const double2 hd = ub1[k / 2];
        const double l0q = hd.x, piv1 = hd.y, rinv1 = dgs[k + 1].y;
        const bool act1 = r > k + 1;
        viol |= act1 & !(fabs(a[k + 1]) <= fabs(piv1));
        double l1 = a[k + 1] * rinv1;
        l1 = fma(fma(-l1, piv1, a[k + 1]), rinv1, l1);
        a[k + 1] = act1 ? l1 : a[k + 1];
        l1 = act1 ? l1 : 0.0;
#pragma unroll
        for (int m = k / 2 + 1; m < WP / 2; ++m) {
            const double2 t0 = ub0[m], t1 = ub1[m];
            const double u1x = fma(-l0q, t0.x, t1.x), u1y = fma(-l0q, t0.y, t1.y);
            a[2 * m] = fma(-l1, u1x, fma(-l0, t0.x, a[2 * m]));
            a[2 * m + 1] = fma(-l1, u1y, fma(-l0, t0.y, a[2 * m + 1]));
        }
        if ((k & 7) == 6 && k + 2 < WP && __any_sync(FULL, viol)) return false;
    }
    __syncwarp();
    const double dg = dgs[r].x;  // a zero or out-of-range pivot: the general loop (info, true divisions) takes over
    viol |= (dg == 0.0) || !rcp_fast_ok(dg);
    return !__any_sync(FULL, viol);
}

// The optimistic elimination for narrower blocks (G = 2..16 lanes per block, several blocks per warp): lane k of each
// group is the pivot row of step k, one publication per step through the group's mailbox.  Same checks, same
// arithmetic as the general loop; the vote is over the whole warp, so one group that has to pivot sends its
// neighbours through the general loop as well (results do not depend on the path).
template <int WP>
__device__ __forceinline__ bool eliminate_diag_small(double (&a)[WP], double (*urow)[WP + 2], double2 *dgs, int r)
{
    bool viol = false;
#pragma unroll
    for (int k = 0; k < WP; ++k) {
        const double rc = rcp_fast(a[k]);
        double2 *ub = reinterpret_cast<double2 *>(urow[k & 1]);
        if (r == k) {
#pragma unroll
            for (int m = (k + 1) / 2; m < WP / 2; ++m) ub[m] = make_double2(a[2 * m], a[2 * m + 1]);
            dgs[k] = make_double2(a[k], rc);
        }
        __syncwarp();
        const double2 d = dgs[k];
        const double piv = d.x, rinv = d.y;
        const bool act = r > k;
        viol |= act & !(fabs(a[k]) <= fabs(piv));
        double l = a[k] * rinv;
        l = fma(fma(-l, piv, a[k]), rinv, l);
        a[k] = act ? l : a[k];
        l = act ? l : 0.0;
#pragma unroll
        for (int m = (k + 1) / 2; m < WP / 2; ++m) {
            const double2 u = ub[m];
            if (2 * m > k) a[2 * m] = fma(-l, u.x, a[2 * m]);
            a[2 * m + 1] = fma(-l, u.y, a[2 * m + 1]);
        }
        if ((k == 1 || k == WP / 2 - 1) && k + 1 < WP && __any_sync(FULL, viol)) return false;  // pivoting blocks leave early
    }
    __syncwarp();
    const double dg = dgs[r].x;
    viol |= (dg == 0.0) || !rcp_fast_ok(dg);
    return !__any_sync(FULL, viol);
}

// The same optimistic elimination on a 2-D lane grid (blockw == ld == 32): lane = i + 16 j holds rows {i, i+16} and
// the columns of parity j, t[a][b] = A[i + 16a][j + 2b].  With one lane per row every lane reads every element of every
// pivot row (the shared-memory pipe bounds that kernel: l1tex lsu wavefronts 69 %, profiles/r02_ncu_full_summary.md);
// here a lane reads only its half of the pivot row and uses each element for two rows, and the two pivot-row lanes
// publish 8 + 8 doubles instead of one lane publishing 32.  The multipliers of column k are formed by the 16 lanes that
// hold it and handed to their row partners by shuffle.  The 2-D layout sank an earlier attempt on the GENERAL
// elimination (two-row pivot bookkeeping, git show 1850672:scratch/factor32x2_experiment.patch); with static diagonal pivots there is no
// bookkeeping left.  Element for element the operations are those of eliminate_diag32 without its pairing (each update
// a -= l*u is one fma with the finished pivot-row value), hence those of the general loop: bit-identical results.
__device__ __forceinline__ bool eliminate_diag32_grid(double (&t)[2][16], double *mb, double2 *dgs, int lane)
{
    const int i = lane & 15, j = lane >> 4;
    bool viol = false;
#pragma unroll
    for (int k = 0; k < 32; ++k) {
        const int pi = k & 15, ak = k >> 4, pj = k & 1, bk = k >> 1;
        const int b0 = (pj == 0) ? bk : bk + 1;  // first column index (of either half) a reader may need
        const double x = t[ak][bk];              // the pivot, in lane (pi, pj)
        const double rc = rcp_fast(x);
        double *buf = mb + (k & 1) * 32;         // double-buffered by step parity: one barrier per step is enough
        if (i == pi) {
            double2 *dst = reinterpret_cast<double2 *>(buf + j * 16);
#pragma unroll
            for (int m = b0 / 2; m < 8; ++m) dst[m] = make_double2(t[ak][2 * m], t[ak][2 * m + 1]);
            if (j == pj) dgs[k] = make_double2(x, rc);
        }
        __syncwarp();
        const double2 d = dgs[k];
        const double piv = d.x, rinv = d.y;
        const bool own = (j == pj);
        double l[2] = {0.0, 0.0};
#pragma unroll
        for (int a = (k >= 15 ? 1 : 0); a < 2; ++a) {
            const double xa = t[a][bk];
            const bool act = (i + 16 * a) > k;
            viol |= own & act & !(fabs(xa) <= fabs(piv));  // a larger candidate below, or a NaN on either side
            double q = xa * rinv;
            q = fma(fma(-q, piv, xa), rinv, q);
            t[a][bk] = (own & act) ? q : xa;
            q = group_bcast<32>(q, i | (pj << 4));  // from the lane of my row pair that holds column k
            l[a] = act ? q : 0.0;
        }
        const double2 *src = reinterpret_cast<const double2 *>(buf + j * 16);
#pragma unroll
        for (int m = b0 / 2; m < 8; ++m) {
            const double2 u = src[m];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int b = 2 * m + e;
                if (b >= b0) {
                    double ub = e ? u.y : u.x;
                    if (pj == 0 && b == bk) ub = (j == 1) ? ub : 0.0;  // column k itself in the even half: multipliers
#pragma unroll
                    for (int a = (k >= 15 ? 1 : 0); a < 2; ++a) t[a][b] = fma(-l[a], ub, t[a][b]);
                }
            }
        }
        // blocks that pivot leave early: almost always at the first vote (a matrix that is not dominant shows it at once)
        if ((k == 1 || (k & 7) == 7) && k + 1 < 32 && __any_sync(FULL, viol)) return false;
    }
    __syncwarp();
    const double dg = dgs[lane].x;  // a zero or out-of-range pivot: the general loop (info, true divisions) takes over
    viol |= (dg == 0.0) || !rcp_fast_ok(dg);
    return !__any_sync(FULL, viol);
}

// The grid elimination with TWO steps per publication (as the one-lane-per-row loop does, and for the same reason: a
// shared-memory store costs the register -> MIO transfer of all 32 lanes whatever the number of active ones, and that
// queue is the busiest unit of this kernel: sm__mio_pq_write 64 % with one publication per step).  Rows k and k+1
// (k even) go out in ONE store sequence with four lanes active; row k+1 is published as it stands, i.e. without step
// k's update in the columns beyond k+1, and every lane redoes that update itself: u1 = r1 - l0(k+1) * u0, the same fma
// the owner would have done.  Results are bit-identical to the unpaired loops.
// MEASURED AND NOT USED (-DBDILU_FACTOR_GRID_PAIRED=1 to build it): stores 168 -> 128 per block, but 16 % more DFMAs and
// 8 % more instructions; cfg3 factor 2.55 ms against 2.41 ms for eliminate_diag32_grid.
__device__ __forceinline__ bool eliminate_diag32_grid2(double (&t)[2][16], double *mb, double2 *dgs, int lane)
{
    const int i = lane & 15, j = lane >> 4;
    bool viol = false;
#pragma unroll
    for (int k = 0; k < 32; k += 2) {
        const int pi0 = k & 15, a0 = k >> 4, bk = k >> 1, pi1 = pi0 + 1;  // rows k, k+1 sit in slot a0 of lanes pi0, pi1
        const int as = (k >= 15) ? 1 : 0;                                   // first slot with rows still active
        // ---- step k: pivot in lane (pi0, 0); column k+1 needs (pivot, 1/pivot) and the pivot row's entry in column k+1
        {
            const double x0 = t[a0][bk];
            const double rc0 = rcp_fast(x0);
            if (i == pi0) {
                if (j == 0) dgs[k] = make_double2(x0, rc0);
                else mb[64] = x0;  // (pi0, 1): row k, column k+1
            }
        }
        __syncwarp();
        double l0[2] = {0.0, 0.0}, l1[2] = {0.0, 0.0};
        {
            const double2 d0 = dgs[k];
            const double piv0 = d0.x, rinv0 = d0.y;
            const double u0n = (j == 1) ? mb[64] : 0.0;
#pragma unroll
            for (int a = as; a < 2; ++a) {
                const double xa = t[a][bk];
                const bool act = (i + 16 * a) > k;
                viol |= (j == 0) & act & !(fabs(xa) <= fabs(piv0));
                double q = xa * rinv0;
                q = fma(fma(-q, piv0, xa), rinv0, q);
                t[a][bk] = ((j == 0) & act) ? q : xa;
                q = group_bcast<32>(q, i);  // from lane (i, 0), which holds column k of my rows
                l0[a] = act ? q : 0.0;
                t[a][bk] = fma(-l0[a], u0n, t[a][bk]);  // column k+1 in the odd half (u0n = 0 in the even half)
            }
        }
        // ---- step k+1: pivot in lane (pi1, 1); rows k and k+1 are published together
        {
            const double x1 = t[a0][bk];
            const double rc1 = rcp_fast(x1);
            const unsigned rk = (unsigned)(i - pi0);
            if (rk < 2u) {
                double2 *dst = reinterpret_cast<double2 *>(mb + rk * 32 + j * 16);
#pragma unroll
                for (int m = (bk + 1) / 2; m < 8; ++m) dst[m] = make_double2(t[a0][2 * m], t[a0][2 * m + 1]);
                if (rk == 1u) {
                    if (j == 1) dgs[k + 1] = make_double2(x1, rc1);
                    else mb[65] = x1;  // (pi1, 0): the multiplier of row k+1 against row k
                }
            }
        }
        __syncwarp();
        {
            const double2 d1 = dgs[k + 1];
            const double piv1 = d1.x, rinv1 = d1.y, l0q = mb[65];
#pragma unroll
            for (int a = as; a < 2; ++a) {
                const double xa = t[a][bk];
                const bool act = (i + 16 * a) > k + 1;
                viol |= (j == 1) & act & !(fabs(xa) <= fabs(piv1));
                double q = xa * rinv1;
                q = fma(fma(-q, piv1, xa), rinv1, q);
                t[a][bk] = ((j == 1) & act) ? q : xa;
                q = group_bcast<32>(q, i | 16);  // from lane (i, 1), which holds column k+1 of my rows
                l1[a] = act ? q : 0.0;
            }
            const double2 *U0 = reinterpret_cast<const double2 *>(mb + j * 16);
            const double2 *R1 = reinterpret_cast<const double2 *>(mb + 32 + j * 16);
#pragma unroll
            for (int m = (bk + 1) / 2; m < 8; ++m) {
                const double2 u0 = U0[m], r1 = R1[m];
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int b = 2 * m + e;
                    if (b > bk) {
                        const double u0v = e ? u0.y : u0.x, r1v = e ? r1.y : r1.x;
                        const double u1 = fma(-l0q, u0v, r1v);
#pragma unroll
                        for (int a = as; a < 2; ++a) t[a][b] = fma(-l1[a], u1, fma(-l0[a], u0v, t[a][b]));
                    }
                }
            }
        }
        if ((k & 7) == 6 && k + 2 < 32 && __any_sync(FULL, viol)) return false;
    }
    __syncwarp();
    const double dg = dgs[lane].x;
    viol |= (dg == 0.0) || !rcp_fast_ok(dg);
    return !__any_sync(FULL, viol);
}

// the source of a fused assembly (SCALED kernels only): J, its system stride, and M = scale*J + diag_add*I
struct ScaledSrc {
    const double *src;
    int64_t sys_stride;
    double scale, diag_add;
    int ahead;  // blockw == ld == 32: L2 look-ahead distance in blocks (0 = off)
};

template <int WP, bool EXACT, bool SCALED>
__global__ void __launch_bounds__(FACTOR_WARPS * 32, WP == 32 ? 4 : (WP <= 8 ? 10 : 6))  // blockw <= 8: measured on cfg5, CTAs per SM 6: 0.69, 8: 0.61, 10: 0.58, 12: 0.62 ms
factor_kernel(double *data, int64_t sys_stride, int nsys, Dims D, int *__restrict__ ipiv,
              int *__restrict__ rowbycol, int *__restrict__ colbyrow, int *__restrict__ info, ScaledSrc SS)
{
    constexpr int G = WP;
    constexpr int BPW = 32 / G;
    // pivot-row mailbox: [warp][group][2 buffers][WP doubles + reciprocal of the pivot (+ pad)]
    __shared__ __align__(16) double urow_s[FACTOR_WARPS][BPW][2][WP + 2];
    // (U[k][k], 1/U[k][k]) of every step, left by the pivot lanes for the lanes that scale column k of sub/bot
    __shared__ __align__(16) double2 diag_s[FACTOR_WARPS][BPW][WP];

    const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
    const int grp = lane / G, r = lane % G;
    const int w = EXACT ? WP : D.w, ld = EXACT ? WP : D.ld;  // EXACT: w == ld == WP, all offsets are immediates
    const int64_t total = (int64_t)nsys * D.nblocks;
    int64_t gb = ((int64_t)blockIdx.x * FACTOR_WARPS + wic) * BPW + grp;
    const bool live = gb < total;  // dead groups run on the last block but store nothing
    if (!live) gb = total - 1;
    const int sys = (int)(gb / D.nblocks), bi = (int)(gb % D.nblocks);
    double *sysdata = data + (int64_t)sys * sys_stride;
    double *blk = sysdata + D.block(bi);
    const bool rowok = r < w;
    SrcView V;
    V.sys = SCALED ? SS.src + (int64_t)sys * SS.sys_stride : sysdata;
    V.scale = SS.scale; V.diag_add = SS.diag_add; V.scaled = SCALED;

    double (*urow)[WP + 2] = urow_s[wic][grp];
    double2 *dgs = diag_s[wic][grp];
    double a[WP];
    if constexpr (G == 32 && EXACT) {
        // L2 prefetch of the block that the warp taking this one's place will load (a CTA lives for one block per warp and
        // starts with a cold read of 8 KB): about one generation of resident warps ahead
        const int64_t ahead = gb + SS.ahead;
        if (SS.ahead > 0 && ahead < total) {
            const int psys = (int)(ahead / D.nblocks), pbi = (int)(ahead % D.nblocks);
            const char *pb = reinterpret_cast<const char *>((SCALED ? SS.src + (int64_t)psys * SS.sys_stride
                                                                    : data + (int64_t)psys * sys_stride) + D.block(pbi));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(pb + 128 * lane));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(pb + 128 * (lane + 32)));
        }
    }
    if constexpr (G == 32 && EXACT && OPTIMISTIC && GRID2D) {
        // optimistic elimination on the 2-D lane grid; on success everything is stored from here
        const int gi = lane & 15, gj = lane >> 4;
        double t[2][16];
#pragma unroll
        for (int b = 0; b < 16; ++b)
#pragma unroll
            for (int aa = 0; aa < 2; ++aa)
                t[aa][b] = V.block_at(D.block(bi) + (int64_t)(gj + 2 * b) * WP + gi + 16 * aa, gj + 2 * b == gi + 16 * aa);
        if (GRID_PAIRED ? eliminate_diag32_grid2(t, &urow[0][0], dgs, lane) : eliminate_diag32_grid(t, &urow[0][0], dgs, lane)) {
            if (live) {
#pragma unroll
                for (int b = 0; b < 16; ++b)
#pragma unroll
                    for (int aa = 0; aa < 2; ++aa) blk[(int64_t)(gj + 2 * b) * WP + gi + 16 * aa] = t[aa][b];
                const double2 dd = dgs[r];
                RowState S;
                S.pos = r; S.ipiv = r + 1; S.info = 0; S.diag = dd.x; S.rinv = dd.y; S.fast = true;
                store_block<WP, false>(a, S, V, sysdata, blk, D, w, ld, sys, bi, r, ipiv, rowbycol, colbyrow, info);
            }
            return;
        }
        __syncwarp();
    }
    // row r of the block; rows/columns beyond w are padded with the identity (they pivot on themselves
    // in the trailing steps k >= w and never disturb the real rows)
#pragma unroll
    for (int c = 0; c < WP; ++c)
        a[c] = (rowok && c < w) ? V.block_at(D.block(bi) + (int64_t)c * ld + r, c == r) : ((c == r) ? 1.0 : 0.0);

    int pos = r;          // current position of my row in the swapped ordering
    bool active = true;   // not yet chosen as a pivot row (padded rows take part: see above)
    int my_ipiv = r + 1;  // lane k keeps ipiv[k]

    bool done = false;
    if constexpr (G == 32 && PAIRED && !SHFL_ROW && OPTIMISTIC && !(EXACT && GRID2D)) {
        done = eliminate_diag32(a, urow, dgs, r);  // pos = r, my_ipiv = r + 1 stand
        if (!done) {
            __syncwarp();
#pragma unroll
            for (int c = 0; c < WP; ++c)
                a[c] = (rowok && c < w) ? V.block_at(D.block(bi) + (int64_t)c * ld + r, c == r) : ((c == r) ? 1.0 : 0.0);
        }
    }
    if constexpr (G < 32 && OPTIMISTIC) {
        done = eliminate_diag_small<WP>(a, urow, dgs, r);
        if (!done) {
            __syncwarp();
#pragma unroll
            for (int c = 0; c < WP; ++c)
                a[c] = (rowok && c < w) ? V.block_at(D.block(bi) + (int64_t)c * ld + r, c == r) : ((c == r) ? 1.0 : 0.0);
        }
    }
    if (done) {
    } else if constexpr (G == 32 && PAIRED && !SHFL_ROW) {
        // ---- two elimination steps per publication (blockw 32) --------------------------------------------------
        // A shared-memory store costs the register->MIO transfer of all 32 lanes whatever the number of active ones
        // (measured: STS.128 4.2 SM-cycles with 1 or 32 lanes, scratch/bw/mio_bench.cu), and that path bounds this
        // kernel.  So the pivot rows of steps k and k+1 go out in ONE store sequence, two lanes active:
        //   step k   : its pivot lane publishes only (pivot, 1/pivot, u0[k+1]); everybody forms l0 and updates
        //              column k+1 alone, which is all the search of step k+1 needs;
        //   step k+1 : its pivot lane publishes its row as it stands, i.e. WITHOUT step k's update in columns
        //              > k+1, together with the finished row of step k; every lane redoes that one update itself,
        //              u1[c] = r1[c] - l0(q1)*u0[c] (the same fma the pivot lane would have done), and then applies
        //              both rank-1 updates in the order of the serial algorithm.  Results are bit-identical.
        double2 *ub0 = reinterpret_cast<double2 *>(urow[0]);
        double2 *ub1 = reinterpret_cast<double2 *>(urow[1]);
        double *nx = &urow[0][WP];  // u0[k+1]
        const unsigned ub0_s = (unsigned)__cvta_generic_to_shared(ub0), ub1_s = (unsigned)__cvta_generic_to_shared(ub1);
#pragma unroll
        for (int k = 0; k < WP; k += 2) {
            // ---- step k: search, then the three numbers column k+1 needs ----
            const int hi0 = active ? (__double2hiint(a[k]) & 0x7fffffff) : -1;
            const unsigned lo0 = (unsigned)__double2loint(a[k]);
            const double rcp0 = rcp_fast(a[k]);  // speculative; only the winner's is used
            const int mhi0 = group_max_i<G>(hi0);
            const bool c10 = (hi0 == mhi0);
            const unsigned mlo0 = group_max_u<G>(c10 ? lo0 : 0u);
            const bool c20 = c10 && (lo0 == mlo0);
            const unsigned q0 = group_min_u<G>(c20 ? (unsigned)pos : 0xffffu);
            const bool isp0 = c20 && ((unsigned)pos == q0);
            if (r == k) my_ipiv = (int)q0 + 1;
            if (pos == k) pos = (int)q0;
            if (isp0) {
                pos = k;
                active = false;
                dgs[k] = make_double2(a[k], rcp0);
                *nx = a[k + 1];
            }
            __syncwarp();
            const double2 d0 = dgs[k];
            const double u0n = *nx;
            const double piv0 = d0.x, rinv0 = d0.y;
            const bool nz0 = (piv0 != 0.0);  // dgetf2 leaves the column of an exactly zero pivot unscaled
            double l0 = a[k] * rinv0;
            l0 = fma(fma(-l0, piv0, a[k]), rinv0, l0);
            const bool upd0 = active && nz0;
            a[k] = upd0 ? l0 : a[k];
            l0 = upd0 ? l0 : 0.0;
            a[k + 1] = fma(-l0, u0n, a[k + 1]);

            // ---- step k+1: search on the updated column ----
            const int hi1 = active ? (__double2hiint(a[k + 1]) & 0x7fffffff) : -1;
            const unsigned lo1 = (unsigned)__double2loint(a[k + 1]);
            const double rcp1 = rcp_fast(a[k + 1]);
            const int mhi1 = group_max_i<G>(hi1);
            const bool c11 = (hi1 == mhi1);
            const unsigned mlo1 = group_max_u<G>(c11 ? lo1 : 0u);
            const bool c21 = c11 && (lo1 == mlo1);
            const unsigned q1 = group_min_u<G>(c21 ? (unsigned)pos : 0xffffu);
            const bool isp1 = c21 && ((unsigned)pos == q1);
            if (r == k + 1) my_ipiv = (int)q1 + 1;
            if (pos == k + 1) pos = (int)q1;
            if (isp1) {
                pos = k + 1;
                active = false;
                dgs[k + 1] = make_double2(a[k + 1], rcp1);
            }
            // ---- both pivot rows, columns k.. (the pair (k, k+1) of row q1 carries l0(q1) and the pivot) ----
            {
                // ONE predicated store per column pair, executed by both pivot lanes at once (a branch here makes
                // ptxas emit one store sequence per lane, which is what this loop exists to avoid)
                const bool pub = isp0 || isp1;
                const unsigned ubw = isp1 ? ub1_s : ub0_s;
#pragma unroll
                for (int m = k / 2; m < WP / 2; ++m) sts128_if(pub, ubw + 16u * m, a[2 * m], a[2 * m + 1]);
            }
            __syncwarp();
            const double2 hd = ub1[k / 2];
            const double l0q = nz0 ? hd.x : 0.0;  // row q1 was active in step k: its a[k] is its multiplier unless !nz0
            const double piv1 = hd.y, rinv1 = dgs[k + 1].y;
            const bool nz1 = (piv1 != 0.0);
            double l1 = a[k + 1] * rinv1;
            l1 = fma(fma(-l1, piv1, a[k + 1]), rinv1, l1);
            const bool upd1 = active && nz1;
            a[k + 1] = upd1 ? l1 : a[k + 1];
            l1 = upd1 ? l1 : 0.0;
#pragma unroll
            for (int m = k / 2 + 1; m < WP / 2; ++m) {
                const double2 t0 = ub0[m], t1 = ub1[m];
                const double u1x = fma(-l0q, t0.x, t1.x), u1y = fma(-l0q, t0.y, t1.y);
                a[2 * m] = fma(-l1, u1x, fma(-l0, t0.x, a[2 * m]));
                a[2 * m + 1] = fma(-l1, u1y, fma(-l0, t0.y, a[2 * m + 1]));
            }
        }
    } else {
#pragma unroll
    for (int k = 0; k < WP; ++k) {
        // ---- pivot search (idamax over the not-yet-chosen rows, ties -> lowest current position) ----
        const int hi = active ? (__double2hiint(a[k]) & 0x7fffffff) : -1;
        const unsigned lo = (unsigned)__double2loint(a[k]);
        const double myrcp = rcp_fast(a[k]);  // speculative; only the winner's is used
        const int mhi = group_max_i<G>(hi);
        const bool c1 = (hi == mhi);
        const unsigned mlo = group_max_u<G>(c1 ? lo : 0u);
        const bool c2 = c1 && (lo == mlo);
        const unsigned q = group_min_u<G>(c2 ? (unsigned)pos : 0xffffu);  // position of the chosen row
        const bool is_p = c2 && ((unsigned)pos == q);
        if (r == k) my_ipiv = (int)q + 1;
        if (pos == k) pos = (int)q;
        if (is_p) { pos = k; active = false; }

        // ---- hand the pivot row (columns k..WP-1) and 1/pivot to the group ----
        double u[WP];
        double rinv;
        if (G == 32 && SHFL_ROW) {
            // by shuffle from the pivot lane: no shared-memory store.  (The store of a whole row by ONE lane pays
            // the operand transfer of all 32 lanes; ncu showed the register -> MIO path 78 % busy with it.)
            const int pl = __ffs(__ballot_sync(FULL, is_p)) - 1;
#pragma unroll
            for (int j = k; j < WP; ++j) u[j] = group_bcast<32>(a[j], pl);
            rinv = group_bcast<32>(myrcp, pl);
            if (is_p) dgs[k] = make_double2(a[k], myrcp);
        } else {
            double *ub = urow[k & 1];
            if (is_p) {
                double2 *u2w = reinterpret_cast<double2 *>(ub);
#pragma unroll
                for (int m = k / 2; m < WP / 2; ++m) u2w[m] = make_double2(a[2 * m], a[2 * m + 1]);
                ub[WP] = myrcp;
                dgs[k] = make_double2(a[k], myrcp);
            }
            __syncwarp();
            const double2 *u2 = reinterpret_cast<const double2 *>(ub);
#pragma unroll
            for (int m = k / 2; m < WP / 2; ++m) {
                const double2 t = u2[m];
                u[2 * m] = t.x;
                u[2 * m + 1] = t.y;
            }
            rinv = ub[WP];
        }
        const double piv = u[k];
        const bool nz = (piv != 0.0);  // dgetf2 leaves the column of an exactly zero pivot unscaled (info, below)

        // ---- multiplier by (exactly rounded) division, rank-1 update of the trailing columns ----
        double l = a[k] * rinv;
        l = fma(fma(-l, piv, a[k]), rinv, l);
        const bool upd = active && nz;
        a[k] = upd ? l : a[k];
        l = upd ? l : 0.0;
#pragma unroll
        for (int j = k + 1; j < WP; ++j) a[j] = fma(-l, u[j], a[j]);
    }
    }

    // lane k picks up U[k][k] and its reciprocal; the per-block bookkeeping happens once, here, not in every step:
    // pivots outside the exact range of the reciprocal send the warp to the slow path, the first exactly zero
    // pivot is dgetrf's info (1-based).
    __syncwarp();
    const double2 dd = dgs[r];
    const double my_diag = dd.x, my_rinv = dd.y;
    const bool weird = (my_diag != 0.0) && !rcp_fast_ok(my_diag);
    const unsigned zb = __ballot_sync(FULL, my_diag == 0.0);
    const unsigned zg = (G == 32) ? zb : ((zb >> (grp * G)) & ((1u << (G & 31)) - 1u));
    const int my_info = zg ? __ffs(zg) : 0;
    if (__any_sync(FULL, weird)) {
        // rare: redo this warp's blocks with true divisions from the pristine data (nothing stored yet)
        factor_block_slow<WP>(V, sysdata, blk, D, w, ld, sys, bi, r, live, urow[0], ipiv, rowbycol, colbyrow, info);
        return;
    }
    if (live && rowok) {
        RowState S;
        S.pos = pos; S.ipiv = my_ipiv; S.info = my_info; S.diag = my_diag; S.rinv = my_rinv; S.fast = true;
        store_block<WP>(a, S, V, sysdata, blk, D, w, ld, sys, bi, r, ipiv, rowbycol, colbyrow, info);
    }
}

// 32 < blockw <= 64: one 64-thread CTA per block, thread r keeps row r (64 doubles) in registers.
// Same algorithm; the pivot search and the pivot-row mailbox cross the two warps through shared
// memory and __syncthreads().
__global__ void __launch_bounds__(64)
factor_wide_kernel(double *__restrict__ data, int64_t sys_stride, int nsys, Dims D,
                   int *__restrict__ ipiv, int *__restrict__ rowbycol, int *__restrict__ colbyrow,
                   int *__restrict__ info)
{
    constexpr int WP = 64;
    __shared__ __align__(16) double urow[2][WP];
    __shared__ double cand_v[2][2];
    __shared__ int cand_pos[2][2], cand_row[2][2];

    const int r = threadIdx.x, lane = r & 31, wid = r >> 5;
    const int w = D.w, ld = D.ld;
    const int64_t gb = blockIdx.x;
    const int sys = (int)(gb / D.nblocks), bi = (int)(gb % D.nblocks);
    double *sysdata = data + (int64_t)sys * sys_stride;
    double *blk = sysdata + D.block(bi);
    const bool rowok = r < w;

    double a[WP];
#pragma unroll
    for (int c = 0; c < WP; ++c)
        a[c] = (rowok && c < w) ? ld_stream(blk + (int64_t)c * ld + r) : ((c == r) ? 1.0 : 0.0);

    int pos = r;
    bool active = rowok;
    int my_ipiv = r + 1;
    double my_diag = 1.0;
    int my_info = 0;

#pragma unroll
    for (int k = 0; k < WP; ++k) {
        if (k < w) {
            const double v = active ? fabs(a[k]) : -1.0;
            const int pl = group_argmax<32>(v, pos, lane);  // winner inside my warp
            if (lane == pl) { cand_v[k & 1][wid] = v; cand_pos[k & 1][wid] = pos; cand_row[k & 1][wid] = r; }
            __syncthreads();
            const double v0 = cand_v[k & 1][0], v1 = cand_v[k & 1][1];
            const int p0 = cand_pos[k & 1][0], p1 = cand_pos[k & 1][1];
            const bool take1 = (v1 > v0) || (v1 == v0 && p1 < p0);
            const int p = take1 ? cand_row[k & 1][1] : cand_row[k & 1][0];
            const int q = take1 ? p1 : p0;
            if (r == k) my_ipiv = q + 1;
            if (pos == k) pos = q;
            if (r == p) { pos = k; active = false; }

            double2 *u2w = reinterpret_cast<double2 *>(urow[k & 1]);
            if (r == p) {
#pragma unroll
                for (int m = k / 2; m < WP / 2; ++m) u2w[m] = make_double2(a[2 * m], a[2 * m + 1]);
            }
            __syncthreads();
            const double2 *u2 = reinterpret_cast<const double2 *>(urow[k & 1]);
            const double piv = urow[k & 1][k];
            if (r == k) my_diag = piv;
            if (piv == 0.0) {
                if (my_info == 0) my_info = k + 1;
            } else if (active) {
                const double l = a[k] / piv;
                a[k] = l;
#pragma unroll
                for (int m = k / 2; m < WP / 2; ++m) {
                    const double2 t = u2[m];
                    if (2 * m > k) a[2 * m] = fma(-l, t.x, a[2 * m]);
                    if (2 * m + 1 > k) a[2 * m + 1] = fma(-l, t.y, a[2 * m + 1]);
                }
            }
        }
    }

    if (rowok) {
#pragma unroll
        for (int c = 0; c < WP; ++c)
            if (c < w) blk[(int64_t)c * ld + pos] = a[c];
        const int64_t ib = ((int64_t)sys * D.nblocks + bi) * w;
        ipiv[ib + r] = my_ipiv;
        rowbycol[ib + pos] = r;
        colbyrow[ib + r] = pos;
        for (int di = 0; di < D.ndiag && bi + di < D.nblocks - 1; ++di) {
            double *s = sysdata + D.off_sub + D.diag_chunk(di, bi) + r;
            *s = *s / my_diag;
        }
        if (bi < D.nsat)
            for (int sati = bi; sati < D.nsat; ++sati) {
                double *s = sysdata + D.off_bot + D.sat_chunk(sati, bi) + r;
                *s = *s / my_diag;
            }
        if (my_info != 0 && r == 0) {
            const int key = (bi << 7) | my_info;
            int old = atomicCAS(&info[sys], 0, key);
            while (old != 0 && (old >> 7) > bi) {
                const int assumed = old;
                old = atomicCAS(&info[sys], assumed, key);
                if (old == assumed) break;
            }
        }
    }
}

// src == nullptr: in place.  Otherwise the fused assembly data = LU(scale*src + diag_add*I) (src may equal data).
#ifndef BDILU_FACTOR_AHEAD
#define BDILU_FACTOR_AHEAD 16  // look-ahead in units of (SMs x warps per CTA): 4 CTAs x 4 warps resident per SM
#endif
static cudaError_t launch_factor_any(double *data, int64_t sys_stride, int nsys, const Dims &D, int *ipiv, int *rowbycol,
                                     int *colbyrow, int *info, const ScaledSrc &SS0, cudaStream_t st)
{
    ScaledSrc SS = SS0;
    SS.ahead = (D.w == 32 && D.ld == 32) ? sm_count() * BDILU_FACTOR_AHEAD : 0;
    const int64_t total = (int64_t)nsys * D.nblocks;
    if (total == 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(info, 0, sizeof(int) * (size_t)nsys, st);
    if (e != cudaSuccess) return e;
    const bool scaled = SS.src != nullptr;
#define BDILU_LAUNCH_FACTOR(WP)                                                                                    \
    {                                                                                                              \
        const int64_t per_cta = (int64_t)FACTOR_WARPS * (32 / WP);                                                 \
        const unsigned grid = (unsigned)((total + per_cta - 1) / per_cta);                                         \
        const bool exact = (D.w == WP && D.ld == WP);                                                              \
        auto kern = scaled ? (exact ? factor_kernel<WP, true, true> : factor_kernel<WP, false, true>)              \
                           : (exact ? factor_kernel<WP, true, false> : factor_kernel<WP, false, false>);           \
        kern<<<grid, FACTOR_WARPS * 32, 0, st>>>(data, sys_stride, nsys, D, ipiv, rowbycol, colbyrow, info, SS);   \
    }
    const int w = D.w;
    if (w <= 2) BDILU_LAUNCH_FACTOR(2)
    else if (w <= 4) BDILU_LAUNCH_FACTOR(4)
    else if (w <= 8) BDILU_LAUNCH_FACTOR(8)
    else if (w <= 16) BDILU_LAUNCH_FACTOR(16)
    else if (w <= 32) BDILU_LAUNCH_FACTOR(32)
    else {
        // 32 < blockw: no fused load in the wide kernel; assemble with the streaming kernel first
        if (scaled) {
            e = launch_scale_diag_add(data, SS.src, sys_stride, SS.sys_stride, nsys, D, SS.scale, SS.diag_add, st);
            if (e != cudaSuccess) return e;
        }
        factor_wide_kernel<<<(unsigned)total, 64, 0, st>>>(data, sys_stride, nsys, D, ipiv, rowbycol, colbyrow, info);
    }
#undef BDILU_LAUNCH_FACTOR
    count_launch(1);
    return cudaGetLastError();
}

cudaError_t launch_factor(double *data, int64_t sys_stride, int nsys, const Dims &D, int *ipiv,
                          int *rowbycol, int *colbyrow, int *info, cudaStream_t st)
{
    const ScaledSrc none = {nullptr, 0, 1.0, 0.0, 0};
    return launch_factor_any(data, sys_stride, nsys, D, ipiv, rowbycol, colbyrow, info, none, st);
}

cudaError_t launch_factor_scaled(double *dst, int64_t dst_stride, const double *src, int64_t src_stride, int nsys,
                                 const Dims &D, double scale, double diag_add, int *ipiv, int *rowbycol, int *colbyrow,
                                 int *info, cudaStream_t st)
{
    const ScaledSrc SS = {src, src_stride, scale, diag_add, 0};
    return launch_factor_any(dst, dst_stride, nsys, D, ipiv, rowbycol, colbyrow, info, SS, st);
}

}  // namespace bdilu
