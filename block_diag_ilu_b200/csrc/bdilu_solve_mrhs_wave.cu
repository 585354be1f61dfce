// Kernel 2 (many right-hand sides per factorisation), wavefront form: ILU_inplace<double>::solve
// (block_diag_ilu.hpp:486-538) for a group of R = 32/Q vectors that share one factor, with K warps of one CTA
// working on K CONSECUTIVE blocks of the chain at once.
//
// bdilu_solve_mrhs.cu walks the chain with one warp per group: every block costs the full dependent sequence of its
// two-column steps plus staging, coupling and publication (~6800 cycles per blockw-64 block on B200).  Row pair p of
// block b needs from its predecessors only y_{b-di}[colbyrow[2p]], y_{b-di}[colbyrow[2p+1]] (:508-511, 524-527), so
// block b+1 can run a few steps behind block b.  Here:
//   * iteration `it` of the chain (forward blocks 0..N-1, then backward blocks N-1..0) belongs to consumer warp
//     it % K; lane layout, two-column steps and the shared-memory reads of L/U are those of bdilu_solve_mrhs.cu;
//   * a dedicated producer warp stages blocks by bulk copy (cp.async.bulk + mbarrier), NSTG deep;
//   * the owner lanes of a pivot pair publish its two finished values into a ring slot [it % NSLOT][vector][row]
//     and then bump that slot's per-vector progress word (iteration << 8 | pairs published): same thread, same
//     shared-memory pipe, so the data is visible before the count;
//   * the coupling terms are merged for G (1, 2 or 4) consecutive pivot pairs at a time, right after the last step of
//     the group before (blocks then trail each other by ~G+1 steps; merging a whole group earlier hides the merge
//     latency but doubles the distance, measured slower: cfg4 7.7 vs 6.6 ms):
//     every lane polls the progress words of its vector until the predecessors have published the furthest pair the
//     permutation of the group's rows points at; inside a group every pair has its own owner lane, so the G pairs
//     are merged by one branch-free instruction sequence (volatile reads of the ring);
//   * a ring slot is reused NSLOT = K + ndiag + 1 iterations later, after checking that its last readers have
//     finished; y between the sweeps goes through x in global memory, with one named barrier at the turn.
// Every wait is on a strictly earlier iteration of the chain, so the earliest unfinished block can always proceed.
// Polls trap after 2^26 failures instead of hanging the device.
#include <stdlib.h>

#include "bdilu_mrhs_common.cuh"

namespace bdilu {
namespace mrhs {

using wave::ld_volatile_s;
using wave::st_volatile_s;

constexpr int KMAX = 8;  // consumer warps per CTA
// pivot pairs per coupling merge at blockw 64 / blockw 32 (narrower blocks merge pair by pair)
#ifndef BDILU_MRW_G32
#define BDILU_MRW_G32 4
#endif
#ifndef BDILU_MRW_G16
#define BDILU_MRW_G16 2
#endif
// 0: group g+1 is merged when group g starts (its latency hides behind G steps, blocks trail by ~2G steps);
// 1: merged after the last step of group g (blocks trail by ~G+1 steps, the merge sits on the chain)
#ifndef BDILU_MRW_LATE
#define BDILU_MRW_LATE 1
#endif

struct WArgs {
    const double *sd;
    const int *rbc_g, *cbr_g;
    const double *bq;
    double *xq;
    unsigned char *stages;
    unsigned long long *full, *empty;
    double *ring;  // this lane's vector: + j * WPP; layout [slot][R][WPP]
    int *prog;     // this lane's vector: + j;       layout [slot][R]
    int *done;     // [K] last iteration each consumer warp has finished
    double *dinv;  // this warp's [WP]
    int *needw;    // this warp's [8]: pairs a predecessor must have published before group g is merged
    Layout SL;
    Dims D;
    int K, NSTG, NSLOT, lane, q, wic;
    bool live;
};

// Publication stores: volatile asm keeps them in program order among themselves (data before count); there is no
// "memory" clobber on purpose, so the compiler may still move the next steps' ordinary shared-memory loads of L/U
// across them (with the clobber every step became a scheduling barrier: 7.7 -> see header for the effect).
__device__ __forceinline__ void publish_pair(double *p, double a, double b)
{
    asm volatile("st.volatile.shared.v2.f64 [%0], {%1, %2};" ::"r"(wave::smem_u32(p)), "d"(a), "d"(b));
}
__device__ __forceinline__ void publish_count(int *p, int v)
{
    asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"(wave::smem_u32(p)), "r"(v));
}

__device__ __forceinline__ void poll_ge(const int *pw, int target)
{
    int spins = 0;
    while (!__all_sync(FULL, ld_volatile_s(pw) >= target)) {
        if (++spins > (1 << 26)) __trap();
    }
}

template <int WP, int Q, bool FWD>
__device__ __forceinline__ void block_iter(const WArgs &A, int it, double (&acc)[WP / Q], double (&pre)[WP / Q],
                                           int (&nidx)[WP / Q], int &nan_first, int &zero_last)
{
    using C = Cfg<WP, Q>;
    constexpr int NP = C::NP, WPP = C::WPP, R = C::R, HW = WP / 2, w = WP;
    constexpr int RING_SLOT = R * WPP;
    const int ld = A.D.ld, N = A.D.nblocks, nd = A.D.ndiag, NIT = 2 * N, K = A.K, NSLOT = A.NSLOT;
    const int q = A.q;
    const int bi = FWD ? it : NIT - 1 - it;
    const int s = it % A.NSTG, sl = it % NSLOT;

    // the slot this block publishes into was read last by iterations it-NSLOT+1 .. it-NSLOT+ndiag: they must be over
    for (int di = 1; di <= nd; ++di) {
        const int itr = it - NSLOT + di;
        if (itr >= 0) poll_ge(A.done + itr % K, itr);
    }
    mbar_wait(&A.full[s], (it / A.NSTG) & 1);
    const unsigned char *stg = A.stages + (size_t)s * A.SL.stage_bytes;
    const double *blk = reinterpret_cast<const double *>(stg);
    const int *cbr_s = reinterpret_cast<const int *>(stg + A.SL.o_cbr);
    const double *dg_s = reinterpret_cast<const double *>(stg + A.SL.o_diag);

#pragma unroll
    for (int i = 0; i < 2 * NP; ++i) acc[i] = pre[i];
    if (FWD) {
        const int *rbc_s = reinterpret_cast<const int *>(stg + A.SL.o_rbc);
#pragma unroll
        for (int i = 0; i < 2 * NP; ++i) {
            if (isnan(acc[i])) {
                const int p = (i >> 1) * Q + q;
                nan_first = min(nan_first, bi * w + rbc_s[2 * p + (i & 1)] + 1);  // :500-501
            }
        }
    }
    // right-hand side of this warp's next block: b through rowbycol fetched one block earlier, or y from x
    {
        const int nit = it + K;
        if (FWD) {
            if (nit < N) {
                const double *bn = A.bq + (int64_t)nit * w;
#pragma unroll
                for (int i = 0; i < 2 * NP; ++i) pre[i] = __ldg(bn + nidx[i]);
            }
            if (nit + K < N) {
                const int *rn = A.rbc_g + (int64_t)(nit + K) * w;
#pragma unroll
                for (int i = 0; i < 2 * NP; ++i) nidx[i] = __ldg(rn + 2 * ((i >> 1) * Q + q) + (i & 1));
            }
        } else if (nit < NIT) {
            const double *yn = A.xq + (int64_t)(NIT - 1 - nit) * w;
#pragma unroll
            for (int i = 0; i < 2 * NP; ++i) pre[i] = __ldcg(yn + 2 * ((i >> 1) * Q + q) + (i & 1));
        }
    }
    if (!FWD) {
        for (int t = A.lane; t < w; t += 32) {
            const double dgv = blk[t * ld + t];
            if (dgv == 0.0) zero_last = max(zero_last, bi * w + t);  // :533-534
            A.dinv[t] = 1.0 / dgv;
        }
        __syncwarp();
    }

    const int dmax = FWD ? min(nd, bi) : min(nd, N - 1 - bi);
    // Coupling terms are merged for G consecutive pivot pairs at a time, one group ahead of the pivot: group g (in
    // processing order) needs the predecessors to have published needw[g] pairs -- the furthest pair the permutation
    // of its rows points at.  Inside a group every pair has its own owner lane, so all G pairs are merged at once.
    constexpr int G0 = HW >= 32 ? BDILU_MRW_G32 : (HW >= 16 ? BDILU_MRW_G16 : 1);
    constexpr int G = G0 < Q ? G0 : Q;
    static_assert(Q % G == 0, "a group of pivot pairs maps to consecutive owner lanes and one register index");
    constexpr int NG = HW / G;
    int ci[2 * NP];
    int spo[MAXD];
    if (dmax > 0) {
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const int p = i * Q + q;
            ci[2 * i] = cbr_s[2 * p];
            ci[2 * i + 1] = cbr_s[2 * p + 1];
        }
        int sp = sl;
#pragma unroll
        for (int di = 0; di < MAXD; ++di) {
            sp = sp == 0 ? NSLOT - 1 : sp - 1;  // slot of iteration it - 1 - di
            spo[di] = sp;
        }
        if (A.lane < HW) {
            const int c0 = cbr_s[2 * A.lane], c1 = cbr_s[2 * A.lane + 1];
            int n = FWD ? (max(c0, c1) >> 1) + 1 : HW - (min(c0, c1) >> 1);
#pragma unroll
            for (int off = 1; off < G; off <<= 1) n = max(n, __shfl_xor_sync(HW >= 32 ? FULL : ((1u << HW) - 1), n, off));
            const int tau = FWD ? A.lane : HW - 1 - A.lane;
            if (tau % G == 0) A.needw[tau / G] = n;
        }
        __syncwarp();
    }
    auto merge_group = [&](int g) {
        if (dmax > 0) {
            const int pk0 = FWD ? g * G : HW - (g + 1) * G;  // lowest pair of the group
            const int i = pk0 / Q, q0 = pk0 % Q;
            const bool mine = q >= q0 && q < q0 + G;
            const int need = A.needw[g];
            int spins = 0, ok;
            do {
                ok = 1;
#pragma unroll
                for (int di = 0; di < MAXD; ++di)
                    if (di < dmax) ok &= ld_volatile_s(A.prog + spo[di] * R) >= ((it - 1 - di) << 8) + need;
                if (++spins > (1 << 26)) __trap();
            } while (!__all_sync(FULL, ok));
            double t0 = acc[2 * i], t1 = acc[2 * i + 1];
            const int c0 = ci[2 * i], c1 = ci[2 * i + 1];
#pragma unroll
            for (int di = 0; di < MAXD; ++di) {
                if (di < dmax) {
                    const double *rg = A.ring + spo[di] * RING_SLOT;
                    const double *dg = dg_s + di * w;
                    t0 = fma(-dg[c0], ld_volatile_s(rg + c0), t0);
                    t1 = fma(-dg[c1], ld_volatile_s(rg + c1), t1);
                }
            }
            acc[2 * i] = mine ? t0 : acc[2 * i];
            acc[2 * i + 1] = mine ? t1 : acc[2 * i + 1];
        }
    };
    double *rgc = A.ring + sl * RING_SLOT;
    int *pgc = A.prog + sl * R;

    if (FWD) {
        merge_group(0);
#pragma unroll
        for (int pk = 0; pk < HW; ++pk) {
            if (!BDILU_MRW_LATE && pk % G == 0 && pk / G + 1 < NG) merge_group(pk / G + 1);
            const int qo = pk % Q, io = (pk / Q) * 2;
            const double *c0 = blk + (2 * pk) * ld;
            const double *c1 = c0 + ld;
            double y0 = acc[io];
            double y1 = fma(-c0[2 * pk + 1], y0, acc[io + 1]);  // unit lower L (:503-507)
            y0 = quad_bcast<Q>(y0, qo);
            y1 = quad_bcast<Q>(y1, qo);
            if (q == qo) {
                acc[io + 1] = y1;
                publish_pair(rgc + 2 * pk, y0, y1);
                publish_count(pgc, (it << 8) | (pk + 1));
            }
#pragma unroll
            for (int i = pk / Q; i < NP; ++i) {
                if (i * Q + Q - 1 > pk) {
                    const int p = i * Q + q;
                    if (i * Q > pk || p > pk) {
                        const double2 a = *reinterpret_cast<const double2 *>(c0 + 2 * p);
                        const double2 b2 = *reinterpret_cast<const double2 *>(c1 + 2 * p);
                        acc[2 * i] = fma(-a.x, y0, acc[2 * i]);
                        acc[2 * i + 1] = fma(-a.y, y0, acc[2 * i + 1]);
                        acc[2 * i] = fma(-b2.x, y1, acc[2 * i]);
                        acc[2 * i + 1] = fma(-b2.y, y1, acc[2 * i + 1]);
                    }
                }
            }
            if (BDILU_MRW_LATE && pk % G == G - 1 && pk / G + 1 < NG) merge_group(pk / G + 1);
        }
    } else {
        merge_group(0);
#pragma unroll
        for (int pkr = 0; pkr < HW; ++pkr) {
            const int pk = HW - 1 - pkr;
            if (!BDILU_MRW_LATE && pkr % G == 0 && pkr / G + 1 < NG) merge_group(pkr / G + 1);
            const int qo = pk % Q, io = (pk / Q) * 2;
            const double *c0 = blk + (2 * pk) * ld;
            const double *c1 = c0 + ld;
            double x1 = acc[io + 1] * A.dinv[2 * pk + 1];
            double x0 = fma(-c1[2 * pk], x1, acc[io]) * A.dinv[2 * pk];  // U[2pk, 2pk+1] (:519-523, 531-532)
            x0 = quad_bcast<Q>(x0, qo);
            x1 = quad_bcast<Q>(x1, qo);
            if (q == qo) {
                acc[io] = x0;
                acc[io + 1] = x1;
                publish_pair(rgc + 2 * pk, x0, x1);
                publish_count(pgc, (it << 8) | (pkr + 1));
            }
#pragma unroll
            for (int i = 0; i <= pk / Q; ++i) {
                if (i * Q < pk) {
                    const int p = i * Q + q;
                    if (i * Q + Q - 1 < pk || p < pk) {
                        const double2 a = *reinterpret_cast<const double2 *>(c0 + 2 * p);
                        const double2 b2 = *reinterpret_cast<const double2 *>(c1 + 2 * p);
                        acc[2 * i] = fma(-b2.x, x1, acc[2 * i]);
                        acc[2 * i + 1] = fma(-b2.y, x1, acc[2 * i + 1]);
                        acc[2 * i] = fma(-a.x, x0, acc[2 * i]);
                        acc[2 * i + 1] = fma(-a.y, x0, acc[2 * i + 1]);
                    }
                }
            }
            if (BDILU_MRW_LATE && pkr % G == G - 1 && pkr / G + 1 < NG) merge_group(pkr / G + 1);
        }
    }

    // x (y between the sweeps, :499) and release of the stage
    if (A.live) {
        double *xo = A.xq + (int64_t)bi * w;
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const int p = i * Q + q;
            xo[2 * p] = acc[2 * i];
            xo[2 * p + 1] = acc[2 * i + 1];
        }
    }
    __syncwarp();
    if (A.lane == 0) {
        mbar_arrive(&A.empty[s]);
        st_volatile_s(A.done + A.wic, it);
    }
}

// Register budget: ptxas' choice under a 288-thread bound measured faster at blockw 64 (6.6 -> 5.9 ms on cfg4) and slower
// below (blockw 32: 14.5 -> 18.3 ms, blockw 16: 1.05 -> 1.72 ms) than under a 224-thread bound, so the bound follows the width.
template <int WP, int Q>
__global__ void __launch_bounds__(WP == 64 ? (KMAX + 1) * 32 : (KMAX - 1) * 32)
solve_mrhs_wave_kernel(const double *__restrict__ data, int64_t sys_stride, Dims D, const int *__restrict__ rowbycol,
                       const int *__restrict__ colbyrow, const double *__restrict__ bvec, double *xvec, VecLayout V,
                       int *__restrict__ flags, int groups_per_sys, Layout SL, int NSTG)
{
    using C = Cfg<WP, Q>;
    constexpr int NP = C::NP, WPP = C::WPP, R = C::R, w = WP;
    extern __shared__ __align__(128) unsigned char smem[];
    const int K = (blockDim.x >> 5) - 1;
    const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
    const int q = lane % Q, j = lane / Q;
    const int N = D.nblocks, nd = D.ndiag, NIT = 2 * N;
    const int NSLOT = K + nd + 1;

    const int sys = blockIdx.x / groups_per_sys, grp = blockIdx.x % groups_per_sys;
    const int rhs_raw = grp * R + j;
    const bool live = rhs_raw < V.nrhs;
    const int rhs = live ? rhs_raw : V.nrhs - 1;

    WArgs A;
    A.sd = data + (int64_t)sys * sys_stride;
    A.rbc_g = rowbycol + (int64_t)sys * N * w;
    A.cbr_g = colbyrow + (int64_t)sys * N * w;
    A.bq = bvec + (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride;
    A.xq = xvec + (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride;
    A.stages = smem;
    A.full = reinterpret_cast<unsigned long long *>(smem + (size_t)NSTG * SL.stage_bytes);
    A.empty = A.full + NSTG;
    double *ring0 = reinterpret_cast<double *>(A.empty + NSTG);
    double *dinv0 = ring0 + (size_t)NSLOT * R * WPP;
    int *prog0 = reinterpret_cast<int *>(dinv0 + (size_t)K * WP);
    int *done0 = prog0 + NSLOT * R;
    int *need0 = done0 + KMAX;      // [K][8]
    int *nanw = need0 + KMAX * 8;   // [R] first NaN per vector
    int *zerow = nanw + R;          // [1] zero U diagonal (same for every vector of the system)
    A.ring = ring0 + (size_t)j * WPP;
    A.prog = prog0 + j;
    A.done = done0;
    A.needw = need0 + (wic < K ? wic : 0) * 8;
    A.dinv = dinv0 + (size_t)(wic < K ? wic : 0) * WP;
    A.SL = SL; A.D = D; A.K = K; A.NSTG = NSTG; A.NSLOT = NSLOT; A.lane = lane; A.q = q; A.wic = wic; A.live = live;

    if (threadIdx.x == 0) {
        for (int s = 0; s < NSTG; ++s) {
            mbar_init(&A.full[s], 1);
            mbar_init(&A.empty[s], 1);
        }
        *zerow = -1;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int t = threadIdx.x; t < NSLOT * R; t += blockDim.x) prog0[t] = -1;
    if (threadIdx.x < R) nanw[threadIdx.x] = 0x7fffffff;
    if (threadIdx.x < KMAX) done0[threadIdx.x] = -1;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();

    if (wic == K) {
        // ---- producer warp: one lane stages the chain's iterations in order ----
        if (lane == 0) {
            for (int pit = 0; pit < NIT; ++pit) {
                const int ps = pit % NSTG;
                const bool pf = pit < N;
                const int pb = pf ? pit : NIT - 1 - pit;
                if (pit >= NSTG) mbar_wait(&A.empty[ps], ((pit / NSTG) - 1) & 1);
                const int pnd = pf ? min(nd, pb) : min(nd, N - 1 - pb);
                unsigned char *st = A.stages + (size_t)ps * SL.stage_bytes;
                mbar_expect_tx(&A.full[ps], SL.blk_bytes + (pf ? 2 : 1) * SL.idx_bytes + pnd * SL.diag_bytes);
                bulk_g2s(st, A.sd + D.block(pb), SL.blk_bytes, &A.full[ps]);
                bulk_g2s(st + SL.o_cbr, A.cbr_g + (int64_t)pb * w, SL.idx_bytes, &A.full[ps]);
                if (pf) bulk_g2s(st + SL.o_rbc, A.rbc_g + (int64_t)pb * w, SL.idx_bytes, &A.full[ps]);
                for (int d = 0; d < pnd; ++d) {
                    // forward: sub(d, b-1-d, :) couples block b to y_{b-1-d}; backward: sup(d, b, :) to x_{b+1+d}
                    const double *src = pf ? A.sd + D.off_sub + D.diag_chunk(d, pb - 1 - d) : A.sd + D.off_sup + D.diag_chunk(d, pb);
                    bulk_g2s(st + SL.o_diag + d * SL.diag_bytes, src, SL.diag_bytes, &A.full[ps]);
                }
            }
        }
        return;
    }

    // ---- consumer warps ----
    int nan_first = 0x7fffffff, zero_last = -1;
    double acc[2 * NP], pre[2 * NP];
    int nidx[2 * NP];
    int it = wic;
#pragma unroll
    for (int i = 0; i < 2 * NP; ++i) {
        pre[i] = 0.0;
        nidx[i] = 0;
    }
    if (it < N) {
        const int *r0 = A.rbc_g + (int64_t)it * w;
#pragma unroll
        for (int i = 0; i < 2 * NP; ++i) nidx[i] = __ldg(r0 + 2 * ((i >> 1) * Q + q) + (i & 1));
#pragma unroll
        for (int i = 0; i < 2 * NP; ++i) pre[i] = __ldg(A.bq + (int64_t)it * w + nidx[i]);
        if (it + K < N) {
            const int *r1 = A.rbc_g + (int64_t)(it + K) * w;
#pragma unroll
            for (int i = 0; i < 2 * NP; ++i) nidx[i] = __ldg(r1 + 2 * ((i >> 1) * Q + q) + (i & 1));
        }
    }
    for (; it < N; it += K) block_iter<WP, Q, true>(A, it, acc, pre, nidx, nan_first, zero_last);

    // the turn: every y is in global memory before any warp starts reading it back
    __threadfence_block();
    asm volatile("bar.sync 1, %0;" ::"r"(K * 32) : "memory");
    if (it < NIT) {
        const double *y0 = A.xq + (int64_t)(NIT - 1 - it) * w;
#pragma unroll
        for (int i = 0; i < 2 * NP; ++i) pre[i] = __ldcg(y0 + 2 * ((i >> 1) * Q + q) + (i & 1));
    }
    for (; it < NIT; it += K) block_iter<WP, Q, false>(A, it, acc, pre, nidx, nan_first, zero_last);

    // reference return value: NaN index wins, else N*w + index of the zero diagonal, else 0 (:491-495)
#pragma unroll
    for (int off = Q / 2; off > 0; off >>= 1) nan_first = min(nan_first, __shfl_xor_sync(FULL, nan_first, off, Q));
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) zero_last = max(zero_last, __shfl_xor_sync(FULL, zero_last, off));
    if (q == 0) atomicMin(&nanw[j], nan_first);
    if (lane == 0) atomicMax(zerow, zero_last);
    asm volatile("bar.sync 1, %0;" ::"r"(K * 32) : "memory");
    if (wic == 0 && live && q == 0 && flags) {
        const int nf = nanw[j], zl = *zerow;
        flags[(int64_t)sys * V.nrhs + rhs] = (nf != 0x7fffffff) ? nf : (zl >= 0 ? N * w + zl : 0);
    }
}

template <int WP, int Q>
static cudaError_t launch_w(unsigned grid, int K, size_t smem, const double *data, int64_t sys_stride, const Dims &D,
                            const int *rowbycol, const int *colbyrow, const double *b, double *x, const VecLayout &V,
                            int *flags, int gps, const Layout &SL, int NSTG, cudaStream_t st)
{
    // per (function, device): set on every launch
    cudaError_t e = cudaFuncSetAttribute(solve_mrhs_wave_kernel<WP, Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    solve_mrhs_wave_kernel<WP, Q><<<grid, (K + 1) * 32, smem, st>>>(data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, gps,
                                                                     SL, NSTG);
    return cudaGetLastError();
}

}  // namespace mrhs

// Wavefront form of the many-vector solve (full power-of-two block widths 8..64).  Q = lanes per vector.
bool try_launch_solve_mrhs_wave(int Q, const double *data, int64_t sys_stride, int nsys, const Dims &D, const int *rowbycol,
                                const int *colbyrow, const double *b, double *x, const VecLayout &V, int *flags,
                                cudaStream_t st, cudaError_t *err)
{
    using namespace mrhs;
    *err = cudaSuccess;
    const int w = D.w;
    if (!(w == 8 || w == 16 || w == 32 || w == 64) || w < 2 * Q) return false;
    // the progress word is (iteration << 8 | pairs) in an int32 and there are 2*nblocks iterations: longer chains go to
    // the other kernels
    if (D.nblocks >= (1 << 21)) return false;
    const int R = 32 / Q;
    const int gps = (V.nrhs + R - 1) / R;
    const Layout SL = make_layout(w, D.ld, D.ndiag);
    // measured (8 x 64 vectors at blockw 32 / 64 vectors at blockw 16): K=4 18.4 / 1.52 ms, K=6 14.4 / 1.05-1.19 ms; blockw 64 fits 4
    int K = (w == 64 ? 4 : 6);
    int NSTG = 0;
    size_t smem = 0;
    for (; K >= 1; --K) {
        const int NSLOT = K + D.ndiag + 1;
        const size_t fixed = (size_t)NSLOT * R * (w + 8) * 8 + (size_t)K * w * 8 + (size_t)NSLOT * R * 4 + (R + 1 + 9 * KMAX) * 4 + 64;
        for (NSTG = 2 * K; NSTG >= K + 2; --NSTG) {
            smem = (size_t)NSTG * (SL.stage_bytes + 16) + fixed;
            if (smem <= 227 * 1024) break;
        }
        if (NSTG >= K + 2) break;
    }
    if (K < 1) return false;
    const int64_t grid = (int64_t)nsys * gps;
    if (grid > 0x7fffffff) return false;
#define BDILU_MRW(WPV, QV) \
    *err = launch_w<WPV, QV>((unsigned)grid, K, smem, data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, gps, SL, NSTG, st)
    if (Q == 4) {
        if (w == 8) BDILU_MRW(8, 4);
        else if (w == 16) BDILU_MRW(16, 4);
        else if (w == 32) BDILU_MRW(32, 4);
        else BDILU_MRW(64, 4);
    } else if (Q == 8) {
        if (w == 16) BDILU_MRW(16, 8);
        else if (w == 32) BDILU_MRW(32, 8);
        else BDILU_MRW(64, 8);
    } else {
        if (w == 32) BDILU_MRW(32, 16);
        else BDILU_MRW(64, 16);
    }
#undef BDILU_MRW
    count_launch(1);
    return true;
}

}  // namespace bdilu
