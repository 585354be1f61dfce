// Kernel 2 (many right-hand sides per factorisation): ILU_inplace<double>::solve (block_diag_ilu.hpp:486-538)
// applied to a batch of vectors that share one factor -- the regime of BASELINE cfg4 (256 RHS, blockw 64) and of
// the preconditioner reuse inside a Newton/Krylov iteration (anyode_iterative.hpp:64-88 calls solve() once per
// Krylov vector; the reference has no multi-RHS entry point).
//
// The one-vector kernels spend a whole SM (quad32) or a whole warp plus a private pass over the factors (chain
// kernel) per vector.  Here a warp carries R = 32/Q vectors at once (Q = 4, 8 or 16 lanes per vector):
//   lane = (j, q): j = vector inside the warp's group of R, q = which share of the block's row PAIRS the lane owns
//   (pair p = i*Q + q holds rows 2p, 2p+1).  A block of the factor is staged ONCE per CTA by one bulk copy
//   (cp.async.bulk + mbarrier, NS-deep ring) together with its permutation vectors and the sub/sup rows it needs,
//   and is consumed from shared memory by every warp of the CTA: the L/U entries of a row pair come from one
//   LDS.128 whose address depends on q only (Q distinct addresses per warp, a broadcast for the R vectors).
//   The triangular solve advances TWO columns per step: the owner of pair p finishes both values locally, two
//   width-Q shuffles hand them to the other lanes of the vector, every lane updates its later (forward) or
//   earlier (backward) pairs with 4 DFMA per pair.  The last ndiag blocks' results live in a per-warp shared ring
//   for the coupling terms; b (forward) and y (backward) are prefetched one block ahead into registers.
//   Backward: 1/U[k,k] is computed once per block per warp (lanes split the diagonal), the zero-diagonal flag of
//   :533-534 is taken there; x = acc * (1/u) differs from the reference's acc / u by at most one rounding.
// Blocks are read once per CTA and sweep (not once per vector): cfg4 moves 2 x 67 MB x (CTAs) through L2 instead
// of 256 x that.
// Measured on B200 (scratch/mrhs_time.py, profiles/r01_ncu_full_summary.md): cfg4 solve 44.9 ms (chain kernel) ->
// 18.9 ms (Q=4, 32 warps) -> 14.2 ms (Q=16, 128 warps).  What bounds it is the dependent sequence of one vector's
// chain -- 2 sweeps x 2048 blocks x 32 two-column steps of shuffle + 3 DFMA (~130 cycles each) plus ~2000 cycles per
// block of staging/coupling/publication -- not bandwidth (60 GB/s of DRAM traffic) and not the FP64 pipe (4 %):
// the next step is the wavefront over consecutive blocks that the blockw-32 one-vector kernel already has.
#include <stdlib.h>

#include "bdilu_mrhs_common.cuh"

namespace bdilu {
namespace mrhs {

template <int WP, int Q, bool FULLW>
__global__ void __launch_bounds__(MAXWARPS * 32)
solve_mrhs_kernel(const double *__restrict__ data, int64_t sys_stride, Dims D, const int *__restrict__ rowbycol,
                  const int *__restrict__ colbyrow, const double *__restrict__ bvec, double *xvec, VecLayout V,
                  int *__restrict__ flags, int groups_per_sys, int ctas_per_sys, Layout SL)
{
    using C = Cfg<WP, Q>;
    constexpr int NP = C::NP, WPP = C::WPP, R = C::R;
    extern __shared__ __align__(128) unsigned char smem[];
    const int nwc = blockDim.x >> 5;
    const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
    const int q = lane % Q, j = lane / Q;
    const int w = D.w, ld = D.ld, N = D.nblocks, nd = D.ndiag;
    const int hw = w >> 1;  // row pairs in a block
    const int Dr = nd > 0 ? nd : 1;

    const int sys = blockIdx.x / ctas_per_sys;
    int grp = (blockIdx.x % ctas_per_sys) * nwc + wic;
    const bool warp_live = grp < groups_per_sys;
    if (!warp_live) grp = groups_per_sys - 1;  // surplus warps shadow the last group (they still release stages)
    const int rhs_raw = grp * R + j;
    const bool live = warp_live && rhs_raw < V.nrhs;
    const int rhs = rhs_raw < V.nrhs ? rhs_raw : V.nrhs - 1;

    const double *sd = data + (int64_t)sys * sys_stride;
    const int *rbc_g = rowbycol + (int64_t)sys * N * w;
    const int *cbr_g = colbyrow + (int64_t)sys * N * w;
    const double *bq = bvec + (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride;
    double *xq = xvec + (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride;

    unsigned char *stages = smem;
    unsigned long long *full = reinterpret_cast<unsigned long long *>(smem + (size_t)NS * SL.stage_bytes);
    unsigned long long *empty = full + NS;
    double *ring = reinterpret_cast<double *>(empty + NS) + (size_t)wic * Dr * R * WPP + (size_t)j * WPP;  // [slot][j][WPP]
    double *dinv = reinterpret_cast<double *>(empty + NS) + (size_t)nwc * Dr * R * WPP + (size_t)wic * WP;
    constexpr int RING_SLOT = R * WPP;

    if (threadIdx.x == 0) {
        for (int s = 0; s < NS; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], nwc);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();

    const int NIT = 2 * N;  // forward blocks 0..N-1, then backward blocks N-1..0
    const bool producer = threadIdx.x == 0;

    // one lane stages iteration `pit`
    auto issue = [&](int pit) {
        const int ps = pit % NS;
        const bool pf = pit < N;
        const int pb = pf ? pit : NIT - 1 - pit;
        if (pit >= NS) mbar_wait(&empty[ps], ((pit / NS) - 1) & 1);
        const int pnd = pf ? min(nd, pb) : min(nd, N - 1 - pb);
        unsigned char *st = stages + (size_t)ps * SL.stage_bytes;
        mbar_expect_tx(&full[ps], SL.blk_bytes + (pf ? 2 : 1) * SL.idx_bytes + pnd * SL.diag_bytes);
        bulk_g2s(st, sd + D.block(pb), SL.blk_bytes, &full[ps]);
        bulk_g2s(st + SL.o_cbr, cbr_g + (int64_t)pb * w, SL.idx_bytes, &full[ps]);
        if (pf) bulk_g2s(st + SL.o_rbc, rbc_g + (int64_t)pb * w, SL.idx_bytes, &full[ps]);
        for (int d = 0; d < pnd; ++d) {
            // forward: sub(d, b-1-d, :) couples block b to y_{b-1-d} (:508-511); backward: sup(d, b, :) to x_{b+1+d} (:524-527)
            const double *src = pf ? sd + D.off_sub + D.diag_chunk(d, pb - 1 - d) : sd + D.off_sup + D.diag_chunk(d, pb);
            bulk_g2s(st + SL.o_diag + d * SL.diag_bytes, src, SL.diag_bytes, &full[ps]);
        }
    };
    if (producer)
        for (int pit = 0; pit < NS - 1 && pit < NIT; ++pit) issue(pit);

    int nan_first = 0x7fffffff;  // 1-based index of the first NaN in b (:500-501)
    int zero_last = -1;          // largest index of a zero U diagonal = the first the reverse loop meets (:533-534)
    double acc[2 * NP], pre[2 * NP];

    // b[rowbycol] of forward block 0
    {
        mbar_wait(&full[0], 0);
        const int *rbc_s = reinterpret_cast<const int *>(stages + SL.o_rbc);
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const int p = i * Q + q;
            pre[2 * i] = pre[2 * i + 1] = 0.0;
            if (FULLW || p < hw) {
                pre[2 * i] = __ldg(bq + rbc_s[2 * p]);
                pre[2 * i + 1] = __ldg(bq + rbc_s[2 * p + 1]);
            }
        }
    }

    int slot = 0;  // ring slot of the current block = bi % Dr
    for (int it = 0; it < NIT; ++it) {
        const bool fwd = it < N;
        const int bi = fwd ? it : NIT - 1 - it;
        const int s = it % NS;
        if (producer && it + NS - 1 < NIT) issue(it + NS - 1);
        __syncwarp();
        mbar_wait(&full[s], (it / NS) & 1);
        const unsigned char *stg = stages + (size_t)s * SL.stage_bytes;
        const double *blk = reinterpret_cast<const double *>(stg);
        const int *cbr_s = reinterpret_cast<const int *>(stg + SL.o_cbr);
        const double *dg_s = reinterpret_cast<const double *>(stg + SL.o_diag);

        // ---- right-hand side of this block: prefetched values minus the coupling terms ----
        if (it != N) {
#pragma unroll
            for (int i = 0; i < 2 * NP; ++i) acc[i] = pre[i];
        }
        if (fwd) {
            const int *rbc_s = reinterpret_cast<const int *>(stg + SL.o_rbc);
#pragma unroll
            for (int i = 0; i < 2 * NP; ++i) {
                if (isnan(acc[i])) {
                    const int p = (i >> 1) * Q + q;
                    nan_first = min(nan_first, bi * w + rbc_s[2 * p + (i & 1)] + 1);
                }
            }
        }
        {
            const int dmax = fwd ? min(nd, bi) : min(nd, N - 1 - bi);
            if (dmax > 0) {
                int ci[2 * NP];
#pragma unroll
                for (int i = 0; i < NP; ++i) {
                    const int p = i * Q + q;
                    const bool ok = FULLW || p < hw;
                    ci[2 * i] = ok ? cbr_s[2 * p] : 0;
                    ci[2 * i + 1] = ok ? cbr_s[2 * p + 1] : 0;
                }
                int sl = slot;
                for (int di = 1; di <= dmax; ++di) {
                    // forward reads block bi-di, backward block bi+di
                    if (fwd) { sl = sl == 0 ? Dr - 1 : sl - 1; } else { sl = sl + 1 == Dr ? 0 : sl + 1; }
                    const double *rg = ring + sl * RING_SLOT;
                    const double *dg = dg_s + (di - 1) * w;
#pragma unroll
                    for (int i = 0; i < 2 * NP; ++i) {
                        const int p = (i >> 1) * Q + q;
                        if (FULLW || p < hw) acc[i] = fma(-dg[ci[i]], rg[ci[i]], acc[i]);
                    }
                }
            }
        }
        if (!fwd) {
            for (int t = lane; t < w; t += 32) {
                const double dgv = blk[t * ld + t];
                if (dgv == 0.0) zero_last = max(zero_last, bi * w + t);
                dinv[t] = 1.0 / dgv;
            }
            __syncwarp();
        }

        // ---- prefetch the next block's right-hand side (b through rowbycol, or y written by this thread) ----
        if (it + 1 < NIT && it + 1 != N) {
            const int nit = it + 1;
            if (nit < N) {
                const int ns = nit % NS;
                mbar_wait(&full[ns], (nit / NS) & 1);
                const int *rbc_n = reinterpret_cast<const int *>(stages + (size_t)ns * SL.stage_bytes + SL.o_rbc);
                const double *bn = bq + (int64_t)nit * w;
#pragma unroll
                for (int i = 0; i < NP; ++i) {
                    const int p = i * Q + q;
                    if (FULLW || p < hw) {
                        pre[2 * i] = __ldg(bn + rbc_n[2 * p]);
                        pre[2 * i + 1] = __ldg(bn + rbc_n[2 * p + 1]);
                    }
                }
            } else {
                const double *yn = xq + (int64_t)(NIT - 1 - nit) * w;
#pragma unroll
                for (int i = 0; i < NP; ++i) {
                    const int p = i * Q + q;
                    if (FULLW || p < hw) {
                        pre[2 * i] = yn[2 * p];
                        pre[2 * i + 1] = yn[2 * p + 1];
                    }
                }
            }
        }

        // ---- triangular solve of the block, two columns per step ----
        if (fwd) {
#pragma unroll
            for (int pk = 0; pk < WP / 2; ++pk) {
                if (FULLW || pk < hw) {
                    const int qo = pk % Q, io = (pk / Q) * 2;
                    const double *c0 = blk + (2 * pk) * ld;
                    const double *c1 = c0 + ld;
                    double y0 = acc[io];
                    double y1 = fma(-c0[2 * pk + 1], y0, acc[io + 1]);  // unit lower L (:503-507)
                    y0 = quad_bcast<Q>(y0, qo);
                    y1 = quad_bcast<Q>(y1, qo);
                    if (q == qo) acc[io + 1] = y1;
#pragma unroll
                    for (int i = pk / Q; i < NP; ++i) {
                        if (i * Q + Q - 1 > pk) {  // some lane of this pair row lies below the pivot pair
                            const int p = i * Q + q;
                            const bool upd = (i * Q > pk || p > pk) && (FULLW || p < hw);
                            if (upd) {
                                const double2 a = *reinterpret_cast<const double2 *>(c0 + 2 * p);
                                const double2 b2 = *reinterpret_cast<const double2 *>(c1 + 2 * p);
                                acc[2 * i] = fma(-a.x, y0, acc[2 * i]);
                                acc[2 * i + 1] = fma(-a.y, y0, acc[2 * i + 1]);
                                acc[2 * i] = fma(-b2.x, y1, acc[2 * i]);
                                acc[2 * i + 1] = fma(-b2.y, y1, acc[2 * i + 1]);
                            }
                        }
                    }
                }
            }
        } else {
#pragma unroll
            for (int pkr = 0; pkr < WP / 2; ++pkr) {
                const int pk = WP / 2 - 1 - pkr;
                if (FULLW || pk < hw) {
                    const int qo = pk % Q, io = (pk / Q) * 2;
                    const double *c0 = blk + (2 * pk) * ld;
                    const double *c1 = c0 + ld;
                    double x1 = acc[io + 1] * dinv[2 * pk + 1];
                    double x0 = fma(-c1[2 * pk], x1, acc[io]) * dinv[2 * pk];  // U[2pk, 2pk+1] (:519-523, 531-532)
                    x0 = quad_bcast<Q>(x0, qo);
                    x1 = quad_bcast<Q>(x1, qo);
                    if (q == qo) {
                        acc[io] = x0;
                        acc[io + 1] = x1;
                    }
#pragma unroll
                    for (int i = 0; i <= pk / Q; ++i) {
                        if (i * Q < pk) {  // some lane of this pair row lies above the pivot pair
                            const int p = i * Q + q;
                            const bool upd = (i * Q + Q - 1 < pk || p < pk);
                            if (upd) {
                                const double2 a = *reinterpret_cast<const double2 *>(c0 + 2 * p);
                                const double2 b2 = *reinterpret_cast<const double2 *>(c1 + 2 * p);
                                acc[2 * i] = fma(-b2.x, x1, acc[2 * i]);
                                acc[2 * i + 1] = fma(-b2.y, x1, acc[2 * i + 1]);
                                acc[2 * i] = fma(-a.x, x0, acc[2 * i]);
                                acc[2 * i + 1] = fma(-a.y, x0, acc[2 * i + 1]);
                            }
                        }
                    }
                }
            }
        }

        // ---- publish: coupling ring (next blocks of this warp) and global x (y between the sweeps, :499) ----
        {
            double *rg = ring + slot * RING_SLOT;
            double *xo = xq + (int64_t)bi * w;
#pragma unroll
            for (int i = 0; i < NP; ++i) {
                const int p = i * Q + q;
                if (FULLW || p < hw) {
                    *reinterpret_cast<double2 *>(rg + 2 * p) = make_double2(acc[2 * i], acc[2 * i + 1]);
                    if (live) {
                        xo[2 * p] = acc[2 * i];
                        xo[2 * p + 1] = acc[2 * i + 1];
                    }
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
        if (fwd) {
            if (it + 1 < N) slot = slot + 1 == Dr ? 0 : slot + 1;
        } else {
            slot = slot == 0 ? Dr - 1 : slot - 1;
        }
    }

    // reference return value: NaN index wins, else N*w + index of the zero diagonal, else 0 (:491-495)
#pragma unroll
    for (int off = Q / 2; off > 0; off >>= 1) nan_first = min(nan_first, __shfl_xor_sync(FULL, nan_first, off, Q));
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) zero_last = max(zero_last, __shfl_xor_sync(FULL, zero_last, off));
    if (live && q == 0 && flags)
        flags[(int64_t)sys * V.nrhs + rhs] = (nan_first != 0x7fffffff) ? nan_first : (zero_last >= 0 ? N * w + zero_last : 0);
}

template <int WP, int Q, bool FULLW>
static cudaError_t launch_t(unsigned grid, int nwc, size_t smem, const double *data, int64_t sys_stride, const Dims &D,
                            const int *rowbycol, const int *colbyrow, const double *b, double *x, const VecLayout &V,
                            int *flags, int gps, int cps, const Layout &SL, cudaStream_t st)
{
    // the attribute belongs to the (function, device) pair: set on every launch (cheap), not once per process
    cudaError_t e = cudaFuncSetAttribute(solve_mrhs_kernel<WP, Q, FULLW>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    solve_mrhs_kernel<WP, Q, FULLW><<<grid, nwc * 32, smem, st>>>(data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, gps,
                                                                cps, SL);
    return cudaGetLastError();
}

}  // namespace mrhs

// Returns true (and launches) when the many-vector path applies; false -> caller falls through to the one-vector kernels.
bool try_launch_solve_mrhs(const double *data, int64_t sys_stride, int nsys, const Dims &D, const int *rowbycol,
                           const int *colbyrow, const double *b, double *x, const VecLayout &V, int *flags,
                           cudaStream_t st, cudaError_t *err)
{
    using namespace mrhs;
    *err = cudaSuccess;
    const int w = D.w;
    const int nsm = sm_count();
    if (g_solve_path != 0 && g_solve_path != 4 && g_solve_path != 6) return false;
    if (w > 64 || w < 4 || (w & 3) || (D.ld & 1) || D.ndiag > MAXD || D.nsat != 0 || D.nblocks < 1) return false;
    if (sys_stride & 1) return false;
    if ((reinterpret_cast<uintptr_t>(data) | reinterpret_cast<uintptr_t>(rowbycol) | reinterpret_cast<uintptr_t>(colbyrow)) & 15)
        return false;
    if (((int64_t)D.nblocks * w) & 3) return false;  // per-system offset of the permutation vectors stays 16-byte aligned
    const int WP = w <= 8 ? 8 : w <= 16 ? 16 : w <= 32 ? 32 : 64;
    const bool fullw = w == WP;
    // Lanes per vector: 4 is the most efficient split (8 vectors share every L/U load); when that leaves most of the
    // GPU idle (cfg4: 256 vectors = 32 groups) the rows of a block are spread over 8 or 16 lanes instead, which shortens
    // each warp's instruction stream per block (the chain over the blocks is what bounds the run time).
    int Q = 4;
    if (fullw) {
        while (Q < 16 && WP >= 4 * Q && (int64_t)nsys * ((V.nrhs + 16 / Q - 1) / (16 / Q)) <= 2 * nsm) Q *= 2;
    }
    const int64_t nvec = (int64_t)nsys * V.nrhs;
    const int64_t groups = (int64_t)nsys * ((V.nrhs + 32 / Q - 1) / (32 / Q));
    // The wavefront form (bdilu_solve_mrhs_wave.cu: one CTA per group, K warps on consecutive blocks) shortens the
    // chain's latency; it pays while the GPU is not full of groups.  With more groups than that the one-warp-per-group
    // form below has the better throughput (blocks staged once for up to 4 groups).
    bool wave = fullw && groups <= 2 * nsm;
    if (g_solve_path == 0) {
        // Measured on B200 (scratch/mrhs_time.py).  blockw 64 has no one-vector wavefront kernel to compete with
        // (cfg4: chain 44.9 ms, one warp per group 13.5 ms, wavefront form 7.7 ms).  At blockw 32 a quad32 CTA finishes
        // one vector per SM in 2.6 ms (N = 8192) where a wavefront group takes 16-29 ms and a one-warp group 34 ms:
        // 1024 vectors take 19.0 ms one by one against 29.3 ms in wavefront groups, so only the one-warp form is used
        // there, from ~14 waves of one-vector CTAs on (2048 vectors: 36 ms against 37.8 ms).  Narrower blocks:
        // the wavefront form matches the generic one-vector kernel at one wave (blockw 16, 64 vectors: 1.19 vs 1.30 ms)
        // and the one-warp form takes over once the one-vector kernels decline the batch (more than 8 waves).
        bool use;
        if (w > 32) use = true;  // even one vector: 5.9 ms in a wavefront group against ~40 ms on the chain kernel
        else if (w == 32) {
            wave = false;
            use = V.nrhs >= 8 && nvec > 14 * nsm;
        } else if (D.nblocks >= 128 && fullw) {
            // blockw 8 / 16, chains long enough to amortise the pipeline fill.  The group kernels also beat the
            // one-vector kernels when every group holds a single live vector (ms, one vector per system):
            //   blockw 16, N = 1024:   1 system  1.05 wavefront / 1.29 one-vector;  296: 1.26 / 2.58;
            //                          600: 3.60 wavefront, 3.27 one warp per group / 6.32;  1200: 6.18, 3.65 / 6.83
            //   blockw 8, N = 512:     64: 0.53 / 0.71;  512: 0.74 / 2.38;  1024: 1.37, 1.49 / 2.45;
            //                          1536: 2.03, 1.58 / 2.41;  2048: 2.58, 1.77 / 2.39;  4096: 4.87, 3.30 / 2.41
            wave = groups <= (w == 8 ? 7 * nsm : 2 * nsm);
            use = wave || groups <= (w == 8 ? 17 * nsm : 16 * nsm) || (V.nrhs >= 8 && nvec > 8 * nsm);
        } else {
            use = V.nrhs >= 8 && nvec > 8 * nsm;
        }
        if (!use) return false;
    }
    if (g_solve_path == 4) wave = false;
    if (g_solve_path == 6) wave = fullw;
    if (wave && try_launch_solve_mrhs_wave(Q, data, sys_stride, nsys, D, rowbycol, colbyrow, b, x, V, flags, st, err))
        return true;
    const int R = 32 / Q;
    const int gps = (V.nrhs + R - 1) / R;
    const int64_t total_groups = (int64_t)nsys * gps;
    int nwc = (int)((total_groups + nsm - 1) / nsm);
    if (nwc > MAXWARPS) nwc = MAXWARPS;
    if (nwc > gps) nwc = gps;
    if (nwc < 1) nwc = 1;
    const int cps = (gps + nwc - 1) / nwc;
    const Layout SL = make_layout(w, D.ld, D.ndiag);
    const int Dr = D.ndiag > 0 ? D.ndiag : 1;
    const size_t smem = (size_t)NS * SL.stage_bytes + 2 * NS * 8 + (size_t)nwc * Dr * R * (WP + 8) * 8 + (size_t)nwc * WP * 8;
    if (smem > 227 * 1024) return false;
    const int64_t grid = (int64_t)nsys * cps;
    if (grid > 0x7fffffff) return false;
#define BDILU_MRHS(WPV, QV, FW)                                                                                      \
    *err = launch_t<WPV, QV, FW>((unsigned)grid, nwc, smem, data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, gps, \
                                 cps, SL, st)
    if (!fullw) {
        if (WP == 8) BDILU_MRHS(8, 4, false);
        else if (WP == 16) BDILU_MRHS(16, 4, false);
        else if (WP == 32) BDILU_MRHS(32, 4, false);
        else BDILU_MRHS(64, 4, false);
    } else if (Q == 4) {
        if (WP == 8) BDILU_MRHS(8, 4, true);
        else if (WP == 16) BDILU_MRHS(16, 4, true);
        else if (WP == 32) BDILU_MRHS(32, 4, true);
        else BDILU_MRHS(64, 4, true);
    } else if (Q == 8) {
        if (WP == 16) BDILU_MRHS(16, 8, true);
        else if (WP == 32) BDILU_MRHS(32, 8, true);
        else BDILU_MRHS(64, 8, true);
    } else {
        if (WP == 32) BDILU_MRHS(32, 16, true);
        else BDILU_MRHS(64, 16, true);
    }
#undef BDILU_MRHS
    count_launch(1);
    return true;
}

}  // namespace bdilu
