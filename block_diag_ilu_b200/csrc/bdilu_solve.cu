// Kernel 2 (general path): ILU_inplace<double>::solve (block_diag_ilu.hpp:486-538).
//
// One G-lane group per (system, right-hand side); the block chain is walked sequentially inside the
// group (forward Ly=b, :502-517, then backward Ux=y, :518-536), lane = row of the current block
// (two rows per lane when 32 < blockw <= 64).  Within a block the triangular solve is a column
// sweep: the lane owning row k broadcasts its finished value, every later row subtracts its L (or
// U) entry times it.  Columns stream from HBM/L2 in chunks of CH columns with one chunk of
// look-ahead that crosses block boundaries, so the loads never sit on the dependency chain.
// y is staged in the x buffer (the reference allocates a scratch y per call, :499); coupling and
// satellite terms read it back from there.  Handles any ndiag, nsat and ld.
#include "bdilu_common.cuh"
#include "bdilu_kernels.h"

namespace bdilu {

constexpr int SOLVE_WARPS = 2;
constexpr int CH = 8;  // columns per streamed chunk

template <int WP>
struct SolveCfg {
    static constexpr int G = WP < 32 ? WP : 32;  // lanes per group
    static constexpr int RPL = WP / G;            // rows per lane
    static constexpr int NCH = WP / CH > 0 ? WP / CH : 1;
    static constexpr int CW = WP < CH ? WP : CH;  // chunk width
};

// Load chunk `ch` (columns ch*CW ..) of block `blk` for my rows.  lower=true: only entries with
// row > col are needed (L); lower=false: only row <= col (U incl. diagonal).
template <int WP, bool LOWER>
__device__ __forceinline__ void load_chunk(double (&dst)[SolveCfg<WP>::RPL][SolveCfg<WP>::CW],
                                           const double *__restrict__ blk, int ch, int r, int w, int ld)
{
    using C = SolveCfg<WP>;
#pragma unroll
    for (int s = 0; s < C::RPL; ++s) {
        const int row = r + s * C::G;
#pragma unroll
        for (int kk = 0; kk < C::CW; ++kk) {
            const int k = ch * C::CW + kk;
            const bool need = row < w && k < w && (LOWER ? (row > k) : (row <= k));
            dst[s][kk] = need ? __ldg(blk + (int64_t)k * ld + row) : 0.0;
        }
    }
}

template <int WP>
__global__ void __launch_bounds__(SOLVE_WARPS * 32)
solve_chain_kernel(const double *__restrict__ data, int64_t sys_stride, int nsys, Dims D,
                   const int *__restrict__ rowbycol, const int *__restrict__ colbyrow,
                   const double *__restrict__ bvec, double *xvec, VecLayout V, int *__restrict__ flags)
{
    using C = SolveCfg<WP>;
    constexpr int G = C::G, RPL = C::RPL, CW = C::CW, NCH = C::NCH;
    constexpr int GPW = 32 / G;
    const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
    const int grp = lane / G, r = lane % G;
    const int w = D.w, ld = D.ld, N = D.nblocks;
    const int64_t total = (int64_t)nsys * V.nrhs;
    int64_t gid = ((int64_t)blockIdx.x * SOLVE_WARPS + wic) * GPW + grp;
    const bool live = gid < total;
    if (!live) gid = total - 1;  // dead groups shadow the last vector and store nothing
    const int sys = (int)(gid / V.nrhs), rhs = (int)(gid % V.nrhs);
    const double *sd = data + (int64_t)sys * sys_stride;
    const int *rbc = rowbycol + (int64_t)sys * N * w;
    const int *cbr = colbyrow + (int64_t)sys * N * w;
    const double *b = bvec + (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride;
    double *x = xvec + (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride;

    int nan_first = 0x7fffffff;  // 1-based index of the first NaN in b (:500-501)
    int zero_last = -1;          // last (= first met by the reverse loop) zero U diagonal (:533-534)

    double cur[RPL][CW], nxt[RPL][CW];

    // ---------------- forward: L y = P b - couplings ----------------
    load_chunk<WP, true>(nxt, sd, 0, r, w, ld);
    for (int bi = 0; bi < N; ++bi) {
        const double *blk = sd + D.block(bi);
        double acc[RPL];
        int ci[RPL];
#pragma unroll
        for (int s = 0; s < RPL; ++s) {
            const int row = r + s * G;
            acc[s] = 0.0;
            ci[s] = 0;
            if (row < w) {
                const int src = bi * w + __ldg(rbc + bi * w + row);
                acc[s] = __ldg(b + src);
                if (isnan(acc[s])) nan_first = min(nan_first, src + 1);
                ci[s] = __ldg(cbr + bi * w + row);
            }
        }
        __syncwarp();  // y of earlier blocks (written by other lanes) is visible below
#pragma unroll
        for (int s = 0; s < RPL; ++s) {
            if (r + s * G < w) {
                const int dmax = min(D.ndiag, bi);
                for (int di = 1; di <= dmax; ++di)
                    acc[s] = fma(-__ldg(sd + D.off_sub + D.diag_chunk(di - 1, bi - di) + ci[s]),
                                 x[(bi - di) * w + ci[s]], acc[s]);
                for (int bci = 0; bci <= D.nsat + bi - N; ++bci)
                    acc[s] = fma(-__ldg(sd + D.off_bot + D.sat_chunk(N + bci - bi - 1, bci) + ci[s]),
                                 x[bci * w + ci[s]], acc[s]);
            }
        }
#pragma unroll
        for (int ch = 0; ch < NCH; ++ch) {
#pragma unroll
            for (int s = 0; s < RPL; ++s)
#pragma unroll
                for (int kk = 0; kk < CW; ++kk) cur[s][kk] = nxt[s][kk];
            if (ch + 1 < NCH) {
                if ((ch + 1) * CW < w) load_chunk<WP, true>(nxt, blk, ch + 1, r, w, ld);
            } else if (bi + 1 < N) {
                load_chunk<WP, true>(nxt, blk + (int64_t)w * ld, 0, r, w, ld);
            }
            if (ch * CW < w) {
#pragma unroll
                for (int kk = 0; kk < CW; ++kk) {
                    const int k = ch * CW + kk;
                    if (k < w) {
                        const double yk = group_bcast<G>(acc[k / G], k % G);
#pragma unroll
                        for (int s = 0; s < RPL; ++s)
                            if (r + s * G > k) acc[s] = fma(-cur[s][kk], yk, acc[s]);
                    }
                }
            }
        }
        if (live) {
#pragma unroll
            for (int s = 0; s < RPL; ++s)
                if (r + s * G < w) x[bi * w + r + s * G] = acc[s];
        }
    }

    // ---------------- backward: U x = y - couplings ----------------
    // chunks are visited in descending column order; chunk index counts from the top
    auto last_chunk = [&]() { return (w - 1) / CW; };
    load_chunk<WP, false>(nxt, sd + D.block(N - 1), last_chunk(), r, w, ld);
    for (int bi = N - 1; bi >= 0; --bi) {
        const double *blk = sd + D.block(bi);
        double acc[RPL];
        __syncwarp();
#pragma unroll
        for (int s = 0; s < RPL; ++s) {
            const int row = r + s * G;
            acc[s] = 0.0;
            if (row < w) {
                acc[s] = x[bi * w + row];
                const int ci = __ldg(cbr + bi * w + row);
                const int dmax = min(N - bi - 1, D.ndiag);
                for (int di = 1; di <= dmax; ++di)
                    acc[s] = fma(-__ldg(sd + D.off_sup + D.diag_chunk(di - 1, bi) + ci),
                                 x[(bi + di) * w + ci], acc[s]);
                for (int sati = D.nsat; sati > bi; --sati)
                    acc[s] = fma(-__ldg(sd + D.off_top + D.sat_chunk(sati - 1, bi) + ci),
                                 x[(N - sati + bi) * w + ci], acc[s]);
            }
        }
#pragma unroll
        for (int chr = 0; chr < NCH; ++chr) {
            const int ch = NCH - 1 - chr;
#pragma unroll
            for (int s = 0; s < RPL; ++s)
#pragma unroll
                for (int kk = 0; kk < CW; ++kk) cur[s][kk] = nxt[s][kk];
            // look-ahead: next lower chunk of this block, or the top chunk of the previous block
            if (ch * CW < w) {
                if (ch > 0) load_chunk<WP, false>(nxt, blk, ch - 1, r, w, ld);
                else if (bi > 0) load_chunk<WP, false>(nxt, blk - (int64_t)w * ld, last_chunk(), r, w, ld);
#pragma unroll
                for (int kr = 0; kr < CW; ++kr) {
                    const int kk = CW - 1 - kr;
                    const int k = ch * CW + kk;
                    if (k < w) {
                        // owner finishes x[k] = acc / U[k][k] (true division, :531-532)
                        if (r == k % G) {
                            const double dg = cur[k / G][kk];
                            if (dg == 0.0) zero_last = max(zero_last, bi * w + k);
                            acc[k / G] = acc[k / G] / dg;
                        }
                        const double xk = group_bcast<G>(acc[k / G], k % G);
#pragma unroll
                        for (int s = 0; s < RPL; ++s)
                            if (r + s * G < k) acc[s] = fma(-cur[s][kk], xk, acc[s]);
                    }
                }
            }
        }
        if (live) {
#pragma unroll
            for (int s = 0; s < RPL; ++s)
                if (r + s * G < w) x[bi * w + r + s * G] = acc[s];
        }
    }

    // reference return value: NaN index wins, else N*w + index of the zero diagonal, else 0
#pragma unroll
    for (int off = G / 2; off > 0; off >>= 1) {
        nan_first = min(nan_first, __shfl_xor_sync(FULL, nan_first, off, G));
        zero_last = max(zero_last, __shfl_xor_sync(FULL, zero_last, off, G));
    }
    if (live && r == 0 && flags)
        flags[gid] = (nan_first != 0x7fffffff) ? nan_first : (zero_last >= 0 ? N * w + zero_last : 0);
}

int g_solve_path = 0;  // test hook: bdilu_debug_force_chain_solve()

cudaError_t launch_solve(const double *data, int64_t sys_stride, int nsys, const Dims &D,
                         const int *rowbycol, const int *colbyrow, const double *b, double *x,
                         const VecLayout &V, int *flags, cudaStream_t st)
{
    const int64_t total = (int64_t)nsys * V.nrhs;
    if (total == 0 || D.nblocks == 0) return cudaSuccess;
    cudaError_t werr;
    // Narrow blocks (bdilu_solve_pf.cu).  Measured on B200 (scratch/pf_time.py; ms, this kernel against the previous
    // choice among the wavefront-group / many-vector / general kernels):
    //   blockw 4, N = 64:    1 system 0.053 / 0.10 (cfg1);  4096: 0.07 / 0.16
    //   blockw 8, N = 512:   1..64 systems 0.39 / 0.54;  256: 0.39 / 0.60;  2048: 0.41 / 1.76;  4096: 0.75 / 2.4 (cfg5);
    //                        4096 x 8 vectors 3.1 / 3.5;  1 system, N = 4096: 2.9 / 4.1
    //   blockw 16, N = 1024: 1..296 systems 1.04 / 1.05-1.25 (cfg2);  1200: 1.18 / 3.6;  16 x 64 vectors 1.09 / 1.26
    // so it is the first choice wherever it fits, except for big batches of many vectors per system, where the many-vector
    // kernels read each block once per group of vectors instead of once per vector, and huge batches of tiny systems.
    // With satellites (nsat > 0, which the wavefront and many-vector kernels decline) it also covers blockw <= 32.
    const bool pf_shape = D.ndiag >= 1 && D.ndiag <= 4 && D.w >= 2 && D.w <= (D.nsat > 0 ? 32 : 16);
    bool pf_first = g_solve_path == 9 || g_solve_path == 11 || g_solve_path == 12;
    if (g_solve_path == 0 && pf_shape) {
        const bool groups_fit = D.nsat == 0 && D.w >= 4 && !(D.w & 3);  // shapes the many-vector kernels accept at all
        const bool many = groups_fit && V.nrhs >= 8 && total > 64 * (int64_t)sm_count();
        pf_first = !many && (D.w > 4 || total <= 20000);
    }
    if (pf_first && try_launch_solve_pf(data, sys_stride, nsys, D, rowbycol, colbyrow, b, x, V, flags, st, &werr)) return werr;
    if (try_launch_solve_mrhs(data, sys_stride, nsys, D, rowbycol, colbyrow, b, x, V, flags, st, &werr)) return werr;
    if (g_solve_path != 1 && try_launch_solve_wave(data, sys_stride, nsys, D, rowbycol, colbyrow, b, x, V, flags, st, &werr))
        return werr;
    if (g_solve_path == 0 && pf_shape && (D.w > 4 || total <= 20000) &&
        try_launch_solve_pf(data, sys_stride, nsys, D, rowbycol, colbyrow, b, x, V, flags, st, &werr))
        return werr;  // faster than the general chain kernel below except for huge batches of tiny systems
#define BDILU_LAUNCH_SOLVE(WP)                                                                       \
    {                                                                                                \
        const int64_t per_cta = (int64_t)SOLVE_WARPS * (32 / SolveCfg<WP>::G);                       \
        const unsigned grid = (unsigned)((total + per_cta - 1) / per_cta);                           \
        solve_chain_kernel<WP><<<grid, SOLVE_WARPS * 32, 0, st>>>(data, sys_stride, nsys, D, rowbycol, \
                                                                  colbyrow, b, x, V, flags);         \
    }
    const int w = D.w;
    if (w <= 2) BDILU_LAUNCH_SOLVE(2)
    else if (w <= 4) BDILU_LAUNCH_SOLVE(4)
    else if (w <= 8) BDILU_LAUNCH_SOLVE(8)
    else if (w <= 16) BDILU_LAUNCH_SOLVE(16)
    else if (w <= 32) BDILU_LAUNCH_SOLVE(32)
    else BDILU_LAUNCH_SOLVE(64)
#undef BDILU_LAUNCH_SOLVE
    count_launch(1);
    return cudaGetLastError();
}

}  // namespace bdilu
