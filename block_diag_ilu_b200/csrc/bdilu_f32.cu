// Real_t = float instantiation of the hot path (block_diag_ilu.hpp:63, 104, 426 take Real_t as a template parameter;
// LAPACK side: sgetrf_, anyode_blas_lapack.hpp:56).  Same packed layout with float elements, same int32 index arrays,
// same info / flag encodings as the double entry points (include/bdilu.h).
//
//   K1f  factor_f32_kernel<WP>      one G-lane group per block (G = blockw rounded up to a power of two, <= 32), lane r
//                                   keeps ROW r in registers, implicit pivoting exactly as bdilu_factor.cu (rows never
//                                   move, every lane tracks the position its row holds in LAPACK's swapped order);
//                                   the pivot search is two group reductions on the integer image of |a| (a float's
//                                   modulus orders like its bit pattern), the pivot row travels by shuffle.
//        factor_wide_f32_kernel     32 < blockw <= 64: one 64-thread CTA per block.
//   K2f  solve_chain_f32_kernel<WP> ILU_inplace<float>::solve (:486-538): the chain walked by one lane group per
//                                   (system, vector), columns streamed with one chunk of look-ahead; any ndiag / nsat / ld.
//   K3f  dot_vec_f32_kernel         BlockDiagMatrix<float>::dot_vec (:295-326), one thread per output row.
//
// These are the straightforward forms: the float path is half the bytes of the double one and is not what BASELINE's
// metric is quoted on; the tuned wavefront kernels stay double-only.
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/bdilu.h"
#include "bdilu_common.cuh"
#include "bdilu_kernels.h"

namespace bdilu {
namespace f32 {

constexpr int FACTOR_WARPS = 4;

template <int G>
__device__ __forceinline__ float bcastf(float v, int src)
{
    return __shfl_sync(FULL, v, src, G);
}
template <int G>
__device__ __forceinline__ int gmax_i(int v)
{
    if (G == 32) return __reduce_max_sync(FULL, v);
#pragma unroll
    for (int off = G / 2; off > 0; off >>= 1) v = max(v, __shfl_xor_sync(FULL, v, off, G));
    return v;
}
template <int G>
__device__ __forceinline__ unsigned gmin_u(unsigned v)
{
    if (G == 32) return __reduce_min_sync(FULL, v);
#pragma unroll
    for (int off = G / 2; off > 0; off >>= 1) v = min(v, __shfl_xor_sync(FULL, v, off, G));
    return v;
}

// keep the failing block with the lowest index (serial loop order of the reference, :469-470)
__device__ __forceinline__ void record_info(int *info, int sys, int bi, int blk_info)
{
    const int key = (bi << 7) | blk_info;
    int old = atomicCAS(&info[sys], 0, key);
    while (old != 0 && (old >> 7) > bi) {
        const int assumed = old;
        old = atomicCAS(&info[sys], assumed, key);
        if (old == assumed) break;
    }
}

template <int WP>
__global__ void __launch_bounds__(FACTOR_WARPS * 32)
factor_f32_kernel(float *data, int64_t sys_stride, int nsys, Dims D, int *__restrict__ ipiv, int *__restrict__ rowbycol,
                  int *__restrict__ colbyrow, int *__restrict__ info)
{
    constexpr int G = WP;
    constexpr int BPW = 32 / G;
    const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
    const int grp = lane / G, r = lane % G;
    const int w = D.w, ld = D.ld;
    const int64_t total = (int64_t)nsys * D.nblocks;
    int64_t gb = ((int64_t)blockIdx.x * FACTOR_WARPS + wic) * BPW + grp;
    const bool live = gb < total;  // dead groups shadow the last block and store nothing
    if (!live) gb = total - 1;
    const int sys = (int)(gb / D.nblocks), bi = (int)(gb % D.nblocks);
    float *sysdata = data + (int64_t)sys * sys_stride;
    float *blk = sysdata + D.block(bi);
    const bool rowok = r < w;

    // row r; rows / columns beyond w are padded with the identity (they pivot on themselves in the steps k >= w)
    float a[WP];
#pragma unroll
    for (int c = 0; c < WP; ++c) a[c] = (rowok && c < w) ? __ldcs(blk + (int64_t)c * ld + r) : ((c == r) ? 1.0f : 0.0f);

    int pos = r;
    bool active = true;
    int my_ipiv = r + 1, my_info = 0;
    float my_diag = 1.0f;
#pragma unroll
    for (int k = 0; k < WP; ++k) {
        // idamax over the rows not yet chosen, ties -> lowest current position
        const int key = active ? (__float_as_int(a[k]) & 0x7fffffff) : -1;
        const int mkey = gmax_i<G>(key);
        const bool c1 = (key == mkey);
        const unsigned q = gmin_u<G>(c1 ? (unsigned)pos : 0xffffu);
        const bool is_p = c1 && ((unsigned)pos == q);
        const unsigned bal = __ballot_sync(FULL, is_p);
        const int pl = __ffs(G == 32 ? bal : ((bal >> (grp * G)) & ((1u << (G & 31)) - 1u))) - 1;
        if (r == k) my_ipiv = (int)q + 1;
        if (pos == k) pos = (int)q;
        if (is_p) { pos = k; active = false; }
        float u[WP];
#pragma unroll
        for (int j = k; j < WP; ++j) u[j] = bcastf<G>(a[j], pl);
        const float piv = u[k];
        if (r == k) my_diag = piv;
        if (piv == 0.0f) {
            if (my_info == 0) my_info = k + 1;  // sgetf2 leaves the column unscaled and records info
        } else if (active) {
            const float l = a[k] / piv;
            a[k] = l;
#pragma unroll
            for (int j = k + 1; j < WP; ++j) a[j] = fmaf(-l, u[j], a[j]);
        }
    }
    if (my_info > w) my_info = 0;  // padded steps never fail; defensive
    if (live && rowok) {
#pragma unroll
        for (int c = 0; c < WP; ++c)
            if (c < w) blk[(int64_t)c * ld + pos] = a[c];
        const int64_t ib = ((int64_t)sys * D.nblocks + bi) * w;
        ipiv[ib + r] = my_ipiv;
        rowbycol[ib + pos] = r;
        colbyrow[ib + r] = pos;
        for (int di = 0; di < D.ndiag && bi + di < D.nblocks - 1; ++di) {  // sub(di,bi,c) /= U[c][c]  (:471-475)
            float *s = sysdata + D.off_sub + D.diag_chunk(di, bi) + r;
            *s = *s / my_diag;
        }
        if (bi < D.nsat)
            for (int sati = bi; sati < D.nsat; ++sati) {  // bot(sati,bi,c) /= U[c][c]  (:476-479)
                float *s = sysdata + D.off_bot + D.sat_chunk(sati, bi) + r;
                *s = *s / my_diag;
            }
        if (my_info != 0 && r == 0) record_info(info, sys, bi, my_info);
    }
}

// 32 < blockw <= 64: thread r of a 64-thread CTA keeps row r; the search and the pivot row cross the two warps through
// shared memory.
__global__ void __launch_bounds__(64)
factor_wide_f32_kernel(float *__restrict__ data, int64_t sys_stride, int nsys, Dims D, int *__restrict__ ipiv,
                       int *__restrict__ rowbycol, int *__restrict__ colbyrow, int *__restrict__ info)
{
    constexpr int WP = 64;
    __shared__ __align__(16) float urow[2][WP];
    __shared__ int cand_key[2][2], cand_pos[2][2], cand_row[2][2];
    const int r = threadIdx.x, lane = r & 31, wid = r >> 5;
    const int w = D.w, ld = D.ld;
    const int64_t gb = blockIdx.x;
    const int sys = (int)(gb / D.nblocks), bi = (int)(gb % D.nblocks);
    float *sysdata = data + (int64_t)sys * sys_stride;
    float *blk = sysdata + D.block(bi);
    const bool rowok = r < w;
    float a[WP];
#pragma unroll
    for (int c = 0; c < WP; ++c) a[c] = (rowok && c < w) ? __ldcs(blk + (int64_t)c * ld + r) : ((c == r) ? 1.0f : 0.0f);
    int pos = r;
    bool active = rowok;
    int my_ipiv = r + 1, my_info = 0;
    float my_diag = 1.0f;
#pragma unroll
    for (int k = 0; k < WP; ++k) {
        if (k < w) {
            const int key = active ? (__float_as_int(a[k]) & 0x7fffffff) : -1;
            const int mkey = __reduce_max_sync(FULL, key);
            const bool c1 = (key == mkey);
            const unsigned q = __reduce_min_sync(FULL, c1 ? (unsigned)pos : 0xffffu);
            if (c1 && (unsigned)pos == q) { cand_key[k & 1][wid] = key; cand_pos[k & 1][wid] = pos; cand_row[k & 1][wid] = r; }
            __syncthreads();
            const int k0 = cand_key[k & 1][0], k1 = cand_key[k & 1][1];
            const int p0 = cand_pos[k & 1][0], p1 = cand_pos[k & 1][1];
            const bool take1 = (k1 > k0) || (k1 == k0 && p1 < p0);
            const int p = take1 ? cand_row[k & 1][1] : cand_row[k & 1][0];
            const int qq = take1 ? p1 : p0;
            if (r == k) my_ipiv = qq + 1;
            if (pos == k) pos = qq;
            if (r == p) { pos = k; active = false; }
            if (r == p) {
#pragma unroll
                for (int j = k; j < WP; ++j) urow[k & 1][j] = a[j];
            }
            __syncthreads();
            const float piv = urow[k & 1][k];
            if (r == k) my_diag = piv;
            if (piv == 0.0f) {
                if (my_info == 0) my_info = k + 1;
            } else if (active) {
                const float l = a[k] / piv;
                a[k] = l;
#pragma unroll
                for (int j = k + 1; j < WP; ++j) a[j] = fmaf(-l, urow[k & 1][j], a[j]);
            }
        }
    }
    (void)lane;
    if (rowok) {
#pragma unroll
        for (int c = 0; c < WP; ++c)
            if (c < w) blk[(int64_t)c * ld + pos] = a[c];
        const int64_t ib = ((int64_t)sys * D.nblocks + bi) * w;
        ipiv[ib + r] = my_ipiv;
        rowbycol[ib + pos] = r;
        colbyrow[ib + r] = pos;
        for (int di = 0; di < D.ndiag && bi + di < D.nblocks - 1; ++di) {
            float *s = sysdata + D.off_sub + D.diag_chunk(di, bi) + r;
            *s = *s / my_diag;
        }
        if (bi < D.nsat)
            for (int sati = bi; sati < D.nsat; ++sati) {
                float *s = sysdata + D.off_bot + D.sat_chunk(sati, bi) + r;
                *s = *s / my_diag;
            }
        if (my_info != 0 && r == 0) record_info(info, sys, bi, my_info);
    }
}

// ---- K2f: chain solve ---------------------------------------------------------------------------------------------
constexpr int SOLVE_WARPS = 2;
constexpr int CH = 8;  // columns per streamed chunk

template <int WP>
struct Cfg {
    static constexpr int G = WP < 32 ? WP : 32;  // lanes per (system, vector)
    static constexpr int RPL = WP / G;            // rows per lane
    static constexpr int NCH = WP / CH > 0 ? WP / CH : 1;
    static constexpr int CW = WP < CH ? WP : CH;
};

// chunk `ch` of a block for my rows; LOWER: entries with row > col (L), else row <= col (U with its diagonal)
template <int WP, bool LOWER>
__device__ __forceinline__ void load_chunk(float (&dst)[Cfg<WP>::RPL][Cfg<WP>::CW], const float *__restrict__ blk, int ch,
                                           int r, int w, int ld)
{
    using C = Cfg<WP>;
#pragma unroll
    for (int s = 0; s < C::RPL; ++s) {
        const int row = r + s * C::G;
#pragma unroll
        for (int kk = 0; kk < C::CW; ++kk) {
            const int k = ch * C::CW + kk;
            const bool need = row < w && k < w && (LOWER ? (row > k) : (row <= k));
            dst[s][kk] = need ? __ldg(blk + (int64_t)k * ld + row) : 0.0f;
        }
    }
}

template <int WP>
__global__ void __launch_bounds__(SOLVE_WARPS * 32)
solve_chain_f32_kernel(const float *__restrict__ data, int64_t sys_stride, int nsys, Dims D,
                       const int *__restrict__ rowbycol, const int *__restrict__ colbyrow,
                       const float *__restrict__ bvec, float *xvec, VecLayout V, int *__restrict__ flags)
{
    using C = Cfg<WP>;
    constexpr int G = C::G, RPL = C::RPL, CW = C::CW, NCH = C::NCH;
    constexpr int GPW = 32 / G;
    const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
    const int grp = lane / G, r = lane % G;
    const int w = D.w, ld = D.ld, N = D.nblocks;
    const int64_t total = (int64_t)nsys * V.nrhs;
    int64_t gid = ((int64_t)blockIdx.x * SOLVE_WARPS + wic) * GPW + grp;
    const bool live = gid < total;
    if (!live) gid = total - 1;
    const int sys = (int)(gid / V.nrhs), rhs = (int)(gid % V.nrhs);
    const float *sd = data + (int64_t)sys * sys_stride;
    const int *rbc = rowbycol + (int64_t)sys * N * w;
    const int *cbr = colbyrow + (int64_t)sys * N * w;
    const float *b = bvec + (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride;
    float *x = xvec + (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride;

    int nan_first = 0x7fffffff;  // 1-based index of the first NaN in b (:500-501)
    int zero_last = -1;          // the zero U diagonal the reverse loop meets last (:533-534)
    float cur[RPL][CW], nxt[RPL][CW];

    // forward: L y = P b - couplings (:502-517); y is staged in the x buffer
    load_chunk<WP, true>(nxt, sd, 0, r, w, ld);
    for (int bi = 0; bi < N; ++bi) {
        const float *blk = sd + D.block(bi);
        float acc[RPL];
        int ci[RPL];
#pragma unroll
        for (int s = 0; s < RPL; ++s) {
            const int row = r + s * G;
            acc[s] = 0.0f;
            ci[s] = 0;
            if (row < w) {
                const int src = bi * w + __ldg(rbc + bi * w + row);
                acc[s] = __ldg(b + src);
                if (isnan(acc[s])) nan_first = min(nan_first, src + 1);
                ci[s] = __ldg(cbr + bi * w + row);
            }
        }
        __syncwarp();  // y of earlier blocks, written by other lanes, is visible below
#pragma unroll
        for (int s = 0; s < RPL; ++s) {
            if (r + s * G < w) {
                const int dmax = min(D.ndiag, bi);
                for (int di = 1; di <= dmax; ++di)
                    acc[s] = fmaf(-__ldg(sd + D.off_sub + D.diag_chunk(di - 1, bi - di) + ci[s]), x[(bi - di) * w + ci[s]],
                                  acc[s]);
                for (int bci = 0; bci <= D.nsat + bi - N; ++bci)
                    acc[s] = fmaf(-__ldg(sd + D.off_bot + D.sat_chunk(N + bci - bi - 1, bci) + ci[s]), x[bci * w + ci[s]],
                                  acc[s]);
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
                        const float yk = bcastf<G>(acc[k / G], k % G);
#pragma unroll
                        for (int s = 0; s < RPL; ++s)
                            if (r + s * G > k) acc[s] = fmaf(-cur[s][kk], yk, acc[s]);
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

    // backward: U x = y - couplings (:518-536); chunks in descending column order
    const int last_chunk = (w - 1) / CW;
    load_chunk<WP, false>(nxt, sd + D.block(N - 1), last_chunk, r, w, ld);
    for (int bi = N - 1; bi >= 0; --bi) {
        const float *blk = sd + D.block(bi);
        float acc[RPL];
        __syncwarp();
#pragma unroll
        for (int s = 0; s < RPL; ++s) {
            const int row = r + s * G;
            acc[s] = 0.0f;
            if (row < w) {
                acc[s] = x[bi * w + row];
                const int ci = __ldg(cbr + bi * w + row);
                const int dmax = min(N - bi - 1, D.ndiag);
                for (int di = 1; di <= dmax; ++di)
                    acc[s] = fmaf(-__ldg(sd + D.off_sup + D.diag_chunk(di - 1, bi) + ci), x[(bi + di) * w + ci], acc[s]);
                for (int sati = D.nsat; sati > bi; --sati)
                    acc[s] = fmaf(-__ldg(sd + D.off_top + D.sat_chunk(sati - 1, bi) + ci), x[(N - sati + bi) * w + ci],
                                  acc[s]);
            }
        }
#pragma unroll
        for (int chr = 0; chr < NCH; ++chr) {
            const int ch = NCH - 1 - chr;
#pragma unroll
            for (int s = 0; s < RPL; ++s)
#pragma unroll
                for (int kk = 0; kk < CW; ++kk) cur[s][kk] = nxt[s][kk];
            if (ch * CW < w) {
                if (ch > 0) load_chunk<WP, false>(nxt, blk, ch - 1, r, w, ld);
                else if (bi > 0) load_chunk<WP, false>(nxt, blk - (int64_t)w * ld, last_chunk, r, w, ld);
#pragma unroll
                for (int kr = 0; kr < CW; ++kr) {
                    const int kk = CW - 1 - kr;
                    const int k = ch * CW + kk;
                    if (k < w) {
                        if (r == k % G) {  // the owner finishes x[k] = acc / U[k][k] (true division, :531-532)
                            const float dg = cur[k / G][kk];
                            if (dg == 0.0f) zero_last = max(zero_last, bi * w + k);
                            acc[k / G] = acc[k / G] / dg;
                        }
                        const float xk = bcastf<G>(acc[k / G], k % G);
#pragma unroll
                        for (int s = 0; s < RPL; ++s)
                            if (r + s * G < k) acc[s] = fmaf(-cur[s][kk], xk, acc[s]);
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
#pragma unroll
    for (int off = G / 2; off > 0; off >>= 1) {
        nan_first = min(nan_first, __shfl_xor_sync(FULL, nan_first, off, G));
        zero_last = max(zero_last, __shfl_xor_sync(FULL, zero_last, off, G));
    }
    if (live && r == 0 && flags)
        flags[gid] = (nan_first != 0x7fffffff) ? nan_first : (zero_last >= 0 ? N * w + zero_last : 0);
}

// ---- K3f: matvec on the unfactored matrix, scale_diag_add ------------------------------------------------------------
constexpr int MV_THREADS = 128;

__global__ void __launch_bounds__(MV_THREADS)
dot_vec_f32_kernel(const float *__restrict__ data, int64_t sys_stride, Dims D, const float *__restrict__ vec,
                   float *__restrict__ out, VecLayout V)
{
    const int sys = blockIdx.z, rhs = blockIdx.y;
    const int w = D.w, N = D.nblocks, n = N * w;
    const int i = blockIdx.x * MV_THREADS + threadIdx.x;
    if (i >= n) return;
    const float *sd = data + (int64_t)sys * sys_stride;
    const int64_t voff = (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride;
    const float *v = vec + voff;
    const int bi = i / w, row = i - bi * w;
    const float *col = sd + D.block(bi) + row;
    float acc = 0.0f;
#pragma unroll 8
    for (int c = 0; c < w; ++c) acc = fmaf(__ldcs(col + (int64_t)c * D.ld), __ldg(v + bi * w + c), acc);
    for (int di = 0; di < D.ndiag; ++di) {
        if (bi + di + 1 < N) acc = fmaf(__ldcs(sd + D.off_sup + D.diag_chunk(di, bi) + row), __ldg(v + (bi + di + 1) * w + row), acc);
        if (bi - di - 1 >= 0)
            acc = fmaf(__ldcs(sd + D.off_sub + D.diag_chunk(di, bi - di - 1) + row), __ldg(v + (bi - di - 1) * w + row), acc);
    }
    for (int sati = 0; sati < D.nsat; ++sati) {
        if (bi <= sati) acc = fmaf(__ldcs(sd + D.off_top + D.sat_chunk(sati, bi) + row), __ldg(v + (N - 1 - sati + bi) * w + row), acc);
        const int bj = bi - N + 1 + sati;
        if (bj >= 0) acc = fmaf(__ldcs(sd + D.off_bot + D.sat_chunk(sati, bj) + row), __ldg(v + bj * w + row), acc);
    }
    out[voff + i] = acc;
}

__global__ void __launch_bounds__(256)
scale_diag_add_f32_kernel(float *dst, const float *src, int64_t sys_stride, Dims D, float scale, float diag_add)
{
    const int64_t base = (int64_t)blockIdx.y * sys_stride;
    const int64_t per_block = (int64_t)D.w * D.ld;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < D.ndata; i += (int64_t)gridDim.x * blockDim.x) {
        float v = __fmul_rn(scale, src[base + i]);  // product and sum rounded separately, like the reference's two statements
        if (i < D.off_sub) {
            const int64_t in_blk = i % per_block;
            if (in_blk / D.ld == in_blk % D.ld) v = __fadd_rn(v, diag_add);
        }
        dst[base + i] = v;
    }
}

constexpr int MAX_GRID_YZ = 65535;

}  // namespace f32
}  // namespace bdilu

// ---- launchers -------------------------------------------------------------------------------------------------------
using namespace bdilu;

namespace {

cudaError_t launch_factor_f32(float *data, int64_t sys_stride, int nsys, const Dims &D, int *ipiv, int *rowbycol,
                              int *colbyrow, int *info, cudaStream_t st)
{
    const int64_t total = (int64_t)nsys * D.nblocks;
    if (total == 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(info, 0, sizeof(int) * (size_t)nsys, st);
    if (e != cudaSuccess) return e;
#define F32_FACTOR(WP)                                                                                             \
    {                                                                                                              \
        const int64_t per_cta = (int64_t)f32::FACTOR_WARPS * (32 / WP);                                            \
        f32::factor_f32_kernel<WP><<<(unsigned)((total + per_cta - 1) / per_cta), f32::FACTOR_WARPS * 32, 0, st>>>( \
            data, sys_stride, nsys, D, ipiv, rowbycol, colbyrow, info);                                            \
    }
    const int w = D.w;
    if (w <= 2) F32_FACTOR(2)
    else if (w <= 4) F32_FACTOR(4)
    else if (w <= 8) F32_FACTOR(8)
    else if (w <= 16) F32_FACTOR(16)
    else if (w <= 32) F32_FACTOR(32)
    else f32::factor_wide_f32_kernel<<<(unsigned)total, 64, 0, st>>>(data, sys_stride, nsys, D, ipiv, rowbycol, colbyrow, info);
#undef F32_FACTOR
    count_launch(1);
    return cudaGetLastError();
}

cudaError_t launch_solve_f32(const float *data, int64_t sys_stride, int nsys, const Dims &D, const int *rowbycol,
                             const int *colbyrow, const float *b, float *x, const VecLayout &V, int *flags, cudaStream_t st)
{
    const int64_t total = (int64_t)nsys * V.nrhs;
    if (total == 0 || D.nblocks == 0) return cudaSuccess;
#define F32_SOLVE(WP)                                                                                                  \
    {                                                                                                                  \
        const int64_t per_cta = (int64_t)f32::SOLVE_WARPS * (32 / f32::Cfg<WP>::G);                                    \
        f32::solve_chain_f32_kernel<WP><<<(unsigned)((total + per_cta - 1) / per_cta), f32::SOLVE_WARPS * 32, 0, st>>>( \
            data, sys_stride, nsys, D, rowbycol, colbyrow, b, x, V, flags);                                            \
    }
    const int w = D.w;
    if (w <= 2) F32_SOLVE(2)
    else if (w <= 4) F32_SOLVE(4)
    else if (w <= 8) F32_SOLVE(8)
    else if (w <= 16) F32_SOLVE(16)
    else if (w <= 32) F32_SOLVE(32)
    else F32_SOLVE(64)
#undef F32_SOLVE
    count_launch(1);
    return cudaGetLastError();
}

#define CU32(call)                                                    \
    do {                                                              \
        cudaError_t e_ = (call);                                      \
        if (e_ != cudaSuccess) return api_cuda_fail(e_, #call);       \
    } while (0)

}  // namespace

// ---- C ABI (include/bdilu.h, "Real_t = float") --------------------------------------------------------------------------
extern "C" {

int bdilu_factor_f32(float *data, int64_t sys_stride, int nsys, int nblocks, int blockw, int ndiag, int nsat, int ld,
                     int *ipiv, int *rowbycol, int *colbyrow, int *info, void *stream)
{
    Dims D;
    int rc = api_check_dims(nsys, sys_stride, nblocks, blockw, ndiag, nsat, ld, D, 4);
    if (rc) return rc;
    if (nsys == 0) return BDILU_OK;
    if (!data || !ipiv || !rowbycol || !colbyrow || !info) return api_fail(BDILU_ERR_ARG, "null pointer");
    CU32(launch_factor_f32(data, sys_stride, nsys, D, ipiv, rowbycol, colbyrow, info, (cudaStream_t)stream));
    return BDILU_OK;
}

int bdilu_solve_f32(const float *data, int64_t sys_stride, int nsys, int nblocks, int blockw, int ndiag, int nsat, int ld,
                    const int *rowbycol, const int *colbyrow, const float *b, float *x, int nrhs, int64_t rhs_stride,
                    int64_t sys_rhs_stride, int *flags, void *stream)
{
    Dims D;
    int rc = api_check_dims(nsys, sys_stride, nblocks, blockw, ndiag, nsat, ld, D, 4);
    if (rc) return rc;
    if ((rc = api_check_vecs(D, nsys, nrhs, rhs_stride, sys_rhs_stride))) return rc;
    if (nsys == 0 || nrhs == 0) return BDILU_OK;
    if (!data || !rowbycol || !colbyrow || !b || !x) return api_fail(BDILU_ERR_ARG, "null pointer");
    if (b == x) return api_fail(BDILU_ERR_ARG, "b and x must not alias");
    VecLayout V{nrhs, rhs_stride, sys_rhs_stride};
    if (nblocks == 0 && flags) CU32(cudaMemsetAsync(flags, 0, sizeof(int) * (size_t)nsys * nrhs, (cudaStream_t)stream));
    CU32(launch_solve_f32(data, sys_stride, nsys, D, rowbycol, colbyrow, b, x, V, flags, (cudaStream_t)stream));
    return BDILU_OK;
}

int bdilu_factor_solve_f32(float *data, int64_t sys_stride, int nsys, int nblocks, int blockw, int ndiag, int nsat, int ld,
                           int *ipiv, int *rowbycol, int *colbyrow, int *info, const float *b, float *x, int nrhs,
                           int64_t rhs_stride, int64_t sys_rhs_stride, int *flags, void *stream)
{
    int rc = bdilu_factor_f32(data, sys_stride, nsys, nblocks, blockw, ndiag, nsat, ld, ipiv, rowbycol, colbyrow, info, stream);
    if (rc) return rc;
    return bdilu_solve_f32(data, sys_stride, nsys, nblocks, blockw, ndiag, nsat, ld, rowbycol, colbyrow, b, x, nrhs,
                           rhs_stride, sys_rhs_stride, flags, stream);
}

int bdilu_dot_vec_f32(const float *data, int64_t sys_stride, int nsys, int nblocks, int blockw, int ndiag, int nsat, int ld,
                      const float *vec, float *out, int nrhs, int64_t rhs_stride, int64_t sys_rhs_stride, void *stream)
{
    Dims D;
    int rc = api_check_dims(nsys, sys_stride, nblocks, blockw, ndiag, nsat, ld, D, 4);
    if (rc) return rc;
    if ((rc = api_check_vecs(D, nsys, nrhs, rhs_stride, sys_rhs_stride))) return rc;
    const int n = nblocks * blockw;
    if (nsys == 0 || nrhs == 0 || n == 0) return BDILU_OK;
    if (!data || !vec || !out) return api_fail(BDILU_ERR_ARG, "null pointer");
    if (nrhs > f32::MAX_GRID_YZ) return api_fail(BDILU_ERR_UNSUPPORTED, "bdilu_dot_vec_f32: more than 65535 vectors per system");
    VecLayout V{nrhs, rhs_stride, sys_rhs_stride};
    for (int s0 = 0; s0 < nsys; s0 += f32::MAX_GRID_YZ) {
        const int ns = nsys - s0 < f32::MAX_GRID_YZ ? nsys - s0 : f32::MAX_GRID_YZ;
        dim3 grid((n + f32::MV_THREADS - 1) / f32::MV_THREADS, nrhs, ns);
        f32::dot_vec_f32_kernel<<<grid, f32::MV_THREADS, 0, (cudaStream_t)stream>>>(
            data + (int64_t)s0 * sys_stride, sys_stride, D, vec + (int64_t)s0 * sys_rhs_stride,
            out + (int64_t)s0 * sys_rhs_stride, V);
        count_launch(1);
    }
    CU32(cudaGetLastError());
    return BDILU_OK;
}

int bdilu_scale_diag_add_f32(float *dst, const float *src, int64_t sys_stride, int nsys, int nblocks, int blockw, int ndiag,
                             int nsat, int ld, float scale, float diag_add, void *stream)
{
    Dims D;
    int rc = api_check_dims(nsys, sys_stride, nblocks, blockw, ndiag, nsat, ld, D, 4);
    if (rc) return rc;
    if (nsys == 0 || D.ndata == 0) return BDILU_OK;
    if (!dst || !src) return api_fail(BDILU_ERR_ARG, "null pointer");
    const int64_t want = (D.ndata + 255) / 256;
    for (int s0 = 0; s0 < nsys; s0 += f32::MAX_GRID_YZ) {
        const int ns = nsys - s0 < f32::MAX_GRID_YZ ? nsys - s0 : f32::MAX_GRID_YZ;
        dim3 grid((unsigned)(want < sm_count() * 8 ? want : sm_count() * 8), ns);
        f32::scale_diag_add_f32_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
            dst + (int64_t)s0 * sys_stride, src + (int64_t)s0 * sys_stride, sys_stride, D, scale, diag_add);
        count_launch(1);
    }
    CU32(cudaGetLastError());
    return BDILU_OK;
}

}  // extern "C"

// ---- host entry points (what ILU_inplace<float> / BlockDiagMatrix<float> of the header shim bind) ----------------------
struct bdilu_ilu_f32 {
    int nsys, nblocks, blockw, ndiag, nsat, ld, device;
    int64_t sys_stride;
    float *d_data = nullptr, *d_b = nullptr, *d_x = nullptr;
    int *d_idx = nullptr, *d_info = nullptr, *d_flags = nullptr;  // d_idx: ipiv | rowbycol | colbyrow
    size_t vec_cap = 0, flag_cap = 0;
    cudaStream_t st = nullptr;
};

namespace {

struct DevGuard {
    int prev = -1;
    bool ok = true;
    explicit DevGuard(int dev)
    {
        if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; ok = false; return; }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~DevGuard()
    {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

void free_handle_f32(bdilu_ilu_f32 *h)
{
    if (!h) return;
    DevGuard g(h->device);
    if (h->st) cudaStreamSynchronize(h->st);
    cudaFree(h->d_data); cudaFree(h->d_b); cudaFree(h->d_x); cudaFree(h->d_idx); cudaFree(h->d_info); cudaFree(h->d_flags);
    if (h->st) cudaStreamDestroy(h->st);
    cudaGetLastError();
    delete h;
}

}  // namespace

extern "C" {

int bdilu_ilu_create_f32(bdilu_ilu_f32 **out, float *host_data, int64_t sys_stride, int nsys, int nblocks, int blockw,
                         int ndiag, int nsat, int ld, int copy_back, int *ipiv, int *rowbycol, int *colbyrow, int *info)
{
    if (!out) return api_fail(BDILU_ERR_ARG, "out is null");
    *out = nullptr;
    if (nsys < 1 || !host_data) return api_fail(BDILU_ERR_ARG, "nsys < 1 or null data");
    Dims D;
    int rc = api_check_dims(nsys, nsys == 1 ? 0 : sys_stride, nblocks, blockw, ndiag, nsat, ld, D, 4);
    if (rc) return rc;
    if (nsys == 1) sys_stride = D.ndata;
    if (bdilu_device_count() < 1) return api_fail(BDILU_ERR_CUDA, "no CUDA device: libbdilu_b200 has no CPU fallback");
    bdilu_ilu_f32 *h = new (std::nothrow) bdilu_ilu_f32;
    if (!h) return api_fail(BDILU_ERR_ARG, "out of host memory");
    h->nsys = nsys; h->nblocks = nblocks; h->blockw = blockw; h->ndiag = ndiag; h->nsat = nsat; h->ld = ld;
    h->sys_stride = sys_stride;
    h->device = 0;
    const size_t nbytes = sizeof(float) * (size_t)((int64_t)(nsys - 1) * sys_stride + D.ndata);
    const size_t nidx = (size_t)nsys * nblocks * blockw;
#define CUC32(call)                                   \
    do {                                              \
        cudaError_t e_ = (call);                      \
        if (e_ != cudaSuccess) {                      \
            free_handle_f32(h);                       \
            return api_cuda_fail(e_, #call);          \
        }                                             \
    } while (0)
    CUC32(cudaGetDevice(&h->device));
    CUC32(cudaStreamCreateWithFlags(&h->st, cudaStreamNonBlocking));
    CUC32(cudaMalloc(&h->d_data, nbytes ? nbytes : 8));
    CUC32(cudaMalloc(&h->d_idx, sizeof(int) * (nidx ? 3 * nidx : 1)));
    CUC32(cudaMalloc(&h->d_info, sizeof(int) * (size_t)nsys));
    CUC32(cudaMemcpyAsync(h->d_data, host_data, nbytes, cudaMemcpyHostToDevice, h->st));
    rc = bdilu_factor_f32(h->d_data, sys_stride, nsys, nblocks, blockw, ndiag, nsat, ld, h->d_idx, h->d_idx + nidx,
                          h->d_idx + 2 * nidx, h->d_info, h->st);
    if (rc) { free_handle_f32(h); return rc; }
    if (copy_back) CUC32(cudaMemcpyAsync(host_data, h->d_data, nbytes, cudaMemcpyDeviceToHost, h->st));
    if (ipiv) CUC32(cudaMemcpyAsync(ipiv, h->d_idx, sizeof(int) * nidx, cudaMemcpyDeviceToHost, h->st));
    if (rowbycol) CUC32(cudaMemcpyAsync(rowbycol, h->d_idx + nidx, sizeof(int) * nidx, cudaMemcpyDeviceToHost, h->st));
    if (colbyrow) CUC32(cudaMemcpyAsync(colbyrow, h->d_idx + 2 * nidx, sizeof(int) * nidx, cudaMemcpyDeviceToHost, h->st));
    if (info) CUC32(cudaMemcpyAsync(info, h->d_info, sizeof(int) * (size_t)nsys, cudaMemcpyDeviceToHost, h->st));
    CUC32(cudaStreamSynchronize(h->st));
#undef CUC32
    *out = h;
    return BDILU_OK;
}

int bdilu_ilu_solve_f32(bdilu_ilu_f32 *h, const float *host_b, float *host_x, int nrhs, int *flags)
{
    if (!h || !host_b || !host_x || nrhs < 1) return api_fail(BDILU_ERR_ARG, "null handle/pointer or nrhs < 1");
    DevGuard guard(h->device);
    if (!guard.ok) return api_fail(BDILU_ERR_CUDA, "cannot select the handle's device");
    const int64_t n = (int64_t)h->nblocks * h->blockw;
    const size_t nvec = (size_t)h->nsys * nrhs, nidx = (size_t)h->nsys * n;
    const size_t vbytes = sizeof(float) * nvec * (size_t)n;
    if (vbytes > h->vec_cap) {
        cudaFree(h->d_b); cudaFree(h->d_x);
        h->d_b = h->d_x = nullptr; h->vec_cap = 0;
        CU32(cudaMalloc(&h->d_b, vbytes ? vbytes : 8));
        CU32(cudaMalloc(&h->d_x, vbytes ? vbytes : 8));
        h->vec_cap = vbytes;
    }
    if (nvec > h->flag_cap) {
        cudaFree(h->d_flags); h->d_flags = nullptr; h->flag_cap = 0;
        CU32(cudaMalloc(&h->d_flags, sizeof(int) * nvec));
        h->flag_cap = nvec;
    }
    std::vector<int> hf;
    try {
        hf.resize(nvec);
    } catch (const std::bad_alloc &) {
        return api_fail(BDILU_ERR_ARG, "out of host memory");
    }
    // copies into the caller's memory may be in flight from here on: every exit goes through the synchronisation
    int rc = BDILU_OK;
    cudaError_t e = cudaMemcpyAsync(h->d_b, host_b, vbytes, cudaMemcpyHostToDevice, h->st);
    if (e == cudaSuccess) {
        rc = bdilu_solve_f32(h->d_data, h->sys_stride, h->nsys, h->nblocks, h->blockw, h->ndiag, h->nsat, h->ld,
                             h->d_idx + nidx, h->d_idx + 2 * nidx, h->d_b, h->d_x, nrhs, n, n * nrhs, h->d_flags, h->st);
        if (rc == BDILU_OK) e = cudaMemcpyAsync(host_x, h->d_x, vbytes, cudaMemcpyDeviceToHost, h->st);
        if (rc == BDILU_OK && e == cudaSuccess)
            e = cudaMemcpyAsync(hf.data(), h->d_flags, sizeof(int) * nvec, cudaMemcpyDeviceToHost, h->st);
    }
    const cudaError_t es = cudaStreamSynchronize(h->st);
    if (rc != BDILU_OK) return rc;
    if (e != cudaSuccess || es != cudaSuccess) return api_cuda_fail(e != cudaSuccess ? e : es, "bdilu_ilu_solve_f32");
    int first = 0;
    for (size_t i = 0; i < nvec; ++i) {
        if (flags) flags[i] = hf[i];
        if (!first) first = hf[i];
    }
    return first;
}

void bdilu_ilu_destroy_f32(bdilu_ilu_f32 *h) { free_handle_f32(h); }

int bdilu_host_dot_vec_f32(const float *host_data, int nblocks, int blockw, int ndiag, int nsat, int ld,
                           const float *host_vec, float *host_out, int nrhs)
{
    if (!host_data || !host_vec || !host_out || nrhs < 1) return api_fail(BDILU_ERR_ARG, "null pointer or nrhs < 1");
    if (bdilu_device_count() < 1) return api_fail(BDILU_ERR_CUDA, "no CUDA device: libbdilu_b200 has no CPU fallback");
    if (ld == 0) ld = bdilu_default_ld_f32(blockw);
    const int64_t nd = bdilu_ndata(nblocks, blockw, ndiag, nsat, ld), n = (int64_t)nblocks * blockw;
    float *d_data = nullptr, *d_v = nullptr, *d_o = nullptr;
    int rc = BDILU_OK;
    cudaError_t e;
    if ((e = cudaMalloc(&d_data, sizeof(float) * (nd ? nd : 1))) != cudaSuccess ||
        (e = cudaMalloc(&d_v, sizeof(float) * (n * nrhs ? n * nrhs : 1))) != cudaSuccess ||
        (e = cudaMalloc(&d_o, sizeof(float) * (n * nrhs ? n * nrhs : 1))) != cudaSuccess ||
        (e = cudaMemcpy(d_data, host_data, sizeof(float) * nd, cudaMemcpyHostToDevice)) != cudaSuccess ||
        (e = cudaMemcpy(d_v, host_vec, sizeof(float) * n * nrhs, cudaMemcpyHostToDevice)) != cudaSuccess) {
        rc = api_cuda_fail(e, "bdilu_host_dot_vec_f32");
    } else {
        rc = bdilu_dot_vec_f32(d_data, nd, 1, nblocks, blockw, ndiag, nsat, ld, d_v, d_o, nrhs, n, n * nrhs, nullptr);
        if (!rc && (e = cudaMemcpy(host_out, d_o, sizeof(float) * n * nrhs, cudaMemcpyDeviceToHost)) != cudaSuccess)
            rc = api_cuda_fail(e, "bdilu_host_dot_vec_f32");
    }
    cudaFree(d_data); cudaFree(d_v); cudaFree(d_o);
    return rc;
}

int bdilu_host_scale_diag_add_f32(float *host_dst, const float *host_src, int nblocks, int blockw, int ndiag, int nsat,
                                  int ld, float scale, float diag_add)
{
    if (!host_dst || !host_src) return api_fail(BDILU_ERR_ARG, "null pointer");
    if (bdilu_device_count() < 1) return api_fail(BDILU_ERR_CUDA, "no CUDA device: libbdilu_b200 has no CPU fallback");
    if (ld == 0) ld = bdilu_default_ld_f32(blockw);
    const int64_t nd = bdilu_ndata(nblocks, blockw, ndiag, nsat, ld);
    float *d = nullptr;
    int rc = BDILU_OK;
    cudaError_t e;
    if ((e = cudaMalloc(&d, sizeof(float) * (nd ? nd : 1))) != cudaSuccess ||
        (e = cudaMemcpy(d, host_src, sizeof(float) * nd, cudaMemcpyHostToDevice)) != cudaSuccess) {
        rc = api_cuda_fail(e, "bdilu_host_scale_diag_add_f32");
    } else {
        rc = bdilu_scale_diag_add_f32(d, d, nd, 1, nblocks, blockw, ndiag, nsat, ld, scale, diag_add, nullptr);
        if (!rc && (e = cudaMemcpy(host_dst, d, sizeof(float) * nd, cudaMemcpyDeviceToHost)) != cudaSuccess)
            rc = api_cuda_fail(e, "bdilu_host_scale_diag_add_f32");
    }
    cudaFree(d);
    return rc;
}

}  // extern "C"
