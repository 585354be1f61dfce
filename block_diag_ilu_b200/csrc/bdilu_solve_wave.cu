// Kernel 2 (fast path): ILU_inplace<double>::solve (block_diag_ilu.hpp:486-538) as a WAVEFRONT.
//
// The reference walks the block chain strictly in order.  The dependency that forces this is thin:
// row li of block b needs only y_{b-d}[colbyrow[li]] (one scalar per sub-diagonal, :508-511) plus the
// earlier rows of its own block.  So consecutive blocks can be in flight at once, each one a few rows
// behind its predecessor -- same arithmetic per element, no algebraic reformulation.
//
// One CTA per (system, right-hand side):
//   * 1 producer warp: one lane streams whole blocks (L\U, w*ld*8 contiguous bytes) and the block's
//     rowbycol / colbyrow / b / sub (or sup) rows from HBM into a shared-memory stage ring with TMA
//     bulk copies (cp.async.bulk + mbarrier complete_tx), NS stages deep, so HBM latency is never on
//     the chain.
//   * NW consumer warps: warp j owns blocks j, j+NW, ... (lane = row).  Per block a column sweep:
//     lane k's finished value is broadcast by shuffle, every later row does one DFMA with its L (U)
//     entry read from the stage.  Finished values are also published to a small ring in shared
//     memory; the successor blocks poll that ring for exactly the entries their rows need.
//     "Not yet written" is encoded in-band as a signalling-NaN bit pattern that no arithmetic result
//     can have, so one 8-byte shared load delivers data and flag together and the producer side
//     needs no fence.
// Forward sweep writes y to the x buffer (the reference heap-allocates y, :499); after a CTA-wide
// barrier the backward sweep runs the mirror image over U and sup and overwrites x.
#include "bdilu_wave_common.cuh"

namespace bdilu {

namespace wave {

// ---- consumers: warp j sweeps blocks of chain order j, j+NW, ... ----------------------------------
template <bool FWD>
__device__ void consume(const Args &A, const Smem &S, int it0, int j, int r, int &nan_first, int &zero_last)
{
    const int N = A.D.nblocks, w = A.D.w, ld = A.D.ld, ND = A.D.ndiag;
    const bool rowok = r < w;
    double ynext = 0.0;
    if (!FWD && j < N && rowok) ynext = A.x[(int64_t)(N - 1 - j) * w + r];

    for (int i = j; i < N; i += NW) {
        const int b = FWD ? i : N - 1 - i;
        const int it = it0 + i, s = it % A.NS, use = it / A.NS;
        // Clean the ring slot my next block (order i+NW) will publish into.  It still holds the values
        // of order i-NW; their consumers are orders i-NW+1 .. i-NW+ND, which must have completed.
        if (i >= NW) {
            for (int d = 0; d < ND; ++d) {
                const int oi = i - NW + 1 + d;
                if (oi < N) {
                    unsigned spins = 0;
                    while (ld_volatile_s(&S.done[oi % NW]) < oi + 1)
                        if (++spins > SPIN_LIMIT) __trap();
                }
            }
            st_volatile_s(&S.ring[((i + NW) % (2 * NW)) * 32 + r], __longlong_as_double((long long)SENT));
        }
        double y_cur = ynext;
        if (!FWD && i + NW < N && rowok) ynext = A.x[(int64_t)(N - 1 - (i + NW)) * w + r];

        mbar_wait(&S.full[s], use & 1);
        const unsigned char *st = S.stages + (size_t)s * A.SL.bytes;
        const double *Ls = reinterpret_cast<const double *>(st);
        const int *cbr_s = reinterpret_cast<const int *>(st + A.SL.o_cbr);
        const int nd = min(ND, i);
        int ci = 0;
        unsigned need = 0;
        double cv[MAXD];
        double acc = 0.0, dg = 1.0;
        if (rowok) {
            ci = cbr_s[r];
            if (FWD) {
                const int src = reinterpret_cast<const int *>(st + A.SL.o_rbc)[r];
                acc = reinterpret_cast<const double *>(st + A.SL.o_vec)[src];
                if (isnan(acc)) nan_first = min(nan_first, b * w + src + 1);
            } else {
                acc = y_cur;
                dg = Ls[r * ld + r];
                if (dg == 0.0) zero_last = max(zero_last, b * w + r);
            }
            need = (1u << nd) - 1u;
        }
#pragma unroll
        for (int d = 0; d < MAXD; ++d)
            cv[d] = (rowok && d < nd)
                        ? reinterpret_cast<const double *>(st + A.SL.o_diag + d * ((A.SL.diag_bytes + 15) & ~15))[ci]
                        : 0.0;
        const double rinv = 1.0 / dg;
        double *myslot = S.ring + (i % (2 * NW)) * 32;

        // Coupling terms are summed into `cacc` in a FIXED order (farthest sub-diagonal first, i.e. the
        // order the values become available) and subtracted from the row at a FIXED step (just before the
        // row's last L/U term), so the rounding does not depend on when the values happen to arrive:
        // results are bit-reproducible from run to run.
        double cacc = 0.0;
        auto poll = [&]() {
            bool go = true;
#pragma unroll
            for (int d = MAXD - 1; d >= 0; --d) {
                if (go && (need & (1u << d))) {
                    const double v = ld_volatile_s(&S.ring[((i - 1 - d) % (2 * NW)) * 32 + ci]);
                    if (!is_sent(v)) {
                        cacc = fma(cv[d], v, cacc);
                        need &= ~(1u << d);
                    } else {
                        go = false;
                    }
                }
            }
        };
        const int kapply = FWD ? max(r - 1, 0) : min(r + 1, w - 1);

#pragma unroll 4
        for (int kk = 0; kk < w; ++kk) {
            const int k = FWD ? kk : w - 1 - kk;
            const double lk = (rowok && (FWD ? r > k : r < k)) ? Ls[k * ld + r] : 0.0;
            if (need) poll();
            const bool apply_now = rowok && k == kapply;
            {
                unsigned spins = 0;
                while (__any_sync(FULL, apply_now && need != 0)) {
                    if (++spins > SPIN_LIMIT) __trap();
                    poll();
                }
            }
            if (apply_now) acc -= cacc;
            if (!FWD && r == k) {
                // x = acc / U[k][k]: reciprocal + one Newton correction gives the correctly rounded
                // quotient off the slow division path; a zero diagonal takes the true division so that
                // inf/NaN come out as in the reference (:531-534)
                double q = acc * rinv;
                const double rem = fma(-q, dg, acc);
                q = fma(rem, rinv, q);
                if (dg == 0.0 || !isfinite(q)) q = acc / dg;
                acc = q;
            }
            const double vk = group_bcast<32>(acc, k);
            if (r == k) {
                double pub = acc;
                if (is_sent(pub)) pub = __longlong_as_double(0x7FF8000000000000ll);
                st_volatile_s(&myslot[k], pub);
            }
            if (FWD ? r > k : r < k) acc = fma(-lk, vk, acc);
        }
        if (rowok) A.x[(int64_t)b * w + r] = acc;
        __syncwarp();
        if (r == 0) {
            mbar_arrive(&S.empty[s]);
            st_volatile_s(&S.done[j], i + 1);
        }
    }
}

__global__ void __launch_bounds__((NW + 1) * 32, 1)
solve_wave_kernel(const double *__restrict__ data, int64_t sys_stride, Dims D, const int *__restrict__ rowbycol,
                  const int *__restrict__ colbyrow, const double *__restrict__ bvec, double *xvec, VecLayout V,
                  int *__restrict__ flags, int NS)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int vecid = blockIdx.x;
    const int sys = vecid / V.nrhs, rhs = vecid % V.nrhs;
    const int N = D.nblocks, w = D.w;

    Args A;
    A.sd = data + (int64_t)sys * sys_stride;
    A.rbc = rowbycol + (int64_t)sys * N * w;
    A.cbr = colbyrow + (int64_t)sys * N * w;
    A.b = bvec + (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride;
    A.x = xvec + (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride;
    A.D = D;
    A.NS = NS;
    A.SL = stage_layout(w, D.ld, D.ndiag);
    A.dbg = 0;

    Smem S;
    S.stages = smem_raw;
    size_t off = (size_t)NS * A.SL.bytes;
    S.full = reinterpret_cast<unsigned long long *>(smem_raw + off);
    S.empty = S.full + NS;
    off += (size_t)2 * NS * 8;
    S.ring = reinterpret_cast<double *>(smem_raw + off);
    off += (size_t)2 * NW * 32 * 8;
    S.done = reinterpret_cast<int *>(smem_raw + off);
    S.flagw = S.done + NW;

    if (tid == 0) {
        for (int s = 0; s < NS; ++s) {
            mbar_init(&S.full[s], 1);
            mbar_init(&S.empty[s], 1);
        }
        S.flagw[0] = 0x7fffffff;
        S.flagw[1] = -1;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int t = tid; t < 2 * NW * 32; t += blockDim.x) S.ring[t] = __longlong_as_double((long long)SENT);
    if (tid < NW) S.done[tid] = 0;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();

    int nan_first = 0x7fffffff, zero_last = -1;
    // -------- forward --------
    if (warp == NW) {
        if (lane == 0) produce<true>(A, S, 0);
        __syncwarp();
    } else {
        consume<true>(A, S, 0, warp, lane, nan_first, zero_last);
    }
    __syncthreads();  // y (in the x buffer) is complete and visible to the whole CTA
    for (int t = tid; t < 2 * NW * 32; t += blockDim.x) S.ring[t] = __longlong_as_double((long long)SENT);
    if (tid < NW) S.done[tid] = 0;
    __syncthreads();
    // -------- backward --------
    if (warp == NW) {
        if (lane == 0) produce<false>(A, S, N);
        __syncwarp();
    } else {
        consume<false>(A, S, N, warp, lane, nan_first, zero_last);
    }
    if (warp < NW) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            nan_first = min(nan_first, __shfl_xor_sync(FULL, nan_first, o));
            zero_last = max(zero_last, __shfl_xor_sync(FULL, zero_last, o));
        }
        if (lane == 0) {
            atomicMin(&S.flagw[0], nan_first);
            atomicMax(&S.flagw[1], zero_last);
        }
    }
    __syncthreads();
    if (tid == 0 && flags) {
        const int nf = S.flagw[0], zl = S.flagw[1];
        flags[vecid] = (nf != 0x7fffffff) ? nf : (zl >= 0 ? N * w + zl : 0);
    }
}


}  // namespace wave

// Returns true (and launches) when the fast path applies; false -> caller uses the general kernel.
bool try_launch_solve_wave(const double *data, int64_t sys_stride, int nsys, const Dims &D, const int *rowbycol,
                           const int *colbyrow, const double *b, double *x, const VecLayout &V, int *flags,
                           cudaStream_t st, cudaError_t *err)
{
    using namespace wave;
    *err = cudaSuccess;
    const int w = D.w;
    const int64_t nvec = (int64_t)nsys * V.nrhs;
    if (w > 32 || w < 4 || (w & 3) || (D.ld & 1) || D.ndiag < 1 || D.ndiag > MAXD || D.nsat != 0) return false;
    if (D.nblocks < 4 * NW) return false;                       // short chains: nothing to pipeline
    if (nvec > 8 * sm_count() && w < 32) return false;                 // wide batches of small systems: chain kernel
    if ((sys_stride & 1) || (V.rhs_stride & 1) || (V.sys_stride & 1)) return false;
    if ((reinterpret_cast<uintptr_t>(data) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(rowbycol) |
         reinterpret_cast<uintptr_t>(colbyrow)) & 15)
        return false;
    if ((D.off_sub & 1) || (D.off_sup & 1)) return false;
    const StageLayout SL = stage_layout(w, D.ld, D.ndiag);
    const size_t fixed = (size_t)2 * NW * 32 * 8 + (NW + 2) * 4 + 64;
    const size_t budget = 227 * 1024;
    int NS = (int)((budget - fixed) / (SL.bytes + 16));
    if (NS > MAX_STAGES) NS = MAX_STAGES;
    if (NS < NW + 2) return false;
    const size_t smem = (size_t)NS * (SL.bytes + 16) + fixed;
    // per (function, device): set on every launch
    *err = cudaFuncSetAttribute(solve_wave_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget);
    if (*err != cudaSuccess) return true;
    if (w == 32 && g_solve_path == 8) {
        // EXPERIMENTAL, opt-in (test hook 8; measured slower than quad32 on B200, see bdilu_solve_skew.cu): systolic
        // kernel first (vectors of systems that never pivoted), the general wavefront kernel behind it for the vectors
        // handed back through `status` (stream-ordered scratch: one int per vector).
        int *status = nullptr;
        if (cudaMallocAsync(reinterpret_cast<void **>(&status), sizeof(int) * (size_t)nvec, st) == cudaSuccess) {
            if (try_launch_solve_skew32((unsigned)nvec, data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, status, st, err)) {
                if (*err == cudaSuccess &&
                    try_launch_solve_quad32((unsigned)nvec, data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, st, err, status)) {
                    count_launch(1);
                    cudaFreeAsync(status, st);
                    return true;
                }
                // the general kernel does not fit this shape either way: fall through with the scratch released; the
                // paths below redo every vector
                cudaFreeAsync(status, st);
                if (*err != cudaSuccess) return true;
            } else {
                cudaFreeAsync(status, st);
            }
        } else {
            (void)cudaGetLastError();
        }
    }
#ifndef BDILU_HALF32_AUTO
#define BDILU_HALF32_AUTO 0
#endif
    if (w == 32 && (g_solve_path == 10 || (g_solve_path == 0 && BDILU_HALF32_AUTO && 2 * nvec <= sm_count()))) {
        // two SMs per chain while the batch leaves half of the GPU idle (cfg3: 64 chains on 148 SMs); the vectors of
        // systems that pivoted across the middle of a block come back through `status` and run on quad32 behind it
        int *scr = nullptr;
        const size_t ints = (size_t)nvec * (1 + (size_t)half32_scratch_ints());
        if (cudaMallocAsync(reinterpret_cast<void **>(&scr), sizeof(int) * ints, st) == cudaSuccess) {
            int *status = scr, *scratch = scr + nvec;
            if (try_launch_solve_half32((unsigned)nvec, data, sys_stride, nsys, D, rowbycol, colbyrow, b, x, V, flags, status,
                                        scratch, st, err)) {
                if (*err == cudaSuccess &&
                    try_launch_solve_quad32((unsigned)nvec, data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, st, err, status)) {
                    count_launch(3);
                    cudaFreeAsync(scr, st);
                    return true;
                }
                cudaFreeAsync(scr, st);
                if (*err != cudaSuccess) return true;
            } else {
                cudaFreeAsync(scr, st);
            }
        } else {
            (void)cudaGetLastError();
        }
    }
    if (w == 32 && (g_solve_path == 0 || g_solve_path == 5 || g_solve_path == 7 || g_solve_path == 8 || g_solve_path == 10) &&
        try_launch_solve_quad32((unsigned)nvec, data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, st, err)) {
        count_launch(1);
        return true;
    }
    if (w == 32 && g_solve_path != 2) {
        *err = launch_solve_wave32((unsigned)nvec, smem, NS, data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, st);
        count_launch(1);
        return true;
    }
    solve_wave_kernel<<<(unsigned)nvec, (NW + 1) * 32, smem, st>>>(data, sys_stride, D, rowbycol, colbyrow, b, x, V,
                                                                    flags, NS);
    count_launch(1);
    *err = cudaGetLastError();
    return true;
}

}  // namespace bdilu
