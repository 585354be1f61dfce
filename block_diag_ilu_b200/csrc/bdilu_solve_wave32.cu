// Kernel 2, blockw == 32 fast path ("wave32"): ILU_inplace<double>::solve (block_diag_ilu.hpp:486-538) as a
// wavefront over consecutive blocks, one CTA per (system, right-hand side).  Same idea as the generic
// wavefront kernel (bdilu_solve_wave.cu) -- row li of block b needs only y_{b-d}[colbyrow[li]] from its
// predecessors (:508-511, :524-527), so block b can run a couple of rows behind block b-1 -- restructured so
// that a column step costs a handful of instructions and nothing slow sits on the dependency chain:
//
//   * producer warp: lanes 0..3+ND each own ONE of the per-block pieces (the 8 KB block, colbyrow, rowbycol,
//     b, the ND sub/sup rows) and issue their TMA bulk copy in the same instruction; pointers advance by a
//     constant stride, stage index and phase are tracked incrementally (no division, no 64-bit multiply).
//   * consumer warps (lane = row, warp j owns blocks j, j+NW, ...): the row of L (U) is copied from the stage
//     into registers up front and the stage is released at once; the sweep itself touches shared memory only
//     to publish finished values and to poll the predecessors' ring slots.
//   * step k: lane k publishes its finished value (one predicated store) and broadcasts it by shuffle; every
//     lane then does one unpredicated DFMA with the block entry in column k.  Lanes that are already finished
//     pick up garbage from that DFMA -- they are dead: a lane's result was captured when it was published.
//   * coupling: two accumulators (even / odd sweep positions) keep the coupling terms off the dependency chain;
//     "not yet written" is an in-band signalling-NaN pattern in the shared-memory ring, polled one step ahead by
//     a warp-uniform vote.  The coupling is applied at a fixed step, so the result is bit-reproducible.
//   * U rows are pre-scaled by 1/U[r][r], so the backward chain has no division on it.
// Blocks "before the first" are virtual ring slots holding zeros, so every block sees exactly ND predecessors.
#include "bdilu_wave_common.cuh"

namespace bdilu {

namespace wave {

// Phase profile (debug builds only, -DBDILU_W32_PROF): cycles summed over consumer warps / the producer warp.
//   [0] wait for the stage  [1] stage -> registers  [2] first coupling poll  [3] the 32-step chain  [4] store
//   [5] producer: wait for an empty stage  [6] producer: issue  [7] blocks counted
#ifdef BDILU_W32_PROF
__device__ unsigned long long g_w32prof[8];
#define W32_T(var) const long long var = clock64()
#define W32_ADD(slot, a, b) if (r == 0) atomicAdd(&g_w32prof[slot], (unsigned long long)((b) - (a)))
#else
#define W32_T(var)
#define W32_ADD(slot, a, b)
#endif

constexpr int RS = 32;  // ring slots (power of two, >= NW + MAXD + 1)
static_assert(NW + MAXD + 1 <= RS, "ring too small for the reset protocol");

__device__ __forceinline__ void ring_reset(const Smem &S, int tid, int nthreads, int nd)
{
    for (int t = tid; t < RS * 32; t += nthreads) {
        const int slot = t >> 5;
        S.ring[t] = (slot >= RS - nd) ? 0.0 : __longlong_as_double((long long)SENT);
    }
}

// ---- producer warp: lane = piece ---------------------------------------------------------------------------
template <bool FWD, int ND>
__device__ void produce32(const Args &A, const Smem &S, int s, unsigned ph, bool first_pass, int lane)
{
    const int N = A.D.nblocks, w = A.D.w;
    const int dstride = (A.SL.diag_bytes + 15) & ~15;
    const long long step = FWD ? 1 : -1;
    const int b0 = FWD ? 0 : N - 1;
    // piece owned by this lane: source pointer at block order 0, bytes per order step, offset in the stage
    const char *src = nullptr;
    long long inc = 0;
    unsigned bytes = 0, dst = 0;
    int dmine = -1;  // sub/sup diagonal index for lanes 4.., -1 otherwise
    if (lane == 0) {
        src = reinterpret_cast<const char *>(A.sd + A.D.block(b0));
        inc = step * (long long)w * A.D.ld * 8;
        bytes = A.SL.blk_bytes;
        dst = 0;
    } else if (lane == 1) {
        src = reinterpret_cast<const char *>(A.cbr + (int64_t)b0 * w);
        inc = step * w * 4;
        bytes = A.SL.idx_bytes;
        dst = A.SL.o_cbr;
    } else if (lane == 2 && FWD) {
        src = reinterpret_cast<const char *>(A.rbc + (int64_t)b0 * w);
        inc = step * w * 4;
        bytes = A.SL.idx_bytes;
        dst = A.SL.o_rbc;
    } else if (lane == 3 && FWD) {
        src = reinterpret_cast<const char *>(A.b + (int64_t)b0 * w);
        inc = step * w * 8;
        bytes = A.SL.vec_bytes;
        dst = A.SL.o_vec;
    } else if (lane >= 4 && lane < 4 + ND) {
        dmine = lane - 4;
        // forward: sub(d, b-1-d, :) couples block b to y_{b-1-d}; backward: sup(d, b, :) to x_{b+1+d}
        const double *p = FWD ? A.sd + A.D.off_sub + A.D.diag_chunk(dmine, 0) - (int64_t)(1 + dmine) * A.D.ld
                              : A.sd + A.D.off_sup + A.D.diag_chunk(dmine, 0) + (int64_t)(N - 1) * A.D.ld;
        src = reinterpret_cast<const char *>(p);
        inc = step * (long long)A.D.ld * 8;
        bytes = A.SL.diag_bytes;
        dst = A.SL.o_diag + dmine * dstride;
    }
    const unsigned tx_fixed = A.SL.blk_bytes + A.SL.idx_bytes + (FWD ? A.SL.idx_bytes + A.SL.vec_bytes : 0);
    bool reuse = !first_pass;  // stages were used before: wait for the consumers' release
    for (int i = 0; i < N; ++i) {
#ifdef BDILU_W32_PROF
        const int r = lane;
#endif
        W32_T(p0);
        if (reuse) mbar_wait(&S.empty[s], ph ^ 1u);
        W32_T(p1);
        unsigned char *stg = S.stages + (size_t)s * A.SL.bytes;
        const int nd = min(ND, i);
        if (lane == 0) mbar_expect_tx(&S.full[s], tx_fixed + nd * A.SL.diag_bytes);
        if (bytes != 0 && dmine < nd) bulk_g2s(stg + dst, src, bytes, &S.full[s]);
        src += inc;
        W32_T(p2);
        W32_ADD(5, p0, p1);
        W32_ADD(6, p1, p2);
        if (++s == A.NS) {
            s = 0;
            ph ^= 1u;
            reuse = true;
        }
    }
}

// ---- consumers -----------------------------------------------------------------------------------------------
// Ring protocol (RS = 32 slots of 32 doubles, slot of block order i = i & 31):
//   * block i publishes row values into slot i, polls slots i-1 .. i-ND ("not yet written" = SENT);
//   * after its LAST coupling read, and BEFORE it publishes its last rows, block i resets slot i-ND to SENT.  When
//     block i has read everything it needs, blocks i-1 .. i-ND+1 have published everything, hence finished reading
//     slot i-ND too (every row is read before the row that needs it is published), so block i is the last reader.
//     The next writer of that slot, block i-ND+32, is started by a warp that first finished block i-ND+32-NW >= i+1
//     (NW + ND + 1 <= 32), which cannot complete before block i published its last rows -> the reset is ordered
//     before the reuse, with no counters and no extra waiting.
//
// Dependency chain of one block = 32 x (shuffle broadcast + one DFMA).  Nothing else touches the register the chain
// runs through: positions of even parity accumulate in accA, odd ones in accB; step kk broadcasts from the register of
// parity kk, updates the register of parity kk+1 first (the next step reads it) and only then the other one, into which
// the coupling terms of position kk+2 are merged while that register has a whole step of slack.  The predecessor
// values for that merge were loaded one step earlier, so the poll costs no latency unless the predecessor is late.
template <bool FWD, int ND>
__device__ void consume32(const Args &A, const Smem &S, int s0, unsigned ph0, int j, int r, int &nan_first,
                          int &zero_last)
{
    constexpr int W = 32;
    const int N = A.D.nblocks, ld = A.D.ld;
    const int dstride = (A.SL.diag_bytes + 15) & ~15;
    double ynext = 0.0;
    if (!FWD && j < N) ynext = A.x[(int64_t)(N - 1 - j) * W + r];
    // stage index / phase of my first block, advanced by NW per block
    int s = s0 + j;
    unsigned ph = ph0;
    while (s >= A.NS) {
        s -= A.NS;
        ph ^= 1u;
    }
    const int q = FWD ? r : W - 1 - r;  // position of my row in sweep order

    for (int i = j; i < N; i += NW) {
        const int b = FWD ? i : N - 1 - i;
        const double y_cur = ynext;
        if (!FWD && i + NW < N) ynext = A.x[(int64_t)(N - 1 - (i + NW)) * W + r];

        W32_T(t0);
        mbar_wait(&S.full[s], ph);
        W32_T(t1);
        const unsigned char *st = S.stages + (size_t)s * A.SL.bytes;
        const double *Ls = reinterpret_cast<const double *>(st);
        double l[W];
#pragma unroll
        for (int k = 0; k < W; ++k) l[k] = Ls[k * ld + r];
        const int ci = reinterpret_cast<const int *>(st + A.SL.o_cbr)[r];
        const int nd = min(ND, i);
        double cv[ND];
#pragma unroll
        for (int d = 0; d < ND; ++d)
            cv[d] = (d < nd) ? reinterpret_cast<const double *>(st + A.SL.o_diag + d * dstride)[ci] : 0.0;
        double acc;
        if (FWD) {
            const int src = reinterpret_cast<const int *>(st + A.SL.o_rbc)[r];
            const double bv = reinterpret_cast<const double *>(st + A.SL.o_vec)[src];
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
        if (r == 0) mbar_arrive(&S.empty[s]);  // everything this block needs is in registers
        s += NW;
        if (s >= A.NS) {
            s -= A.NS;
            ph ^= 1u;
        }

        double *myslot = S.ring + (i & (RS - 1)) * 32;
        double *resetslot = S.ring + ((i - ND) & (RS - 1)) * 32 + r;
        const double *pd[ND];
#pragma unroll
        for (int d = 0; d < ND; ++d) pd[d] = S.ring + ((i - 1 - d) & (RS - 1)) * 32 + ci;

        double v[ND];
        auto load_pred = [&]() {
#pragma unroll
            for (int d = 0; d < ND; ++d) v[d] = ld_volatile_s(pd[d]);
        };
        // Warp-uniform wait until the lanes selected by `mine` hold real predecessor values in v[] (no lane ever
        // spins on its own, so the warp never splits); the first test uses the values loaded a step earlier.
        auto await_pred = [&](bool mine) {
            unsigned spins = 0;
            for (;;) {
                bool bad = false;
#pragma unroll
                for (int d = 0; d < ND; ++d) bad |= is_sent(v[d]);
                if (!__any_sync(FULL, mine && bad)) break;
                if (++spins > SPIN_LIMIT) __trap();  // bounded: traps instead of hanging
                load_pred();
            }
        };
        auto coupled = [&](double a) {
#pragma unroll
            for (int d = ND - 1; d >= 0; --d) a = fma(-cv[d], v[d], a);
            return a;
        };

        // positions 0 and 1 take their coupling terms before the chain starts
        W32_T(t2);
        load_pred();
        await_pred(q < 2);
        W32_T(t3);
        if (q < 2) acc = coupled(acc);
        double accA = acc, accB = acc, mine_final = acc;
        load_pred();
#pragma unroll
        for (int kk = 0; kk < W; ++kk) {
            const int k = FWD ? kk : W - 1 - kk;
            double &cur = (kk & 1) ? accB : accA;   // holds the finished value of position kk (in lane k)
            double &nxt = (kk & 1) ? accA : accB;   // read by the next step: update it first
            if (r == k) {
                st_volatile_s(&myslot[k], cur);
                mine_final = cur;
            }
            const double vk = group_bcast<32>(cur, k);
            if (kk + 1 < W) nxt = fma(-l[k], vk, nxt);
            if (kk + 2 < W) {
                cur = fma(-l[k], vk, cur);
                await_pred(q == kk + 2);
                if (q == kk + 2) cur = coupled(cur);
                if (kk + 3 < W) load_pred();
                else st_volatile_s(resetslot, __longlong_as_double((long long)SENT));  // all my reads are done
            }
        }
        W32_T(t4);
        A.x[(int64_t)b * W + r] = mine_final;
        W32_T(t5);
        W32_ADD(0, t0, t1);
        W32_ADD(1, t1, t2);
        W32_ADD(2, t2, t3);
        W32_ADD(3, t3, t4);
        W32_ADD(4, t4, t5);
        W32_ADD(7, 0, 1);
    }
}

template <int ND>
__global__ void __launch_bounds__((NW + 1) * 32, 1)
solve_wave32_kernel(const double *__restrict__ data, int64_t sys_stride, Dims D, const int *__restrict__ rowbycol,
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
    off += (size_t)RS * 32 * 8;
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
    ring_reset(S, tid, blockDim.x, ND);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();

    // stage index and phase at which the backward sweep starts (the forward sweep used N stage slots)
    const int s_bwd = N % NS;
    const unsigned ph_bwd = (unsigned)(N / NS) & 1u;

    int nan_first = 0x7fffffff, zero_last = -1;
    if (warp == NW) {
        produce32<true, ND>(A, S, 0, 0u, true, lane);
    } else {
        consume32<true, ND>(A, S, 0, 0u, warp, lane, nan_first, zero_last);
    }
    __syncthreads();  // y (in the x buffer) is complete and visible to the whole CTA
    ring_reset(S, tid, blockDim.x, ND);
    __syncthreads();
    if (warp == NW) {
        produce32<false, ND>(A, S, s_bwd, ph_bwd, N < NS, lane);
    } else {
        consume32<false, ND>(A, S, s_bwd, ph_bwd, warp, lane, nan_first, zero_last);
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

template <int ND>
static cudaError_t launch_cfg(unsigned nvec, size_t smem, cudaStream_t st, const double *data, int64_t sys_stride,
                              const Dims &D, const int *rowbycol, const int *colbyrow, const double *b, double *x,
                              const VecLayout &V, int *flags, int NS)
{
    cudaError_t e = cudaFuncSetAttribute(solve_wave32_kernel<ND>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         227 * 1024);
    if (e != cudaSuccess) return e;
    solve_wave32_kernel<ND><<<nvec, (NW + 1) * 32, smem, st>>>(data, sys_stride, D, rowbycol, colbyrow, b, x, V,
                                                                       flags, NS);
    return cudaGetLastError();
}

}  // namespace wave

#ifdef BDILU_W32_PROF
extern "C" __attribute__((visibility("default"))) int bdilu_debug_w32prof(unsigned long long *out8, int reset)
{
    if (out8) cudaMemcpyFromSymbol(out8, wave::g_w32prof, sizeof(unsigned long long) * 8);
    if (reset) {
        unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        cudaMemcpyToSymbol(wave::g_w32prof, z, sizeof z);
    }
    return 0;
}
#endif

// blockw == 32, 1 <= ndiag <= 4; the caller has checked alignment and sizes (try_launch_solve_wave).
cudaError_t launch_solve_wave32(unsigned nvec, size_t smem, int NS, const double *data, int64_t sys_stride, const Dims &D,
                                const int *rowbycol, const int *colbyrow, const double *b, double *x,
                                const VecLayout &V, int *flags, cudaStream_t st)
{
    using namespace wave;
#define BDILU_W32(ND_) return launch_cfg<ND_>(nvec, smem, st, data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, NS);
    switch (D.ndiag) {
    case 1: BDILU_W32(1)
    case 2: BDILU_W32(2)
    case 3: BDILU_W32(3)
    default: BDILU_W32(4)
    }
#undef BDILU_W32
}

}  // namespace bdilu
