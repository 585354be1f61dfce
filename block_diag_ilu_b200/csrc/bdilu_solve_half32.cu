// Kernel 2, blockw == 32, TWO SMs per chain ("half32"): ILU_inplace<double>::solve (block_diag_ilu.hpp:486-538).
// EXPERIMENTAL, opt-in (test hook 10 of bdilu_debug_force_chain_solve); parity-tested and bit-identical to quad32, NOT the
// default: on B200 it runs cfg3's solve in 3.5-4.1 ms (64 chains) / 3.0-3.4 ms (16 chains) against 2.70 ms for quad32.
// Measured (scratch/half_time.py; variants by -DH32_QW / -DH32_PG / -DH32_LEAD / -DH32_PUBL): without throttle and with
// one publication per 16 blocks 3.47 ms; coupling merged every row 3.52 / 2.97 ms (64 / 16 chains); 10 or 8 consumer warps
// are slower still.  Why it does not pay: the wavefront of a CTA advances one block per "lag" (the time a block needs
// from its start to its first publication, ~2.3 row times), whatever the number of rows per block, so halving the rows
// halves a warp's work per block but not the pace of the chain; the per-block fixed cost (stage hand-shake, row load,
// scaling, pointers: ~190 of quad32's ~600 instructions) is paid by both CTAs, and both stage every block in full (twice
// the L2 -> shared-memory traffic: the 64-chain case is the slower one).  What would be needed: per-half staging (a
// quarter / half of each block) and a shorter lag -- DESIGN.md section 8.
//
// quad32 (bdilu_solve_quad32.cu) runs one chain per SM: cfg3's 64 chains occupy 64 of the 148 SMs and each is bound by
// the instruction stream of its SM.  Splitting a chain between two SMs by alternating blocks would put a cross-SM
// handshake on the critical path of every hand-over.  This kernel splits the ROWS instead, which makes the dependency
// between the two SMs one-directional:
//   forward sweep (L y = P b):  rows 0..15 of every block need nothing from rows 16..31 (L is lower triangular), so the
//     CTA of rank 0 runs the whole chain on the upper-left 16 x 16 triangles; the CTA of rank 1 follows at any distance,
//     subtracts the square L[16..31][0..15] * y[0..15] (read back from the x buffer) and runs the lower-right triangles;
//   backward sweep (U x = y):   the same with the roles swapped (rows 31..16 lead, rows 15..0 follow).
// The leader never waits for the follower (except a throttle LEAD blocks ahead that keeps the blocks it fetched in L2
// for the follower), the follower waits only for data produced long before: nothing latency-critical crosses the SMs.
// Each CTA is a quad32 wavefront over ALL blocks with half the rows: 16 row steps per block instead of 32, half the
// coupling merges -- the per-block instruction stream that bounds quad32 halves, on twice as many SMs.  The two CTAs
// form a cluster only to be co-scheduled (no distributed shared memory is used).
//
// Condition: the coupling terms of a row must stay inside its half: colbyrow maps {0..15} to {0..15} in every block (no
// row interchange crossed the middle of a block; always true for the diagonally dominant matrices this library is made
// for).  A scan kernel checks it per system; the others are handed to quad32 through `status`.
// Arithmetic per row: the operations of quad32's lean chain in the same order (a - l0*y0 - l1*y1 - ...), coupling terms
// merged at the same points, U rows pre-scaled by 1/U[r][r]: results are bit-identical to quad32's.
#include "bdilu_wave_common.cuh"

namespace bdilu {

namespace half32 {

using namespace wave;

constexpr int W = 32, H = 16;
constexpr int G = 4;     // blocks per stage group
constexpr int NP = 2;    // producer warps
#ifndef H32_QW
#define H32_QW 14
#endif
#ifndef H32_PG
#define H32_PG 2
#endif
constexpr int QW = H32_QW;   // consumer warps
constexpr int RS = 32;   // ring slots
constexpr int PG = H32_PG;    // rows per coupling group
#ifndef H32_LEAD
#define H32_LEAD 128
#endif
#ifndef H32_PUBL
#define H32_PUBL 4
#endif
#ifndef H32_PUBF
#define H32_PUBF 2
#endif
constexpr int LEAD = H32_LEAD;     // blocks the leading half may run ahead of the following one (what it fetched stays in L2)
constexpr int PUB_LEAD = H32_PUBL;   // the leader publishes its progress every PUB_LEAD-th block of a warp (one fence each)
constexpr int PUB_FOLL = H32_PUBF;   // ... the follower every PUB_FOLL-th
constexpr int PW = 16;        // progress words per direction and role (one per consumer warp)
static_assert(QW + MAXD + 1 <= RS, "ring too small for the reset protocol");
static_assert(QW <= PW && LEAD > 2 * PUB_FOLL * QW, "throttle distance too short for the publication grain");

struct GroupLayout {
    int blk_bytes, ld8;
    int o_cbr, o_rbc, o_vec, o_diag, bytes;
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

// per-vector scratch in global memory (ints): progress words [direction][role][warp] (last block + 1 the warp has
// finished; warp j of both CTAs owns the same blocks j, j+QW, ...), then the words of the flag reduction
constexpr int SCR_FLAGS = 4 * PW;        // max(0x7fffffff - nan_first), max(zero_last + 1), arrivals
constexpr int SCR_INTS = 4 * PW + 4;

struct Ctx {
    const double *sd;
    const int *rbc, *cbr;
    const double *b;
    double *x;
    Dims D;
    GroupLayout GL;
    int NS, NG;
    unsigned char *stages;
    unsigned long long *full, *empty, *started, *sweep_done;
    double *ring;   // [RS][32]
    double *ytmp;   // [QW][16]: the partner half's values of the block a warp works on
    int *flagw;
    int *scr;       // this vector's global scratch
};

__device__ __forceinline__ void ring_reset(const Ctx &C, int tid, int nthreads, int nd)
{
    for (int t = tid; t < RS * 32; t += nthreads)
        C.ring[t] = ((t >> 5) >= RS - nd) ? 0.0 : __longlong_as_double((long long)SENT);
}

__device__ __forceinline__ int ld_volatile_g(const int *p)
{
    int v;
    asm volatile("ld.volatile.global.s32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_volatile_g(int *p, int v) { asm volatile("st.volatile.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }

// ---- producer warps: as quad32 (lane = piece, one bulk copy per piece per group of G blocks) ---------------------------
template <bool FWD, int ND>
__device__ void produce(const Ctx &C, int p, int lane)
{
    const int N = C.D.nblocks;
    const GroupLayout &GL = C.GL;
    for (int g = p; g < C.NG; g += NP) {
        const int gseq = (FWD ? 0 : C.NG) + g;
        const int s = gseq % C.NS, use = gseq / C.NS;
        if (use >= 1) mbar_wait_timed(&C.empty[s], (unsigned)(use - 1) & 1u);
        const int nb = min(G, N - g * G);
        const int b_lo = FWD ? g * G : N - g * G - nb;
        const int mo = FWD ? 0 : G - nb;
        const char *src = nullptr;
        unsigned bytes = 0, dst = 0;
        if (lane == 0) {
            src = reinterpret_cast<const char *>(C.sd + C.D.block(b_lo));
            bytes = nb * GL.blk_bytes;
            dst = mo * GL.blk_bytes;
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
            const int bs = FWD ? max(b_lo, d + 1) : b_lo;
            const int be = FWD ? b_lo + nb - 1 : min(b_lo + nb - 1, N - 2 - d);
            if (be >= bs) {
                const double *q = FWD ? C.sd + C.D.off_sub + C.D.diag_chunk(d, bs - 1 - d) : C.sd + C.D.off_sup + C.D.diag_chunk(d, bs);
                src = reinterpret_cast<const char *>(q);
                bytes = (be - bs + 1) * GL.ld8;
                dst = GL.o_diag + d * G * GL.ld8 + (mo + bs - b_lo) * GL.ld8;
            }
        }
        const unsigned total = __reduce_add_sync(FULL, bytes);
        if (lane == 0) mbar_expect_tx(&C.full[s], total);
        __syncwarp();
        if (bytes != 0) bulk_g2s(C.stages + (size_t)s * GL.bytes + dst, src, bytes, &C.full[s]);
    }
    if (p == 0) mbar_wait_timed(C.sweep_done, FWD ? 0u : 1u);  // watchdog of the consumers, as in quad32
}

// ---- consumers: lane = row, warp j owns sweep orders j, j+QW, ...; SECOND = my half comes second in sweep order ---------
template <bool FWD, int ND, bool SECOND, bool LD32>
__device__ void consume(const Ctx &C, int j, int r, int &nan_first, int &zero_last)
{
    const int N = C.D.nblocks, ld = LD32 ? W : C.D.ld;
    const GroupLayout &GL = C.GL;
    const int q = FWD ? r : W - 1 - r;        // position of my row in sweep order
    constexpr int K0 = SECOND ? H : 0;         // first sweep position of my half
    constexpr int G0 = K0 / PG, G1 = G0 + H / PG;  // my coupling groups
    const bool mine_half = (q >= K0) && (q < K0 + H);
    const int myquad = q / PG;
    const int ibase = FWD ? 0 : ((N + RS - 1) & ~(RS - 1));
    int *lead_word = C.scr + (FWD ? 0 : 2) * PW + j;  // written by the leading half's warp j
    int *foll_word = C.scr + (FWD ? 1 : 3) * PW + j;  // written by the following half's warp j
    int known = 0;       // blocks below `known` are finished by the other CTA's warp j (as far as I have looked)
    int count = 0;       // my blocks since the last publication
    double pv_next = 0.0;
    bool have_next = false;
    double *ytmp = C.ytmp + j * H;
    double ynext = 0.0;
    if (!FWD && j < N) ynext = C.x[(int64_t)(N - 1 - j) * W + r];
    int gcur = (FWD ? 0 : C.NG) + j / G;
    int s = gcur % C.NS;
    unsigned ph = (unsigned)(gcur / C.NS) & 1u;

    for (int i = j; i < N; i += QW) {
        const int b = FWD ? i : N - 1 - i;
        const double y_cur = ynext;
        if (!FWD && i + QW < N) ynext = C.x[(int64_t)(N - 1 - (i + QW)) * W + r];

        if (!SECOND && i - known >= LEAD) {
            // throttle: stay within LEAD blocks of the follower (keeps what I fetched in L2 for it)
            unsigned n = 0;
            while ((known = ld_volatile_g(foll_word)) + LEAD <= i)
                if (++n > SPIN_LIMIT) __trap();
        }
        const int mo = i % G;
        mbar_wait_unbounded(&C.full[s], ph);
        const int m = FWD ? mo : G - 1 - mo;
        const unsigned char *st = C.stages + (size_t)s * GL.bytes;
        const double *Ls = reinterpret_cast<const double *>(st + (size_t)m * GL.blk_bytes);
        double l[W];
#pragma unroll
        for (int k = 0; k < W; ++k) {
            const bool first_half_col = FWD ? (k < H) : (k >= H);  // column k sits in the first half of the sweep order
            l[k] = (SECOND || first_half_col) ? Ls[k * ld + r] : 0.0;
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
            if (mine_half && isnan(bv)) nan_first = min(nan_first, b * W + src + 1);
            asm volatile("add.rn.f64 %0, %1, %2;" : "=d"(acc) : "d"(bv), "d"(-0.0));  // quiets a signalling NaN (see quad32)
        } else {
            const double dg = Ls[r * ld + r];
            if (mine_half && dg == 0.0) zero_last = max(zero_last, b * W + r);
            const double rinv = rcp_fast_ok(dg) ? rcp_fast(dg) : 1.0 / dg;
            acc = y_cur * rinv;
#pragma unroll
            for (int k = 0; k < W; ++k) l[k] *= rinv;
#pragma unroll
            for (int d = 0; d < ND; ++d) cv[d] *= rinv;
        }
        __syncwarp();
        if (r == 0) {
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
        if (SECOND) {
            // the first half of this block, finished by the other CTA's warp j (long ago, normally: its progress word is
            // polled only when my cached copy does not cover the block); the 16 values are fetched one block ahead
            double pv = pv_next;
            if (!have_next) {
                if (known < i + 1) {
                    unsigned n = 0;
                    while ((known = ld_volatile_g(lead_word)) < i + 1)
                        if (++n > SPIN_LIMIT) __trap();
                    __threadfence();
                }
                pv = (r < H) ? __ldcg(C.x + (int64_t)b * W + (FWD ? r : H + r)) : 0.0;
            }
            have_next = (i + QW < N) && (known >= i + QW + 1);
            if (have_next) {
                const int bn = FWD ? i + QW : N - 1 - (i + QW);
                pv_next = (r < H) ? __ldcg(C.x + (int64_t)bn * W + (FWD ? r : H + r)) : 0.0;
            }
            if (r < H) ytmp[r] = pv;
            __syncwarp();
#pragma unroll
            for (int p = 0; p < H; ++p) {
                const int k = FWD ? p : W - 1 - p;
                acc = fma(-l[k], ytmp[FWD ? p : H - 1 - p], acc);
            }
        }

        double *myslot = C.ring + (i & (RS - 1)) * 32 + r;
        double *resetslot = C.ring + ((i - ND) & (RS - 1)) * 32 + r;
        const double *pd[ND];
#pragma unroll
        for (int d = 0; d < ND; ++d) pd[d] = C.ring + ((i - 1 - d) & (RS - 1)) * 32 + ci;
        double v[ND];
#pragma unroll
        for (int d = 0; d < ND; ++d) v[d] = 0.0;
        auto load_pred = [&](int qd) {
            if (myquad == qd) {
#pragma unroll
                for (int d = 0; d < ND; ++d) v[d] = ld_volatile_s(pd[d]);
            }
        };
        auto merge_quad = [&](int qd) {
            const bool mine = (myquad == qd);
            double t;
            for (;;) {
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
        mbar_wait_unbounded(&C.started[(i - 1) & (RS - 1)], (unsigned)((ibase + i - 1) >> 5) & 1u);
        load_pred(G0);
        merge_quad(G0);
        const double *slotbase = C.ring + (i & (RS - 1)) * 32;
#pragma unroll
        for (int kk = K0; kk < K0 + H; ++kk) {
            const int k = FWD ? kk : W - 1 - kk;
            const int g = kk / PG;
            if (kk % PG == 0 && g + 1 < G1) load_pred(g + 1);
            if (r == k) st_volatile_s(myslot, acc);
            if (kk + 1 < K0 + H) {
                const double vk = ld_volatile_s(slotbase + k);
                acc = fma(-l[k], vk, acc);
            }
            if (kk == K0 + PG - 1) {
                __syncwarp();
                if (r == 0) mbar_arrive(&C.started[i & (RS - 1)]);
            }
            if (kk % PG == PG - 1 && g + 1 < G1) {
                merge_quad(g + 1);
                if (g + 2 == G1) st_volatile_s(resetslot, __longlong_as_double((long long)SENT));
            }
        }
        if (mine_half) C.x[(int64_t)b * W + r] = ld_volatile_s(myslot);
        // tell the other CTA how far this warp is (every few blocks: the fence is not free), and at my last block
        if (++count == (SECOND ? PUB_FOLL : PUB_LEAD) || i + QW >= N) {
            count = 0;
            __syncwarp();
            if (r == 0) {
                __threadfence();
                st_volatile_g(SECOND ? foll_word : lead_word, i + 1);
            }
        }
    }
    __syncwarp();
    if (r == 0) mbar_arrive(C.sweep_done);
}

template <int ND, bool LD32>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__((QW + NP) * 32, 1)
solve_half32_kernel(const double *__restrict__ data, int64_t sys_stride, Dims D, const int *__restrict__ rowbycol,
                    const int *__restrict__ colbyrow, const double *__restrict__ bvec, double *xvec, VecLayout V,
                    int *__restrict__ flags, int NS, const int *__restrict__ status, int *__restrict__ scratch)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int vecid = blockIdx.x >> 1, half = blockIdx.x & 1;  // cluster = the two halves of one vector
    if (status[vecid] != 0) return;  // a row interchange crossed the middle of a block: quad32 solves this vector
    const int sys = vecid / V.nrhs, rhs = vecid % V.nrhs;
    const int N = D.nblocks;

    Ctx C;
    C.sd = data + (int64_t)sys * sys_stride;
    C.rbc = rowbycol + (int64_t)sys * N * W;
    C.cbr = colbyrow + (int64_t)sys * N * W;
    C.b = bvec + (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride;
    C.x = xvec + (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride;
    C.D = D;
    C.GL = group_layout(D.ld, ND);
    C.NS = NS;
    C.NG = (N + G - 1) / G;
    C.scr = scratch + (int64_t)vecid * SCR_INTS;
    C.stages = smem_raw;
    size_t off = (size_t)NS * C.GL.bytes;
    C.full = reinterpret_cast<unsigned long long *>(smem_raw + off);
    C.empty = C.full + NS;
    C.started = C.empty + NS;
    C.sweep_done = C.started + RS;
    off += (size_t)(2 * NS + RS + 1) * 8;
    C.ring = reinterpret_cast<double *>(smem_raw + off);
    off += (size_t)RS * 32 * 8;
    C.ytmp = reinterpret_cast<double *>(smem_raw + off);
    off += (size_t)QW * H * 8;
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
    else if (half == 0) consume<true, ND, false, LD32>(C, warp, lane, nan_first, zero_last);
    else consume<true, ND, true, LD32>(C, warp, lane, nan_first, zero_last);
    __syncthreads();
    ring_reset(C, tid, blockDim.x, ND);
    if (tid < RS && (N & (RS - 1)) != 0 && tid >= (N & (RS - 1))) mbar_arrive(&C.started[tid]);
    __syncthreads();
    if (warp >= QW) produce<false, ND>(C, warp - QW, lane);
    else if (half == 1) consume<false, ND, false, LD32>(C, warp, lane, nan_first, zero_last);
    else consume<false, ND, true, LD32>(C, warp, lane, nan_first, zero_last);
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
    if (tid == 0) {
        // the two CTAs of a vector combine their findings in its scratch; the second one to arrive writes the flag
        int *fw = C.scr + SCR_FLAGS;
        atomicMax(&fw[0], 0x7fffffff - C.flagw[0]);
        atomicMax(&fw[1], C.flagw[1] + 1);
        __threadfence();
        if (atomicAdd(&fw[2], 1) == 1 && flags) {
            __threadfence();
            const int nf = 0x7fffffff - ld_volatile_g(&fw[0]), zl = ld_volatile_g(&fw[1]) - 1;
            flags[vecid] = (nf != 0x7fffffff) ? nf : (zl >= 0 ? N * W + zl : 0);
        }
    }
}

// status[v] = 1 for the vectors of systems in which some block's colbyrow maps a row across the middle (or is damaged)
__global__ void __launch_bounds__(256)
scan_halves_kernel(const int *__restrict__ colbyrow, int64_t per_sys, int nrhs, int *__restrict__ status)
{
    const int sys = blockIdx.y;
    const int *c = colbyrow + (int64_t)sys * per_sys;
    bool bad = false;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < per_sys; t += (int64_t)gridDim.x * blockDim.x) {
        const unsigned ci = (unsigned)__ldg(c + t);
        bad |= (ci >= (unsigned)W) || ((ci < (unsigned)H) != ((t & (W - 1)) < H));
    }
    if (__any_sync(FULL, bad) && (threadIdx.x & 31) == 0)
        for (int rh = 0; rh < nrhs; ++rh) status[sys * nrhs + rh] = 1;
}

template <int ND>
static cudaError_t launch_nd(unsigned nvec, size_t smem, int NS, cudaStream_t st, const double *data, int64_t sys_stride,
                             const Dims &D, const int *rowbycol, const int *colbyrow, const double *b, double *x,
                             const VecLayout &V, int *flags, const int *status, int *scratch)
{
#define H32_GO(LD32)                                                                                                    \
    {                                                                                                                   \
        cudaError_t e = cudaFuncSetAttribute(solve_half32_kernel<ND, LD32>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                             227 * 1024);                                                               \
        if (e != cudaSuccess) return e;                                                                                 \
        solve_half32_kernel<ND, LD32><<<2 * nvec, (QW + NP) * 32, smem, st>>>(data, sys_stride, D, rowbycol, colbyrow, b, \
                                                                             x, V, flags, NS, status, scratch);         \
    }
    if (D.ld == W) H32_GO(true) else H32_GO(false)
#undef H32_GO
    return cudaGetLastError();
}

}  // namespace half32

// blockw == 32, 1 <= ndiag <= 4, nsat == 0, even ld, 16-byte aligned arrays (checked by try_launch_solve_wave).
// Solves the vectors whose systems satisfy the half condition and marks the others in `status` (int per vector,
// zeroed here) for quad32.  `scratch`: nvec * scratch_ints() ints.  Returns false when the shape does not fit.
int half32_scratch_ints() { return half32::SCR_INTS; }

bool try_launch_solve_half32(unsigned nvec, const double *data, int64_t sys_stride, int nsys, const Dims &D,
                             const int *rowbycol, const int *colbyrow, const double *b, double *x, const VecLayout &V,
                             int *flags, int *status, int *scratch, cudaStream_t st, cudaError_t *err)
{
    using namespace half32;
    const GroupLayout GL = group_layout(D.ld, D.ndiag);
    const size_t fixed = (size_t)RS * 32 * 8 + (size_t)(RS + 1) * 8 + (size_t)QW * H * 8 + 64;
    const size_t budget = 227 * 1024;
    int NS = (int)((budget - fixed) / (GL.bytes + 16));
    if (NS > 16) NS = 16;
    if (NS * G < QW + 2 * G) return false;
    const size_t smem = (size_t)NS * (GL.bytes + 16) + fixed;
    *err = cudaMemsetAsync(status, 0, sizeof(int) * (size_t)nvec, st);
    if (*err != cudaSuccess) return true;
    *err = cudaMemsetAsync(scratch, 0, sizeof(int) * (size_t)nvec * SCR_INTS, st);
    if (*err != cudaSuccess) return true;
    const int64_t per_sys = (int64_t)D.nblocks * W;
    {
        const int64_t want = (per_sys + 255) / 256;
        dim3 grid((unsigned)(want < 64 ? want : 64), (unsigned)nsys);
        scan_halves_kernel<<<grid, 256, 0, st>>>(colbyrow, per_sys, V.nrhs, status);
        *err = cudaGetLastError();
        if (*err != cudaSuccess) return true;
    }
    switch (D.ndiag) {
    case 1: *err = launch_nd<1>(nvec, smem, NS, st, data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, status, scratch); break;
    case 2: *err = launch_nd<2>(nvec, smem, NS, st, data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, status, scratch); break;
    case 3: *err = launch_nd<3>(nvec, smem, NS, st, data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, status, scratch); break;
    default: *err = launch_nd<4>(nvec, smem, NS, st, data, sys_stride, D, rowbycol, colbyrow, b, x, V, flags, status, scratch); break;
    }
    return true;
}

}  // namespace bdilu
