// Kernel 2, blockw == 32, EXPERIMENTAL systolic form ("skew32"): ILU_inplace<double>::solve (block_diag_ilu.hpp:486-538)
// with lane = block.  Opt-in (test hook 8 of bdilu_debug_force_chain_solve); parity-tested, NOT the default: on B200
// it runs cfg3's sweeps in 4.3-6.1 ms against 2.65 ms of the wavefront-across-warps kernel (bdilu_solve_quad32.cu).
// DESIGN.md section 9 records the measurements and why: what bounds it is the shared-memory pipe of the one SM a
// chain lives on, the same 128 B/cycle that bounds quad32 -- not the instruction count this design was built to cut.
//
// The chain over the blocks is sequential (:502, :518), but row li of block B needs from its predecessors only
// y_{B-1-d}[colbyrow[li]] (:508-511, :524-527).  When the permutations are the identity (no row interchange was
// needed -- the diagonally dominant matrices this library is made for, README.rst:9-15) that is row li itself, so
// block B can run exactly ONE ROW behind block B-1.  A warp therefore holds 32 consecutive blocks, lane j on the
// blocks j, j+32, ..., lane j one row behind lane j-1; every lane keeps the partial sums of its block in registers,
// and one step of the warp finishes one row of each of the 32 blocks in flight:
//     y   = (b[q] + R[q]) - sub1[q]*y'' - sub0[q]*y'        y', y'': the values lanes j-1 / j-2 finished one and two
//     R[r] -= L[r][q] * y   for r > q                        steps ago -- two shuffles, no polling
// The register that holds "the current row" must have the same NAME in all lanes although the lanes sit on different
// rows: row r of lane j lives in R[(r + j) mod 32], so at step u the current row is R[u mod 32] everywhere and the
// 32 steps of the unrolled loop body are straight-line code with static register indices (~80 warp instructions per
// block and sweep instead of ~650).  Two consumer warps share a step: the NEAR warp owns the chain and the 16 rows
// behind the current one, the FAR warp the 15 rows beyond, two group steps behind, handing its sums over through a
// ring.
//
// Staging.  A lane reads column q of ITS block, so the blocks in flight have to sit in shared memory; 32 whole blocks
// do not fit, and the lanes need only the triangle.  Time is cut into group steps of 8 rows.  Four 8-column PANELS
// per block are needed for 8 steps each; panel p of the 8 blocks of group G is first needed in group step G + p, and
// all panels first needed in group step t share TIME SLOT t, together with the matching 8-row pieces of b / y, of the
// sub / sup rows and (backward) of 1/U[r][r].  The ring has 4 slots, so its period equals the 32 steps of the loop
// body and every shared-memory address a lane touches is a function of (step mod 32, lane) alone: a 1024-entry table,
// one LDS.64 per step, no address arithmetic.
// The slots are filled by four LOADER warps through registers (LDG.128 -> STS.64).  TMA was tried first and cannot do
// it: a request costs ~60 cycles of issue whatever its size, and tensor boxes with 64..288-byte rows ran at
// ~10 B/cycle (scratch/ub; the TMA version: git show 1850672:scratch/skew_tma_version.cu.txt).  The register path stores each column from the row
// below the diagonal on (backward: from the row above it, reversed), 16-byte aligned and skewed so that the consumers'
// LDS.128 are conflict free, with zero padding that never has to be rewritten.  Blocks outside the system are
// clamped to a real block and given zero right-hand sides and coefficients, which makes the ramp-up and the tail
// harmless without a special case in the loop.  HBM traffic: the triangle of every block per sweep (53 %).
//
// Scope: identity permutations, finite b, U diagonals inside the exact range of the fast reciprocal.  Everything else
// (a system that pivoted, NaN in b, a zero diagonal) is handed back through `status` and solved by the general
// wavefront kernel in a second launch, which also produces the reference's flag values for those cases.
#include <algorithm>

#include "bdilu_wave_common.cuh"

namespace bdilu {

namespace skew {

using namespace wave;  // mbarrier helpers

constexpr int GB = 8;       // blocks per group
constexpr int NSLOT = 4;    // time slots
constexpr int MAXND = 2;
constexpr int NLOAD = 4;    // loader warps
constexpr int NLANES = NLOAD * 32;
constexpr int PFG = 6;      // L2 prefetch distance in groups, counted from the group being loaded

__device__ unsigned long long g_stats[3];  // vectors solved here / handed back at the end / skipped at the start

// Staged layout of one slot, in 16-byte chunks.  Panel p: 8 blocks x 8 columns; a column holds CH(p) = 16 - 4p
// chunks (the rows from the diagonal of the panel's first column on), odd panels one chunk more, and the blocks of
// even panels are skewed by 5 chunks: with these numbers the eight lanes of a quarter warp -- eight consecutive
// blocks, one row apart -- always hit eight different 16-byte bank groups (exhaustive search).
__host__ __device__ constexpr int pan_ch(int p) { return 16 - 4 * p; }                         // chunks loaded per column
__host__ __device__ constexpr int pan_cc(int p) { return pan_ch(p) + (p & 1); }                 // chunks per staged column
__host__ __device__ constexpr int pan_bc(int p) { return 8 * pan_cc(p) + ((p & 1) ? 0 : 5); }   // chunks per staged block
__host__ __device__ constexpr int pan_off(int p)  // bytes before panel p: 16 * GB * (sum of pan_bc)
{
    return p == 0 ? 0 : p == 1 ? 17024 : p == 2 ? 30336 : p == 3 ? 39168 : 44288;
}
static_assert(pan_off(1) == 16 * GB * pan_bc(0) && pan_off(2) == pan_off(1) + 16 * GB * pan_bc(1) &&
                  pan_off(3) == pan_off(2) + 16 * GB * pan_bc(2) && pan_off(4) == pan_off(3) + 16 * GB * pan_bc(3),
              "panel offsets");
constexpr int PAN_BYTES = pan_off(4);   // 44288
constexpr int VEC_BYTES = 4 * GB * 64;  // one 8-row piece per panel and block: [panel][block][row] doubles
constexpr int NARR = 2 + MAXND;         // b / y / x, sub or sup rows, reciprocal diagonals
constexpr int RINV_OFF = (1 + MAXND) * VEC_BYTES;
constexpr int SLOT_BYTES = PAN_BYTES + NARR * VEC_BYTES;  // 52480
constexpr int TAB_BYTES = 32 * 32 * 8;
static_assert(SLOT_BYTES % 128 == 0, "slot size");

struct SkewArgs {
    const double *data, *b;
    double *x;
    int64_t sys_stride, rhs_stride, vec_sys_stride;
    int64_t off_sub[MAXND], off_sup[MAXND];  // sub(d, 0, 0) / sup(d, 0, 0) inside one system
    int N, nrhs, ld;
};

__device__ __forceinline__ double lds64(unsigned a)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
    return v;
}
template <int OFF>
__device__ __forceinline__ void lds128o(unsigned a, double &x, double &y)
{
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2+%3];" : "=d"(x), "=d"(y) : "r"(a), "n"(OFF));
}
__device__ __forceinline__ void sts64(unsigned a, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v) : "memory"); }
__device__ __forceinline__ uint2 lds_u2(unsigned a)
{
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
    return v;
}
// 16-byte global load that does not allocate in L1 (with 227 KB of shared memory the L1 has no lines to give)
__device__ __forceinline__ double2 ldg_na(const char *p)
{
    double2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
}

// A wait that cannot hang: a protocol failure ends in a trap = cudaErrorLaunchFailure at the next synchronisation,
// which the C ABI reports as BDILU_ERR_CUDA.
__device__ __forceinline__ unsigned mbar_test(unsigned bar, unsigned parity)
{
    unsigned ok;
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(ok)
                 : "r"(bar), "r"(parity)
                 : "memory");
    return ok;
}
__device__ __forceinline__ void mbar_wait_bounded(unsigned bar, unsigned parity)
{
    for (int spin = 0; !mbar_test(bar, parity); ++spin)
        if (spin > (1 << 21)) __trap();
}
__device__ __forceinline__ void mbar_arrive_a(unsigned bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }

// lanes whose row position q = (u - lane) mod 32 is <= qmax
__host__ __device__ constexpr unsigned lanes_q_le(int u, int qmax)
{
    unsigned m = 0;
    for (int j = 0; j < 32; ++j)
        if (((u - j) & 31) <= qmax) m |= 1u << j;
    return m;
}

// ---- consumers ---------------------------------------------------------------------------------------------------
// Two warps share the work of a step.  The NEAR warp owns the chain: it finishes the current row and updates the 16
// rows right behind it.  The FAR warp updates rows 17..31 behind; those sums are not needed for another 16 steps, so
// it runs up to two group steps (8..16 rows) behind the near warp, reads the values the near warp left in the row
// arrays, and hands every row's far part over through a 32-entry ring (`hand`) one step before the row enters the
// near band.  The two meet only at group-step granularity (mbarriers `near_done` / `far_done`); the far warp is the
// last reader of a slot and releases it.
struct Cons {
    double R[32];     // partial sums: row r of my block in R[(r + lane) & 31]
    double v1;        // near: the value I finished one step ago
    double s1, s2;    // near: what lanes j-1 / j-2 finished one / two steps ago
    unsigned col, vec;  // shared-memory addresses of this step (from the table)
    unsigned zcol;    // a column of zeros
    unsigned tab;     // my table column
    unsigned hand;    // my column of the hand-over ring
    unsigned lanebit;
    int src1, src2;
    unsigned okfull;
    bool last;        // last round of the sweep
    unsigned bar_full, bar_empty, bar_near, bar_far;  // addresses of full[0], empty[0], near_done[0], far_done[0]
};

__device__ __forceinline__ double shfl_d(double v, int src)
{
    const int lo = __shfl_sync(FULL, __double2loint(v), src);
    const int hi = __shfl_sync(FULL, __double2hiint(v), src);
    return __hiloint2double(hi, lo);
}

// coefficients of the 4 rows 4g .. 4g+3 behind the row of step U in sweep order (two 16-byte chunks).  Lanes whose
// block has no such rows left read a zero area instead (one select per group, no predicated arithmetic): their
// updates add 0.  The rows a valid group reads beyond the block are zero padding.
template <int U, int G>
__device__ __forceinline__ void group_load(const Cons &C, double (&a)[32])
{
    constexpr unsigned MASK = lanes_q_le(U, 30 - 4 * G);
    const unsigned base = (C.lanebit & MASK) ? C.col : C.zcol;
    lds128o<32 * G>(base, a[4 * G], a[4 * G + 1]);
    lds128o<32 * G + 16>(base, a[4 * G + 2], a[4 * G + 3]);
}

constexpr int NEARG = 4;  // groups of 4 rows in the near band: rows 1..16 behind the current one

// the slot of the next group step: tested early, waited for one step before its first use
template <int U>
__device__ __forceinline__ void full_wait(Cons &C, int M)
{
    if constexpr ((U & 7) == 3 || (U & 7) == 7) {
        constexpr int SLN = ((U >> 3) + 1) & 3;
        const unsigned par = ((U >> 3) == 3) ? ((unsigned)(M + 1) & 1u) : ((unsigned)M & 1u);
        const bool more = ((U >> 3) != 3) || !C.last;  // nothing is staged beyond the last round
        if (more) {
            if constexpr ((U & 7) == 3) C.okfull = mbar_test(C.bar_full + 8 * SLN, par);  // early, non-blocking
            else if (!C.okfull) mbar_wait_bounded(C.bar_full + 8 * SLN, par);
        }
    }
}

template <bool FWD, int ND, int U>
__device__ __forceinline__ void near_step(Cons &C, int M)
{
    if constexpr ((U & 7) == 0) {
        // the far parts I am about to consume were handed over during the far warp's group step k-2 at the latest
        const int k = 4 * M + (U >> 3);
        if (k >= 2) mbar_wait_bounded(C.bar_far + 8 * (((U >> 3) + 2) & 3), (unsigned)((k - 2) >> 2) & 1u);
    }
    full_wait<U>(C, M);
    const uint2 nxt = lds_u2(C.tab + ((U + 1) & 31) * 256);

    double a[32];
    group_load<U, 0>(C, a);
    group_load<U, 1>(C, a);
    group_load<U, 2>(C, a);
    group_load<U, 3>(C, a);

    const double bv = lds64(C.vec);
    const double hv = lds64(C.hand + U * 256);  // far part of this row
    double cf[ND];
#pragma unroll
    for (int d = 0; d < ND; ++d) cf[d] = lds64(C.vec + VEC_BYTES * (1 + d));
    double rinv = 1.0;
    if constexpr (!FWD) rinv = lds64(C.vec + RINV_OFF);  // 1/U[li][li], formed by the loaders (0 outside the system)

    double t = (bv + hv) + C.R[U];
    if constexpr (ND == 2) t = fma(-cf[1], C.s2, t);
    t = fma(-cf[0], C.s1, t);
    const double v = FWD ? t : t * rinv;
    C.R[U] = 0.0;
    // hand my value to the lanes behind me (used in their next step)
    C.s1 = shfl_d(v, C.src1);
    if constexpr (ND == 2) C.s2 = shfl_d(C.v1, C.src2);
    sts64(C.vec, v);
    C.v1 = v;
    const double nv = -v;
#pragma unroll
    for (int sidx = 0; sidx < 4 * NEARG; ++sidx) C.R[(U + 1 + sidx) & 31] = fma(a[sidx], nv, C.R[(U + 1 + sidx) & 31]);

    C.col = nxt.x;
    C.vec = nxt.y;
    if constexpr ((U & 7) == 7) {  // group step done: the far warp may read my values
        __syncwarp();
        if (C.lanebit == 1u) mbar_arrive_a(C.bar_near + 8 * (U >> 3));
    }
}

template <int U>
__device__ __forceinline__ void far_step(Cons &C, int M)
{
    if constexpr ((U & 7) == 0) {
        // the previous group step is finished: its older slot is free (I am its last reader); then wait for the near
        // warp to have finished this group step
        __syncwarp();
        if (C.lanebit == 1u) mbar_arrive_a(C.bar_empty + 8 * (((U >> 3) + 2) & 3));
        mbar_wait_bounded(C.bar_near + 8 * (U >> 3), (unsigned)M & 1u);
        mbar_wait_bounded(C.bar_full + 8 * (U >> 3), (unsigned)M & 1u);  // my own acquire of the slot (long complete)
    }
    const uint2 nxt = lds_u2(C.tab + ((U + 1) & 31) * 256);
    double a[32];
    group_load<U, 4>(C, a);
    group_load<U, 5>(C, a);
    group_load<U, 6>(C, a);
    group_load<U, 7>(C, a);
    const double nv = -lds64(C.vec);  // the value the near warp finished at its step U
#pragma unroll
    for (int sidx = 4 * NEARG; sidx < 31; ++sidx) C.R[(U + 1 + sidx) & 31] = fma(a[sidx], nv, C.R[(U + 1 + sidx) & 31]);
    // the row 17 behind enters the near band at the next step: its far part is complete
    sts64(C.hand + ((U + 1 + 4 * NEARG) & 31) * 256, C.R[(U + 1 + 4 * NEARG) & 31]);
    C.R[(U + 1 + 4 * NEARG) & 31] = 0.0;
    C.col = nxt.x;
    C.vec = nxt.y;
    if constexpr ((U & 7) == 7) {
        __syncwarp();
        if (C.lanebit == 1u) mbar_arrive_a(C.bar_far + 8 * (U >> 3));
    }
}

template <bool FWD, int ND, bool NEAR, int U>
struct Steps {
    static __device__ __forceinline__ void run(Cons &C, int M)
    {
        if constexpr (NEAR) near_step<FWD, ND, U>(C, M);
        else far_step<U>(C, M);
        Steps<FWD, ND, NEAR, U + 1>::run(C, M);
    }
};
template <bool FWD, int ND, bool NEAR>
struct Steps<FWD, ND, NEAR, 32> {
    static __device__ __forceinline__ void run(Cons &, int) {}
};

struct ConsArgs {
    unsigned zero_area, tabbase, handbase, bar_full, bar_empty, bar_near, bar_far;
    int Mtot;
};

template <bool FWD, int ND, bool NEAR>
__device__ __forceinline__ void consume(const ConsArgs &A, int lane)
{
    Cons C;
#pragma unroll
    for (int r = 0; r < 32; ++r) C.R[r] = 0.0;
    C.v1 = C.s1 = C.s2 = 0.0;
    C.tab = A.tabbase + lane * 8;
    C.hand = A.handbase + lane * 8;
    C.lanebit = 1u << lane;
    C.src1 = (lane + 31) & 31;
    C.src2 = (lane + 30) & 31;
    C.okfull = 0;
    C.bar_full = A.bar_full; C.bar_empty = A.bar_empty; C.bar_near = A.bar_near; C.bar_far = A.bar_far;
    C.zcol = A.zero_area;
    const uint2 e0 = lds_u2(C.tab);
    C.col = e0.x;
    C.vec = e0.y;
    if (NEAR) mbar_wait_bounded(A.bar_full, 0);  // slot of group step 0
    for (int M = 0; M < A.Mtot; ++M) {
        C.last = (M == A.Mtot - 1);
        Steps<FWD, ND, NEAR, 0>::run(C, M);
    }
    if (!NEAR) {  // the two slots still held
        __syncwarp();
        if (lane == 0) {
            mbar_arrive_a(A.bar_empty + 8 * 2);
            mbar_arrive_a(A.bar_empty + 8 * 3);
        }
    }
}

// ---- loaders ------------------------------------------------------------------------------------------------------
struct Load {
    unsigned sbase, bar_full, bar_empty;
    int N, ld;
    const double *blocks;                   // this system's packed data (block 0)
    const double *b;                        // this vector's right-hand side
    double *x;                              // this vector's solution (forward: y out; backward: y in, x out)
    const double *sub[MAXND], *sup[MAXND];  // this system's sub(d, 0, 0) / sup(d, 0, 0)
    int *puntw;                             // shared: set when the vector has to go to the general kernel
};

// Work split of the 128 loader lanes over one panel of one group (8 blocks x 8 columns x CH(p) chunks): a lane keeps
// its (column f, chunk mm) for the whole sweep and walks over the blocks, so everything but the block index is a
// per-lane constant.  Panel 0: 8 x 16 lanes per block, 8 iterations; panel 1: 8 x 12 = 96 lanes (32 idle), 8
// iterations; panel 2: 2 blocks x 8 x 8 per iteration, 4 iterations; panel 3: 4 x 8 x 4, 2 iterations: 22 chunks.
constexpr int NITEM = 22;
__host__ __device__ constexpr int pan_iters(int p) { return p < 2 ? 8 : p == 2 ? 4 : 2; }
__host__ __device__ constexpr int pan_bpi(int p) { return 8 / pan_iters(p); }  // blocks per iteration

struct LaneMap {
    const char *src[4];  // my chunk in block 0 of the system
    unsigned soff[4];    // bytes from the start of a slot to where row r0 = 2m of my chunk goes (my first block)
    int e0[4];           // my first block inside a group
    unsigned flags;      // bit 3p: row r0 of my chunk is needed, 3p+1: row r1, 3p+2: I prefetch my chunk's line
};

template <bool FWD>
__device__ __forceinline__ LaneMap make_lane_map(const double *blocks, int gl, int ld)
{
    LaneMap M;
    M.flags = 0;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        int e0, f, mm;
        bool act = true;
        if (p == 0) { e0 = 0; f = (gl >> 4) & 7; mm = gl & 15; }
        else if (p == 1) { e0 = 0; f = gl / 12; mm = gl - 12 * f; act = gl < 96; f &= 7; }
        else if (p == 2) { e0 = (gl >> 6) & 1; f = (gl >> 3) & 7; mm = gl & 7; }
        else { e0 = (gl >> 5) & 3; f = (gl >> 2) & 7; mm = gl & 3; }
        const int col = FWD ? 8 * p + f : 31 - 8 * p - f;
        const int m = FWD ? 4 * p + mm : mm;
        const int r0 = 2 * m, r1 = 2 * m + 1;
        // forward: rows below the diagonal, row r at position r - col - 1 of the staged column; backward: rows above
        // it, reversed: position col - 1 - r
        const bool n0 = act && (FWD ? r0 > col : r0 < col), n1 = act && (FWD ? r1 > col : r1 < col);
        const bool pf = (n0 || n1) && ((m & 7) == 0 || mm == 0);  // one lane per 128-byte line
        M.flags |= ((n0 ? 1u : 0u) | (n1 ? 2u : 0u) | (pf ? 4u : 0u)) << (3 * p);
        M.src[p] = reinterpret_cast<const char *>(blocks + (col * ld + r0));
        M.soff[p] = (unsigned)(pan_off(p) + (e0 * pan_bc(p) + f * pan_cc(p)) * 16 + 8 * (FWD ? r0 - col - 1 : col - 1 - r0));
        M.e0[p] = e0;
    }
    return M;
}

// what one loader lane holds for one time slot: 22 chunks of panel data and two rows of each small array
struct Set {
    double2 pan[NITEM];
    double sm[NARR][2];
};

// memory block of sweep-order block i
template <bool FWD>
__device__ __forceinline__ int mem_block(const Load &L, int i) { return FWD ? i : L.N - 1 - i; }

// Blocks outside the system (ramp-up, tail, partial groups) are CLAMPED to a real block: their right-hand sides,
// coupling coefficients and reciprocal diagonals are zeros (small arrays below), so their values are exactly 0
// whatever finite coefficients they are given, and nothing they compute reaches a real block.
template <bool FWD, int ND>
__device__ __forceinline__ void load_slot(const Load &L, const LaneMap &M, int tau, int gl, Set &S)
{
    const int bsb = 256 * L.ld;  // bytes per block
    const int nm1 = L.N - 1;
    int n = 0;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const unsigned fl = M.flags >> (3 * p);
        if (fl & 3u) {
#pragma unroll
            for (int i = 0; i < pan_iters(p); ++i) {
                const int blk = mem_block<FWD>(L, 8 * (tau - p) + pan_bpi(p) * i + M.e0[p]);
                const int bc = min(max(blk, 0), nm1);
                S.pan[n + i] = ldg_na(M.src[p] + (int64_t)bc * bsb);
                if (fl & 4u) {  // the same line PFG groups further along the sweep -> L2
                    const int b2 = min(max(FWD ? blk + 8 * PFG : blk - 8 * PFG, 0), nm1);
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(M.src[p] + (int64_t)b2 * bsb));
                }
            }
        }
        n += pan_iters(p);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int t = i * NLANES + gl;
        const int p = t >> 6, e = (t >> 3) & 7, f = t & 7;
        const int blk = mem_block<FWD>(L, 8 * (tau - p) + e);
        const int row = FWD ? 8 * p + f : 31 - 8 * p - f;
        const bool in = (unsigned)blk < (unsigned)L.N;
        // y was written by this kernel's forward sweep: L2-coherent load, not the read-only path
        double v = 0.0;
        if (in) v = FWD ? __ldg(L.b + (int64_t)blk * 32 + row) : __ldcg(L.x + (int64_t)blk * 32 + row);
        S.sm[0][i] = v;
#pragma unroll
        for (int d = 0; d < ND; ++d) {
            // forward: sub(d, blk-1-d, row) couples block blk to y_{blk-1-d}; backward: sup(d, blk, row) to x_{blk+1+d}
            const int cr = FWD ? blk - 1 - d : blk;
            const bool cin = in && cr >= 0 && cr <= L.N - 2 - d;
            S.sm[1 + d][i] = cin ? __ldg((FWD ? L.sub[d] : L.sup[d]) + (int64_t)cr * L.ld + row) : 0.0;
        }
        if constexpr (!FWD) S.sm[1 + MAXND][i] = in ? __ldg(L.blocks + ((int64_t)blk * 32 + row) * L.ld + row) : 0.0;
    }
}

template <bool FWD, int ND>
__device__ __forceinline__ void store_slot(const Load &L, const LaneMap &M, int tau, int gl, const Set &S)
{
    const unsigned slot = L.sbase + (tau & 3) * SLOT_BYTES;
    int n = 0;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const unsigned fl = M.flags >> (3 * p);
        const unsigned d0 = slot + M.soff[p];
#pragma unroll
        for (int i = 0; i < pan_iters(p); ++i) {
            const unsigned dst = d0 + i * (pan_bpi(p) * pan_bc(p) * 16);
            if (fl & 1u) sts64(dst, S.pan[n + i].x);
            if (fl & 2u) sts64(FWD ? dst + 8 : dst - 8, S.pan[n + i].y);
        }
        n += pan_iters(p);
    }
    bool punt = false;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int t = i * NLANES + gl;
        const unsigned dst = slot + PAN_BYTES + t * 8;
        const double v = S.sm[0][i];
        if (FWD) punt |= (v != v);  // NaN in b: the general kernel reports the reference's flag
        sts64(dst, v);
#pragma unroll
        for (int d = 0; d < ND; ++d) sts64(dst + VEC_BYTES * (1 + d), S.sm[1 + d][i]);
        if constexpr (!FWD) {
            // reciprocal of the U diagonal (0 outside the system: see load_slot).  Zero, denormal, huge or NaN
            // diagonals (outside the exact range of the fast reciprocal): general kernel.
            const int p = t >> 6, e = (t >> 3) & 7;
            const int blk = mem_block<FWD>(L, 8 * (tau - p) + e);
            const bool in = (unsigned)blk < (unsigned)L.N;
            const double dg = S.sm[1 + MAXND][i];
            punt |= in && !rcp_fast_ok(dg);
            sts64(dst + RINV_OFF, in ? rcp_fast(dg) : 0.0);
        }
    }
    if (punt) *L.puntw = 1;
}

// finished values of time slot tau -> x
template <bool FWD>
__device__ __forceinline__ void drain_slot(const Load &L, int tau, int gl)
{
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int t = i * NLANES + gl;
        const int p = t >> 6, e = (t >> 3) & 7, f = t & 7;
        const int blk = mem_block<FWD>(L, 8 * (tau - p) + e);
        const int row = FWD ? 8 * p + f : 31 - 8 * p - f;
        const double v = lds64(L.sbase + (tau & 3) * SLOT_BYTES + PAN_BYTES + t * 8);
        if ((unsigned)blk < (unsigned)L.N) L.x[(int64_t)blk * 32 + row] = v;
    }
}

// One loader warp.  The data of slot tau is loaded into registers one group step before the slot is released (with
// the L2 prefetch PFG groups ahead the loads are L2 hits and land within that group step) and stored as soon as the
// consumer lets go of the slot.
template <bool FWD, int ND>
__device__ __forceinline__ void loader(const Load &L, int Mtot, int gl, int lane)
{
    const int kfill = 4 * Mtot;  // group steps of the sweep (>= 4)
    const LaneMap M = make_lane_map<FWD>(L.blocks, gl, L.ld);
    Set S;
    for (int tau = 0; tau < 2; ++tau) {  // slots 0 and 1 are free from the start
        load_slot<FWD, ND>(L, M, tau, gl, S);
        store_slot<FWD, ND>(L, M, tau, gl, S);
        __syncwarp();
        if (lane == 0) mbar_arrive_a(L.bar_full + 8 * tau);
    }
    load_slot<FWD, ND>(L, M, 2, gl, S);
    for (int k = 0; k <= kfill + 1; ++k) {
        mbar_wait_bounded(L.bar_empty + 8 * ((k + 2) & 3), (unsigned)(k >> 2) & 1u);
        if (k >= 2) drain_slot<FWD>(L, k - 2, gl);
        if (k + 2 < kfill) {
            store_slot<FWD, ND>(L, M, k + 2, gl, S);
            __syncwarp();
            if (lane == 0) mbar_arrive_a(L.bar_full + 8 * ((k + 2) & 3));
            if (k + 3 < kfill) load_slot<FWD, ND>(L, M, k + 3, gl, S);
        }
    }
}

// ---- kernel --------------------------------------------------------------------------------------------------------
// table[u][lane] = {address of my staged column, address of my row of the small arrays} at step u of a round; the
// same for both sweeps (the loaders stage everything in sweep order)
__device__ __forceinline__ void build_table(unsigned char *tabp, unsigned sbase, int tid, int nthreads)
{
    uint2 *tab = reinterpret_cast<uint2 *>(tabp);
    for (int idx = tid; idx < 1024; idx += nthreads) {
        const int u = idx >> 5, j = idx & 31;
        const int q = (u - j) & 31, p = q >> 3, f = q & 7, e = j & 7, a = j >> 3;
        const unsigned slot = sbase + ((a + p) & 3) * SLOT_BYTES;
        tab[idx] = make_uint2(slot + pan_off(p) + (e * pan_bc(p) + f * pan_cc(p)) * 16,
                              slot + PAN_BYTES + (p * 64 + e * 8 + f) * 8);
    }
}

// Eight warps, two per scheduler: warp 0 = near consumer, warp 1 = far consumer, warps 2, 3, 6, 7 = loaders (they share
// schedulers 2 and 3, away from the consumers), warps 4 and 5 idle.  256 threads x 255 registers fit the register file.
constexpr int NWARPS = 8;
constexpr int HAND_BYTES = 32 * 32 * 8;

template <int ND>
__global__ void __launch_bounds__(NWARPS * 32, 1)
solve_skew32_kernel(const SkewArgs A, int *__restrict__ flags, int *__restrict__ status)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned char *smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int vecid = blockIdx.x;
    if (status[vecid] != 0) {  // permutations not the identity: left to the general kernel
        if (tid == 0) atomicAdd(&g_stats[2], 1ull);
        return;
    }
    const int N = A.N;
    const int sys = vecid / A.nrhs, rhs = vecid % A.nrhs;
    unsigned char *tabp = smem + NSLOT * SLOT_BYTES;
    double *hand = reinterpret_cast<double *>(tabp + TAB_BYTES);                       // [32 rows][32 lanes]
    unsigned long long *bars = reinterpret_cast<unsigned long long *>(tabp + TAB_BYTES + HAND_BYTES);
    // bars: full[4], empty[4], near_done[4], far_done[4]
    double *zeros = reinterpret_cast<double *>(bars + 16);                             // 32 doubles
    int *puntw = reinterpret_cast<int *>(zeros + 32);
    const unsigned sbase = smem_u32(smem);
    const int Mtot = (N + 31 + 31) / 32;  // rounds of 32 steps: N + 31 steps in all

    ConsArgs CA;
    CA.zero_area = smem_u32(zeros); CA.tabbase = smem_u32(tabp); CA.handbase = smem_u32(hand);
    CA.bar_full = smem_u32(bars); CA.bar_empty = smem_u32(bars + 4);
    CA.bar_near = smem_u32(bars + 8); CA.bar_far = smem_u32(bars + 12);
    CA.Mtot = Mtot;

    Load L;
    L.sbase = sbase; L.bar_full = CA.bar_full; L.bar_empty = CA.bar_empty;
    L.N = N; L.ld = A.ld;
    L.blocks = A.data + (int64_t)sys * A.sys_stride;
    L.b = A.b + (int64_t)sys * A.vec_sys_stride + (int64_t)rhs * A.rhs_stride;
    L.x = A.x + (int64_t)sys * A.vec_sys_stride + (int64_t)rhs * A.rhs_stride;
#pragma unroll
    for (int d = 0; d < MAXND; ++d) {
        L.sub[d] = L.blocks + A.off_sub[d];
        L.sup[d] = L.blocks + A.off_sup[d];
    }
    L.puntw = puntw;

    // everything zero once: the padding of the staged columns is never written again, and slot 3 stands for "the
    // group before the first one" during the first group step
    for (int t = tid; t < (NSLOT * SLOT_BYTES) / 16; t += NWARPS * 32) reinterpret_cast<uint4 *>(smem)[t] = make_uint4(0, 0, 0, 0);
    if (tid < 32) zeros[tid] = 0.0;
    if (tid == 0) *puntw = 0;
    build_table(tabp, sbase, tid, NWARPS * 32);
    const bool is_loader = (warp & 2) != 0;                       // warps 2, 3, 6, 7
    const int gl = ((warp >> 2) * 2 + (warp & 1)) * 32 + lane;    // loader lane 0..127

#pragma unroll 1
    for (int sweep = 0; sweep < 2; ++sweep) {
        // fresh barriers (full: one arrival per loader warp; the others: one warp), no far parts yet, and the y
        // values of the forward sweep (global memory) visible before the backward sweep reads them
        if (tid == 0) {
            for (int s = 0; s < 16; ++s) {
                if (sweep) asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(CA.bar_full + 8 * s) : "memory");
                mbar_init(bars + s, s < 4 ? NLOAD : 1);
            }
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        for (int t = tid; t < 32 * 32; t += NWARPS * 32) hand[t] = 0.0;
        __threadfence();
        __syncthreads();
        if (warp == 0) {
            if (sweep == 0) consume<true, ND, true>(CA, lane);
            else consume<false, ND, true>(CA, lane);
        } else if (warp == 1) {
            if (sweep == 0) consume<true, ND, false>(CA, lane);
            else consume<false, ND, false>(CA, lane);
        } else if (is_loader) {
            if (sweep == 0) loader<true, ND>(L, Mtot, gl, lane);
            else loader<false, ND>(L, Mtot, gl, lane);
        }
        __syncthreads();
    }
    if (tid == 0) {
        const bool any = *puntw != 0;
        if (any) status[vecid] = 2;  // redo on the general kernel (it also writes the flag)
        else if (flags) flags[vecid] = 0;
        atomicAdd(&g_stats[any ? 1 : 0], 1ull);
    }
}

// status[v] = 1 for every vector of a system whose colbyrow is not the identity (rowbycol is its inverse)
__global__ void perm_scan_kernel(const int *__restrict__ colbyrow, int64_t n_per_sys, int nrhs, int *__restrict__ status)
{
    const int sys = blockIdx.y;
    const int4 *p = reinterpret_cast<const int4 *>(colbyrow + (int64_t)sys * n_per_sys);
    const int64_t n4 = n_per_sys >> 2;
    bool bad = false;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const int4 c = __ldg(p + i);
        const int e0 = (int)((i * 4) & 31);
        bad |= (c.x != e0) | (c.y != e0 + 1) | (c.z != e0 + 2) | (c.w != e0 + 3);
    }
    if (__any_sync(FULL, bad) && (threadIdx.x & 31) == 0)
        for (int r = 0; r < nrhs; ++r) status[(int64_t)sys * nrhs + r] = 1;
}

}  // namespace skew

cudaError_t skew_stats(unsigned long long out[3], bool reset)
{
    cudaError_t e = cudaMemcpyFromSymbol(out, skew::g_stats, sizeof(unsigned long long) * 3);
    if (e == cudaSuccess && reset) {
        const unsigned long long z[3] = {0, 0, 0};
        e = cudaMemcpyToSymbol(skew::g_stats, z, sizeof z);
    }
    return e;
}

bool try_launch_solve_skew32(unsigned nvec, const double *data, int64_t sys_stride, const Dims &D, const int *rowbycol,
                             const int *colbyrow, const double *b, double *x, const VecLayout &V, int *flags, int *status,
                             cudaStream_t st, cudaError_t *err)
{
    using namespace skew;
    (void)rowbycol;
    *err = cudaSuccess;
    if (D.w != 32 || D.nsat != 0 || D.ndiag < 1 || D.ndiag > MAXND || (D.ld & 1) || D.nblocks < 8) return false;
    if ((sys_stride & 1) || (V.rhs_stride & 1) || (V.sys_stride & 1)) return false;
    if ((reinterpret_cast<uintptr_t>(data) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(x) |
         reinterpret_cast<uintptr_t>(colbyrow)) & 15)
        return false;
    const int nsys = (int)(nvec / V.nrhs);
    if (nsys > 65535) return false;
    SkewArgs A;
    A.data = data; A.b = b; A.x = x;
    A.sys_stride = sys_stride; A.rhs_stride = V.rhs_stride; A.vec_sys_stride = V.sys_stride;
    for (int d = 0; d < MAXND; ++d) {
        const int dd = d < D.ndiag ? d : 0;
        A.off_sub[d] = D.off_sub + D.diag_chunk(dd, 0);
        A.off_sup[d] = D.off_sup + D.diag_chunk(dd, 0);
    }
    A.N = D.nblocks; A.nrhs = V.nrhs; A.ld = D.ld;

    // which systems pivoted
    *err = cudaMemsetAsync(status, 0, sizeof(int) * nvec, st);
    if (*err != cudaSuccess) return true;
    const int64_t n_per_sys = (int64_t)D.nblocks * 32;
    const unsigned gx = (unsigned)std::min<int64_t>((n_per_sys / 4 + 255) / 256, 64);
    perm_scan_kernel<<<dim3(gx, nsys), 256, 0, st>>>(colbyrow, n_per_sys, V.nrhs, status);
    const size_t smem = (size_t)NSLOT * SLOT_BYTES + TAB_BYTES + HAND_BYTES + 128 + 256 + 16 + 128;
    if (D.ndiag == 1) {
        *err = cudaFuncSetAttribute(solve_skew32_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (*err != cudaSuccess) return true;
        solve_skew32_kernel<1><<<nvec, NWARPS * 32, smem, st>>>(A, flags, status);
    } else {
        *err = cudaFuncSetAttribute(solve_skew32_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (*err != cudaSuccess) return true;
        solve_skew32_kernel<2><<<nvec, NWARPS * 32, smem, st>>>(A, flags, status);
    }
    count_launch(2);
    *err = cudaGetLastError();
    return true;
}

}  // namespace bdilu
