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
// ~10 B/cycle (scratch/ub, scratch/skew_tma_version.cu.txt).  The register path stores each column from the row
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

__device__ unsigned long long g_stats[3];

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
