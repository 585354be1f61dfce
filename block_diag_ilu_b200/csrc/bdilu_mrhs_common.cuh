// Pieces shared by the many-vector solve kernels (bdilu_solve_mrhs.cu, bdilu_solve_mrhs_wave.cu): stage layout,
// lane layout, width-Q broadcast.
#pragma once
#include "bdilu_wave_common.cuh"

namespace bdilu {
namespace mrhs {

using wave::bulk_g2s;
using wave::mbar_arrive;
using wave::mbar_expect_tx;
using wave::mbar_init;
using wave::mbar_wait;

// Q = lanes sharing one vector (4, 8 or 16; template parameter), R = 32 / Q vectors per warp
constexpr int NS = 4;      // stages (blocks in flight)
constexpr int MAXD = 4;    // sub/super-diagonals handled here
constexpr int MAXWARPS = 4;

struct Layout {
    int blk_bytes, idx_bytes, diag_bytes;  // bulk-copy sizes
    int o_rbc, o_cbr, o_diag, stage_bytes;
};

static inline Layout make_layout(int w, int ld, int ndiag)
{
    Layout L;
    L.blk_bytes = w * ld * 8;
    L.idx_bytes = w * 4;
    L.diag_bytes = w * 8;
    L.o_rbc = L.blk_bytes;
    L.o_cbr = L.o_rbc + L.idx_bytes;
    L.o_diag = L.o_cbr + L.idx_bytes;
    L.stage_bytes = (L.o_diag + ndiag * L.diag_bytes + 127) & ~127;
    return L;
}

template <int WP, int Q>
struct Cfg {
    static_assert(WP >= 2 * Q, "every lane owns at least one row pair");
    static constexpr int R = 32 / Q;         // vectors per warp
    static constexpr int NP = WP / (2 * Q);  // row pairs per lane
    static constexpr int WPP = WP + 8;       // pitch of one vector in the coupling ring (bank spread)
};

template <int Q>
__device__ __forceinline__ double quad_bcast(double v, int src)
{
    int lo = __shfl_sync(FULL, __double2loint(v), src, Q);
    int hi = __shfl_sync(FULL, __double2hiint(v), src, Q);
    return __hiloint2double(hi, lo);
}

}  // namespace mrhs
}  // namespace bdilu
