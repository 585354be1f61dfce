// Shared device-side helpers for the bdilu kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bdilu {

constexpr unsigned FULL = 0xffffffffu;

// Packed-layout arithmetic of BlockDiagMatrix (block_diag_ilu.hpp:98-102, 203-221), passed by value.
struct Dims {
    int nblocks, w, ndiag, nsat, ld;
    int64_t off_sub, off_sup, off_bot, off_top;  // offsets into one system's m_data
    int64_t ndata;
    __host__ __device__ int64_t block(int b) const { return (int64_t)b * w * ld; }
    __host__ __device__ int64_t diag_chunk(int di, int b) const {
        return ((int64_t)di * nblocks - (di * di + di) / 2 + b) * ld;
    }
    __host__ __device__ int64_t sat_chunk(int si, int b) const { return (int64_t)((si * si + si) / 2 + b) * ld; }
};

inline Dims make_dims(int nblocks, int w, int ndiag, int nsat, int ld)
{
    Dims d;
    d.nblocks = nblocks; d.w = w; d.ndiag = ndiag; d.nsat = nsat; d.ld = ld;
    const int64_t blk = (int64_t)nblocks * w * ld;
    const int64_t dh = ((int64_t)ndiag * nblocks - (ndiag * ndiag + ndiag) / 2) * ld;
    const int64_t sh = (int64_t)((nsat * nsat + nsat) / 2) * ld;
    d.off_sub = blk; d.off_sup = blk + dh; d.off_bot = blk + 2 * dh; d.off_top = blk + 2 * dh + sh;
    d.ndata = blk + 2 * (dh + sh);
    return d;
}

// Broadcast a double from lane `src` of the G-wide group the caller sits in.
template <int G>
__device__ __forceinline__ double group_bcast(double v, int src)
{
    int lo = __shfl_sync(FULL, __double2loint(v), src, G);
    int hi = __shfl_sync(FULL, __double2hiint(v), src, G);
    return __hiloint2double(hi, lo);
}

__device__ __forceinline__ double ld_stream(const double *p)
{
    // read-once data: keep it out of L1 so the chain's small working set stays there
    return __ldcs(p);
}

// 1/x for normal-range x: MUFU.RCP64H seed + the refinement CUDA's own IEEE reciprocal uses on its fast
// path (correctly rounded for |x| in [2^-990, 2^990]; callers test the exponent and take a true division
// otherwise).
__device__ __forceinline__ double rcp_fast(double x)
{
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    double e = fma(-x, y, 1.0);
    e = fma(e, e, e);
    y = fma(y, e, y);
    e = fma(-x, y, 1.0);
    return fma(y, e, y);
}
// true when the exponent of x is inside the range where rcp_fast is exact
__device__ __forceinline__ bool rcp_fast_ok(double x)
{
    const int ehi = (__double2hiint(x) >> 20) & 0x7ff;
    return ehi >= 0x020 && ehi <= 0x7dd;
}

}  // namespace bdilu
