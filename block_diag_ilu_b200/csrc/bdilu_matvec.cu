// Kernel 3: BlockDiagMatrix<double>::dot_vec (block_diag_ilu.hpp:295-326), the residual b - A x and
// its squared 2-norm per (system, vector), plus scale_diag_add (:276-287).
//
// Purely streaming: every matrix element is read once per group of up to 8 vectors.  One thread per output row; consecutive
// threads read consecutive rows of a column-major block column, so each column load is one
// contiguous 8*blockw-byte segment.  The reference scatters sub/sup/top/bot contributions into
// `out`; here each row gathers the (at most 2*ndiag + satellite) couplings that land on it, so no
// zero-fill pass and no atomics are needed.
#include <mutex>

#include "../../include/bdilu.h"
#include "bdilu_common.cuh"
#include "bdilu_kernels.h"

namespace bdilu {

constexpr int MV_THREADS = 128;

// One output row against NR vectors at once: every matrix element is loaded once per NR vectors (the vectors of one
// system share the matrix, so a batch of right-hand sides re-reads it nrhs/NR times instead of nrhs times).
// `vs`: this block's slice of the NR vectors staged in shared memory as [column][NR] (NR > 1), else nullptr.
template <int NR>
__device__ __forceinline__ void row_dot(double (&acc)[NR], const double *__restrict__ sd, const Dims &D,
                                        const double *const (&vec)[NR], const double *vs, int bi, int row)
{
    const int w = D.w, N = D.nblocks;
    const double *col = sd + D.block(bi) + row;
#pragma unroll
    for (int r = 0; r < NR; ++r) acc[r] = 0.0;
    if (NR > 1) {
#pragma unroll 4
        for (int c = 0; c < w; ++c) {
            const double a = ld_stream(col + (int64_t)c * D.ld);
            const double2 *v2 = reinterpret_cast<const double2 *>(vs + c * NR);  // one broadcast LDS.128 per two vectors
#pragma unroll
            for (int r = 0; r < NR / 2; ++r) {
                const double2 v = v2[r];
                acc[2 * r] = fma(a, v.x, acc[2 * r]);
                acc[2 * r + 1] = fma(a, v.y, acc[2 * r + 1]);
            }
        }
    } else {
#pragma unroll 8
        for (int c = 0; c < w; ++c) acc[0] = fma(ld_stream(col + (int64_t)c * D.ld), __ldg(vec[0] + bi * w + c), acc[0]);
    }
    for (int di = 0; di < D.ndiag; ++di) {
        if (bi + di + 1 < N) {  // sup(di,bi,row) sits at (bi*w+row, (bi+di+1)*w+row)
            const double a = ld_stream(sd + D.off_sup + D.diag_chunk(di, bi) + row);
#pragma unroll
            for (int r = 0; r < NR; ++r) acc[r] = fma(a, __ldg(vec[r] + (bi + di + 1) * w + row), acc[r]);
        }
        if (bi - di - 1 >= 0) {  // sub(di,bi-di-1,row) sits at (bi*w+row, (bi-di-1)*w+row)
            const double a = ld_stream(sd + D.off_sub + D.diag_chunk(di, bi - di - 1) + row);
#pragma unroll
            for (int r = 0; r < NR; ++r) acc[r] = fma(a, __ldg(vec[r] + (bi - di - 1) * w + row), acc[r]);
        }
    }
    for (int sati = 0; sati < D.nsat; ++sati) {
        if (bi <= sati) {  // top(sati,bi,row) sits at (bi*w+row, (N-1-sati+bi)*w+row)
            const double a = ld_stream(sd + D.off_top + D.sat_chunk(sati, bi) + row);
#pragma unroll
            for (int r = 0; r < NR; ++r) acc[r] = fma(a, __ldg(vec[r] + (N - 1 - sati + bi) * w + row), acc[r]);
        }
        const int bj = bi - N + 1 + sati;  // bot(sati,bj,row) sits at ((N-1-sati+bj)*w+row, bj*w+row)
        if (bj >= 0) {
            const double a = ld_stream(sd + D.off_bot + D.sat_chunk(sati, bj) + row);
#pragma unroll
            for (int r = 0; r < NR; ++r) acc[r] = fma(a, __ldg(vec[r] + bj * w + row), acc[r]);
        }
    }
}

// MODE 0: out = A*vec.  MODE 1: out[(sys*nrhs + rhs)*gridDim.x + blockIdx.x] = this CTA's part of sum (b - A*x)^2
// (vec = x, rhsb = b); norm2_finish_kernel adds the parts in a fixed order, so norms are bit-reproducible.
// blockIdx.y = group of NR vectors.
template <int MODE, int NR>
__global__ void __launch_bounds__(MV_THREADS)
matvec_kernel(const double *__restrict__ data, int64_t sys_stride, Dims D, const double *__restrict__ vec,
              const double *__restrict__ rhsb, double *__restrict__ out, VecLayout V)
{
    const int sys = blockIdx.z, rhs0 = blockIdx.y * NR;
    const int n = D.nblocks * D.w;
    const int i = blockIdx.x * MV_THREADS + threadIdx.x;
    const double *sd = data + (int64_t)sys * sys_stride;
    int64_t voff[NR];
    const double *vp[NR];
#pragma unroll
    for (int r = 0; r < NR; ++r) {
        const int rhs = rhs0 + r < V.nrhs ? rhs0 + r : V.nrhs - 1;  // a ragged last group repeats its last vector
        voff[r] = (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride;
        vp[r] = vec + voff[r];
    }
    double val[NR];
#pragma unroll
    for (int r = 0; r < NR; ++r) val[r] = 0.0;
    // NR > 1: the slice of the vectors under this CTA's blocks goes to shared memory once, transposed to [column][NR]
    constexpr int MAXW = 64;
    __shared__ __align__(16) double vs[NR > 1 ? (MV_THREADS + 2 * MAXW) * NR : 1];
    const int i0 = blockIdx.x * MV_THREADS;
    const int b_lo = i0 / D.w;
    if (NR > 1) {
        const int last = min(n - 1, i0 + MV_THREADS - 1);
        const int len = (last / D.w + 1 - b_lo) * D.w;  // whole blocks: < MV_THREADS + 2 w
        for (int idx = threadIdx.x; idx < len * NR; idx += MV_THREADS) {
            const int r = idx / len, c = idx - r * len;
            const int rhs = rhs0 + r < V.nrhs ? rhs0 + r : V.nrhs - 1;
            vs[c * NR + r] = __ldg(vec + (int64_t)sys * V.sys_stride + (int64_t)rhs * V.rhs_stride + b_lo * D.w + c);
        }
        __syncthreads();
    }
    if (i < n) {
        const int bi = i / D.w, row = i - bi * D.w;
        row_dot<NR>(val, sd, D, vp, vs + (bi - b_lo) * D.w * NR, bi, row);
#pragma unroll
        for (int r = 0; r < NR; ++r) {
            if (MODE == 0) {
                if (rhs0 + r < V.nrhs) out[voff[r] + i] = val[r];
            } else {
                val[r] = __ldg(rhsb + voff[r] + i) - val[r];
                val[r] *= val[r];
            }
        }
    }
    if (MODE == 1) {
        __shared__ double part[NR][MV_THREADS / 32];
#pragma unroll
        for (int r = 0; r < NR; ++r) {
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) val[r] += group_bcast<32>(val[r], (threadIdx.x & 31) ^ off);
            if ((threadIdx.x & 31) == 0) part[r][threadIdx.x >> 5] = val[r];
        }
        __syncthreads();
        if (threadIdx.x < NR && rhs0 + threadIdx.x < V.nrhs) {
            double s = 0.0;
            for (int k = 0; k < MV_THREADS / 32; ++k) s += part[threadIdx.x][k];
            out[((int64_t)sys * V.nrhs + rhs0 + threadIdx.x) * gridDim.x + blockIdx.x] = s;
        }
    }
}

// norm2[v] = sum of the `parts` partial sums of vector v, always in the same order: lane l adds parts l, l+32, ...
// serially, then a fixed shuffle tree.  One warp per vector.
__global__ void __launch_bounds__(128)
norm2_finish_kernel(const double *__restrict__ partial, int parts, int64_t nvec, double *__restrict__ norm2)
{
    const int64_t v = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
    if (v >= nvec) return;
    const int lane = threadIdx.x & 31;
    double s = 0.0;
    for (int k = lane; k < parts; k += 32) s += partial[v * parts + k];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) s += group_bcast<32>(s, lane ^ off);
    if (lane == 0) norm2[v] = s;
}

__global__ void __launch_bounds__(256)
scale_diag_add_kernel(double *dst, const double *src, int64_t dst_stride, int64_t src_stride, Dims D,
                      double scale, double diag_add)
{
    const int sys = blockIdx.y;
    const int64_t base = (int64_t)sys * dst_stride, sbase = (int64_t)sys * src_stride;
    const int64_t blk_end = D.off_sub;
    const int64_t per_block = (int64_t)D.w * D.ld;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < D.ndata;
         i += (int64_t)gridDim.x * blockDim.x) {
        // product and sum rounded separately, like the reference's two statements (no FMA contraction)
        double v = __dmul_rn(scale, src[sbase + i]);
        if (i < blk_end) {
            const int64_t in_blk = i % per_block;
            const int c = (int)(in_blk / D.ld), r = (int)(in_blk % D.ld);
            if (r == c) v = __dadd_rn(v, diag_add);
        }
        dst[base + i] = v;
    }
}

// grid.z / grid.y carry the system index: batches beyond 65535 systems go out in slabs
constexpr int MAX_GRID_YZ = 65535;

cudaError_t launch_dot_vec(const double *data, int64_t sys_stride, int nsys, const Dims &D,
                           const double *vec, double *out, const VecLayout &V, cudaStream_t st)
{
    const int n = D.nblocks * D.w;
    if (n == 0 || nsys == 0 || V.nrhs == 0) return cudaSuccess;
    const int nr = D.w > 64 ? 1 : (V.nrhs >= 8 ? 8 : (V.nrhs >= 4 ? 4 : 1));  // the staged slice is sized for blockw <= 64
    if ((V.nrhs + nr - 1) / nr > MAX_GRID_YZ) return cudaErrorInvalidValue;  // > 65535 groups of vectors per system
    for (int s0 = 0; s0 < nsys; s0 += MAX_GRID_YZ) {
        const int ns = nsys - s0 < MAX_GRID_YZ ? nsys - s0 : MAX_GRID_YZ;
        const double *d0 = data + (int64_t)s0 * sys_stride, *v0 = vec + (int64_t)s0 * V.sys_stride;
        double *o0 = out + (int64_t)s0 * V.sys_stride;
        dim3 grid((n + MV_THREADS - 1) / MV_THREADS, (V.nrhs + nr - 1) / nr, ns);
        if (nr == 8) matvec_kernel<0, 8><<<grid, MV_THREADS, 0, st>>>(d0, sys_stride, D, v0, nullptr, o0, V);
        else if (nr == 4) matvec_kernel<0, 4><<<grid, MV_THREADS, 0, st>>>(d0, sys_stride, D, v0, nullptr, o0, V);
        else matvec_kernel<0, 1><<<grid, MV_THREADS, 0, st>>>(d0, sys_stride, D, v0, nullptr, o0, V);
        count_launch(1);
    }
    return cudaGetLastError();
}

cudaError_t launch_residual_norm2(const double *data, int64_t sys_stride, int nsys, const Dims &D,
                                  const double *b, const double *x, const VecLayout &V, double *norm2,
                                  cudaStream_t st)
{
    const int n = D.nblocks * D.w;
    if (nsys == 0 || V.nrhs == 0) return cudaSuccess;
    if (n == 0) return cudaMemsetAsync(norm2, 0, sizeof(double) * (size_t)nsys * V.nrhs, st);
    const int nr = D.w > 64 ? 1 : (V.nrhs >= 8 ? 8 : (V.nrhs >= 4 ? 4 : 1));  // the staged slice is sized for blockw <= 64
    if ((V.nrhs + nr - 1) / nr > MAX_GRID_YZ) return cudaErrorInvalidValue;
    const int gx = (n + MV_THREADS - 1) / MV_THREADS;
    // one partial sum per (vector, CTA) in stream-ordered scratch, added up in a fixed order afterwards: no atomics, so
    // the norms are bit-reproducible
    const int slab = nsys < MAX_GRID_YZ ? nsys : MAX_GRID_YZ;
    double *partial = nullptr;
    cudaError_t e = cudaMallocAsync(reinterpret_cast<void **>(&partial), sizeof(double) * (size_t)slab * V.nrhs * gx, st);
    if (e != cudaSuccess) return e;
    for (int s0 = 0; s0 < nsys; s0 += MAX_GRID_YZ) {
        const int ns = nsys - s0 < MAX_GRID_YZ ? nsys - s0 : MAX_GRID_YZ;
        const double *d0 = data + (int64_t)s0 * sys_stride;
        const double *b0 = b + (int64_t)s0 * V.sys_stride, *x0 = x + (int64_t)s0 * V.sys_stride;
        dim3 grid(gx, (V.nrhs + nr - 1) / nr, ns);
        if (nr == 8) matvec_kernel<1, 8><<<grid, MV_THREADS, 0, st>>>(d0, sys_stride, D, x0, b0, partial, V);
        else if (nr == 4) matvec_kernel<1, 4><<<grid, MV_THREADS, 0, st>>>(d0, sys_stride, D, x0, b0, partial, V);
        else matvec_kernel<1, 1><<<grid, MV_THREADS, 0, st>>>(d0, sys_stride, D, x0, b0, partial, V);
        const int64_t nvec = (int64_t)ns * V.nrhs;
        norm2_finish_kernel<<<(unsigned)((nvec + 3) / 4), 128, 0, st>>>(partial, gx, nvec, norm2 + (int64_t)s0 * V.nrhs);
        count_launch(2);
    }
    e = cudaGetLastError();
    cudaFreeAsync(partial, st);
    return e;
}

// ---- residual norms without any allocation (bdilu_residual_norm2_static, include/bdilu.h) -------------------------------
// The per-CTA partial sums go to a file-scope device array instead of stream-ordered scratch, for callers that may not
// allocate (CUDA graph capture, serving loops).  Every device that loads the library holds the array until exit.  All
// calls share it: calls that overlap on one GPU (two streams, two threads) race.  The kernels overwrite every slot they
// later read, so nothing has to start at zero and the previous call's contents are harmless.
constexpr int64_t STATIC_PARTIALS = BDILU_STATIC_NORM2_PARTIALS;  // doubles: 1 MiB per device
static __device__ __align__(256) double g_norm2_partials[STATIC_PARTIALS];

// The array's name is only a host-side stand-in in host code; its device address is looked up once per device, with
// that device current.
static constexpr int MAX_DEVICES = 64;
static double *g_partials_addr[MAX_DEVICES];
static std::mutex g_partials_mutex;

static cudaError_t static_partials(double **out)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_partials_mutex);
    if (!g_partials_addr[dev]) {
        double *p = nullptr;
        e = cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_norm2_partials);
        if (e != cudaSuccess) return e;
        g_partials_addr[dev] = p;
    }
    *out = g_partials_addr[dev];
    return cudaSuccess;
}

int64_t residual_norm2_static_partials(int nsys, const Dims &D, const VecLayout &V)
{
    const int64_t n = (int64_t)D.nblocks * D.w;
    return (int64_t)nsys * V.nrhs * ((n + MV_THREADS - 1) / MV_THREADS);
}

cudaError_t launch_residual_norm2_static(const double *data, int64_t sys_stride, int nsys, const Dims &D,
                                         const double *b, const double *x, const VecLayout &V, double *norm2,
                                         cudaStream_t st)
{
    const int n = D.nblocks * D.w;
    if (nsys == 0 || V.nrhs == 0) return cudaSuccess;
    if (n == 0) return cudaMemsetAsync(norm2, 0, sizeof(double) * (size_t)nsys * V.nrhs, st);
    const int nr = D.w > 64 ? 1 : (V.nrhs >= 8 ? 8 : (V.nrhs >= 4 ? 4 : 1));
    // one launch covers the batch (the caller checked the partial-sum count, which bounds nsys far below 65535)
    if (nsys > MAX_GRID_YZ || (V.nrhs + nr - 1) / nr > MAX_GRID_YZ) return cudaErrorInvalidValue;
    if (residual_norm2_static_partials(nsys, D, V) > STATIC_PARTIALS) return cudaErrorInvalidValue;
    double *partial = nullptr;
    cudaError_t e = static_partials(&partial);
    if (e != cudaSuccess) return e;
    const int gx = (n + MV_THREADS - 1) / MV_THREADS;
    dim3 grid(gx, (V.nrhs + nr - 1) / nr, nsys);
    if (nr == 8) matvec_kernel<1, 8><<<grid, MV_THREADS, 0, st>>>(data, sys_stride, D, x, b, partial, V);
    else if (nr == 4) matvec_kernel<1, 4><<<grid, MV_THREADS, 0, st>>>(data, sys_stride, D, x, b, partial, V);
    else matvec_kernel<1, 1><<<grid, MV_THREADS, 0, st>>>(data, sys_stride, D, x, b, partial, V);
    const int64_t nvec = (int64_t)nsys * V.nrhs;
    norm2_finish_kernel<<<(unsigned)((nvec + 3) / 4), 128, 0, st>>>(partial, gx, nvec, norm2);
    count_launch(2);
    return cudaGetLastError();
}

cudaError_t launch_scale_diag_add(double *dst, const double *src, int64_t dst_stride, int64_t src_stride, int nsys,
                                  const Dims &D, double scale, double diag_add, cudaStream_t st)
{
    if (nsys == 0 || D.ndata == 0) return cudaSuccess;
    const int64_t want = (D.ndata + 255) / 256;
    for (int s0 = 0; s0 < nsys; s0 += MAX_GRID_YZ) {
        const int ns = nsys - s0 < MAX_GRID_YZ ? nsys - s0 : MAX_GRID_YZ;
        dim3 grid((unsigned)(want < sm_count() * 8 ? want : sm_count() * 8), ns);
        scale_diag_add_kernel<<<grid, 256, 0, st>>>(dst + (int64_t)s0 * dst_stride, src + (int64_t)s0 * src_stride, dst_stride,
                                                    src_stride, D, scale, diag_add);
        count_launch(1);
    }
    return cudaGetLastError();
}

}  // namespace bdilu
