// C ABI of libbdilu_b200.so (see include/bdilu.h for the contract of every entry point).
#include <atomic>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <string>

#include "../../include/bdilu.h"
#include "../../include/bdilu_internal.h"
#include "bdilu_kernels.h"

namespace bdilu {
static std::atomic<long long> g_launches{0};
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
int sm_count()
{
    static std::atomic<int> cache[64];  // 0 = not asked yet
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
        cudaGetLastError();
        return 148;
    }
    int n = cache[dev].load(std::memory_order_relaxed);
    if (n == 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n < 1) {
            cudaGetLastError();
            n = 148;
        }
        cache[dev].store(n, std::memory_order_relaxed);
    }
    return n;
}
}  // namespace bdilu

using namespace bdilu;

namespace {

thread_local std::string t_err;

}  // namespace
namespace bdilu {
void set_last_error(const char *msg) { t_err = msg; }
}  // namespace bdilu
namespace {

int fail(int code, const char *msg)
{
    t_err = msg;
    return code;
}

int cuda_fail(cudaError_t e, const char *where)
{
    t_err = std::string(where) + ": " + cudaGetErrorString(e);
    cudaGetLastError();  // clear the sticky-less error state
    return BDILU_ERR_CUDA;
}

#define CU(call)                                               \
    do {                                                       \
        cudaError_t e_ = (call);                               \
        if (e_ != cudaSuccess) return cuda_fail(e_, #call);    \
    } while (0)

int check_dims(int nsys, int64_t sys_stride, int nblocks, int blockw, int ndiag, int nsat, int &ld, Dims &D,
               int esize = (int)sizeof(double))
{
    if (nsys < 0 || nblocks < 0 || blockw < 1 || ndiag < 0 || nsat < 0 || ld < 0)
        return fail(BDILU_ERR_ARG, "negative dimension or blockw < 1");
    if (ld == 0) ld = esize == 4 ? bdilu_default_ld_f32(blockw) : bdilu_default_ld(blockw);
    if (ld < blockw) return fail(BDILU_ERR_ARG, "ld < blockw");
    if (nblocks > 0 && (ndiag > nblocks - 1 && ndiag > 0)) return fail(BDILU_ERR_ARG, "ndiag > nblocks - 1");
    if (nsat > nblocks) return fail(BDILU_ERR_ARG, "nsat > nblocks");
    if (blockw > BDILU_MAX_BLOCKW) return fail(BDILU_ERR_UNSUPPORTED, "blockw > 64 is not supported");
    if (nblocks >= (1 << 24)) return fail(BDILU_ERR_UNSUPPORTED, "nblocks >= 2^24 is not supported");
    D = make_dims(nblocks, blockw, ndiag, nsat, ld);
    if (D.ndata > 0x7fffffffLL) return fail(BDILU_ERR_ARG, "one system exceeds the reference's int m_ndata");
    if (nsys > 1 && sys_stride < D.ndata) return fail(BDILU_ERR_ARG, "sys_stride < ndata");
    return BDILU_OK;
}

int check_vecs(const Dims &D, int nsys, int nrhs, int64_t rhs_stride, int64_t sys_rhs_stride)
{
    const int64_t n = (int64_t)D.nblocks * D.w;
    if (nrhs < 0) return fail(BDILU_ERR_ARG, "nrhs < 0");
    if (nrhs > 1 && rhs_stride < n) return fail(BDILU_ERR_ARG, "rhs_stride < nblocks*blockw");
    if (nsys > 1 && sys_rhs_stride < n) return fail(BDILU_ERR_ARG, "sys_rhs_stride < nblocks*blockw");
    return BDILU_OK;
}

}  // namespace

// the same argument checks and error channel for the other translation units with entry points (bdilu_f32.cu)
namespace bdilu {
int api_check_dims(int nsys, int64_t sys_stride, int nblocks, int blockw, int ndiag, int nsat, int &ld, Dims &D, int esize)
{
    return check_dims(nsys, sys_stride, nblocks, blockw, ndiag, nsat, ld, D, esize);
}
int api_check_vecs(const Dims &D, int nsys, int nrhs, int64_t rhs_stride, int64_t sys_rhs_stride)
{
    return check_vecs(D, nsys, nrhs, rhs_stride, sys_rhs_stride);
}
int api_fail(int code, const char *msg) { return fail(code, msg); }
int api_cuda_fail(cudaError_t e, const char *where) { return cuda_fail(e, where); }
}  // namespace bdilu

extern "C" {

int bdilu_version(void) { return BDILU_VERSION; }
const char *bdilu_last_error(void) { return t_err.c_str(); }

int bdilu_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int bdilu_default_ld(int blockw)
{
    return (int)((((size_t)blockw * sizeof(double) + 63) & ~(size_t)63) / sizeof(double));
}

int bdilu_default_ld_f32(int blockw)
{
    return (int)((((size_t)blockw * sizeof(float) + 63) & ~(size_t)63) / sizeof(float));
}

int64_t bdilu_ndata(int nblocks, int blockw, int ndiag, int nsat, int ld)
{
    if (ld == 0) ld = bdilu_default_ld(blockw);
    return make_dims(nblocks, blockw, ndiag, nsat, ld).ndata;
}

int bdilu_debug_force_chain_solve(int on)
{
    const int old = g_solve_path;
    g_solve_path = on;
    return old;
}

int bdilu_debug_skew_stats(int64_t *out3, int reset)
{
    unsigned long long v[3] = {0, 0, 0};
    cudaError_t e = skew_stats(v, reset != 0);
    if (e != cudaSuccess) return fail(BDILU_ERR_CUDA, cudaGetErrorString(e));
    if (out3)
        for (int i = 0; i < 3; ++i) out3[i] = (int64_t)v[i];
    return BDILU_OK;
}

int64_t bdilu_launch_count(void) { return g_launches.load(); }
void bdilu_launch_count_reset(void) { g_launches.store(0); }

size_t bdilu_workspace_bytes(int, int, int, int, int, int, int) { return 0; }

int bdilu_factor(double *data, int64_t sys_stride, int nsys, int nblocks, int blockw, int ndiag, int nsat,
                 int ld, int *ipiv, int *rowbycol, int *colbyrow, int *info, void *stream)
{
    Dims D;
    int rc = check_dims(nsys, sys_stride, nblocks, blockw, ndiag, nsat, ld, D);
    if (rc) return rc;
    if (nsys == 0) return BDILU_OK;
    if (!data || !ipiv || !rowbycol || !colbyrow || !info) return fail(BDILU_ERR_ARG, "null pointer");
    CU(launch_factor(data, sys_stride, nsys, D, ipiv, rowbycol, colbyrow, info, (cudaStream_t)stream));
    return BDILU_OK;
}

int bdilu_solve(const double *data, int64_t sys_stride, int nsys, int nblocks, int blockw, int ndiag, int nsat,
                int ld, const int *rowbycol, const int *colbyrow, const double *b, double *x, int nrhs,
                int64_t rhs_stride, int64_t sys_rhs_stride, double *, int *flags, void *stream)
{
    Dims D;
    int rc = check_dims(nsys, sys_stride, nblocks, blockw, ndiag, nsat, ld, D);
    if (rc) return rc;
    if ((rc = check_vecs(D, nsys, nrhs, rhs_stride, sys_rhs_stride))) return rc;
    if (nsys == 0 || nrhs == 0) return BDILU_OK;
    if (!data || !rowbycol || !colbyrow || !b || !x) return fail(BDILU_ERR_ARG, "null pointer");
    if (b == x) return fail(BDILU_ERR_ARG, "b and x must not alias");
    VecLayout V{nrhs, rhs_stride, sys_rhs_stride};
    if (nblocks == 0 && flags) CU(cudaMemsetAsync(flags, 0, sizeof(int) * (size_t)nsys * nrhs, (cudaStream_t)stream));
    CU(launch_solve(data, sys_stride, nsys, D, rowbycol, colbyrow, b, x, V, flags, (cudaStream_t)stream));
    return BDILU_OK;
}

int bdilu_factor_solve(double *data, int64_t sys_stride, int nsys, int nblocks, int blockw, int ndiag, int nsat,
                       int ld, int *ipiv, int *rowbycol, int *colbyrow, int *info, const double *b, double *x,
                       int nrhs, int64_t rhs_stride, int64_t sys_rhs_stride, int *flags, void *stream)
{
    int rc = bdilu_factor(data, sys_stride, nsys, nblocks, blockw, ndiag, nsat, ld, ipiv, rowbycol, colbyrow,
                          info, stream);
    if (rc) return rc;
    return bdilu_solve(data, sys_stride, nsys, nblocks, blockw, ndiag, nsat, ld, rowbycol, colbyrow, b, x, nrhs,
                       rhs_stride, sys_rhs_stride, nullptr, flags, stream);
}

int bdilu_dot_vec(const double *data, int64_t sys_stride, int nsys, int nblocks, int blockw, int ndiag, int nsat,
                  int ld, const double *vec, double *out, int nrhs, int64_t rhs_stride, int64_t sys_rhs_stride,
                  void *stream)
{
    Dims D;
    int rc = check_dims(nsys, sys_stride, nblocks, blockw, ndiag, nsat, ld, D);
    if (rc) return rc;
    if ((rc = check_vecs(D, nsys, nrhs, rhs_stride, sys_rhs_stride))) return rc;
    if (nsys == 0 || nrhs == 0) return BDILU_OK;
    if (!data || !vec || !out) return fail(BDILU_ERR_ARG, "null pointer");
    VecLayout V{nrhs, rhs_stride, sys_rhs_stride};
    CU(launch_dot_vec(data, sys_stride, nsys, D, vec, out, V, (cudaStream_t)stream));
    return BDILU_OK;
}

int bdilu_residual_norm2(const double *data, int64_t sys_stride, int nsys, int nblocks, int blockw, int ndiag,
                         int nsat, int ld, const double *b, const double *x, int nrhs, int64_t rhs_stride,
                         int64_t sys_rhs_stride, double *norm2, void *stream)
{
    Dims D;
    int rc = check_dims(nsys, sys_stride, nblocks, blockw, ndiag, nsat, ld, D);
    if (rc) return rc;
    if ((rc = check_vecs(D, nsys, nrhs, rhs_stride, sys_rhs_stride))) return rc;
    if (nsys == 0 || nrhs == 0) return BDILU_OK;
    if (!data || !b || !x || !norm2) return fail(BDILU_ERR_ARG, "null pointer");
    VecLayout V{nrhs, rhs_stride, sys_rhs_stride};
    CU(launch_residual_norm2(data, sys_stride, nsys, D, b, x, V, norm2, (cudaStream_t)stream));
    return BDILU_OK;
}

int bdilu_residual_norm2_static(const double *data, int64_t sys_stride, int nsys, int nblocks, int blockw, int ndiag,
                                int nsat, int ld, const double *b, const double *x, int nrhs, int64_t rhs_stride,
                                int64_t sys_rhs_stride, double *norm2, void *stream)
{
    Dims D;
    int rc = check_dims(nsys, sys_stride, nblocks, blockw, ndiag, nsat, ld, D);
    if (rc) return rc;
    if ((rc = check_vecs(D, nsys, nrhs, rhs_stride, sys_rhs_stride))) return rc;
    if (nsys == 0 || nrhs == 0) return BDILU_OK;
    if (!data || !b || !x || !norm2) return fail(BDILU_ERR_ARG, "null pointer");
    VecLayout V{nrhs, rhs_stride, sys_rhs_stride};
    if (residual_norm2_static_partials(nsys, D, V) > BDILU_STATIC_NORM2_PARTIALS) {
        char msg[200];
        snprintf(msg, sizeof msg,
                 "bdilu_residual_norm2_static: nsys*nrhs*ceil(nblocks*blockw/128) = %lld exceeds the static scratch of %d "
                 "partial sums; use bdilu_residual_norm2",
                 (long long)residual_norm2_static_partials(nsys, D, V), BDILU_STATIC_NORM2_PARTIALS);
        return fail(BDILU_ERR_UNSUPPORTED, msg);
    }
    CU(launch_residual_norm2_static(data, sys_stride, nsys, D, b, x, V, norm2, (cudaStream_t)stream));
    return BDILU_OK;
}

int bdilu_scale_diag_add(double *dst, const double *src, int64_t sys_stride, int nsys, int nblocks, int blockw,
                         int ndiag, int nsat, int ld, double scale, double diag_add, void *stream)
{
    Dims D;
    int rc = check_dims(nsys, sys_stride, nblocks, blockw, ndiag, nsat, ld, D);
    if (rc) return rc;
    if (nsys == 0) return BDILU_OK;
    if (!dst || !src) return fail(BDILU_ERR_ARG, "null pointer");
    CU(launch_scale_diag_add(dst, src, sys_stride, sys_stride, nsys, D, scale, diag_add, (cudaStream_t)stream));
    return BDILU_OK;
}

int bdilu_factor_scaled(double *dst, int64_t dst_stride, const double *src, int64_t src_stride, int nsys, int nblocks,
                        int blockw, int ndiag, int nsat, int ld, double scale, double diag_add, int *ipiv,
                        int *rowbycol, int *colbyrow, int *info, void *stream)
{
    Dims D;
    int rc = check_dims(nsys, dst_stride, nblocks, blockw, ndiag, nsat, ld, D);
    if (rc) return rc;
    if (nsys > 1 && src_stride < D.ndata) return fail(BDILU_ERR_ARG, "src_stride < ndata");
    if (nsys == 0) return BDILU_OK;
    if (!dst || !src || !ipiv || !rowbycol || !colbyrow || !info) return fail(BDILU_ERR_ARG, "null pointer");
    CU(launch_factor_scaled(dst, dst_stride, src, src_stride, nsys, D, scale, diag_add, ipiv, rowbycol, colbyrow, info,
                            (cudaStream_t)stream));
    return BDILU_OK;
}

}  // extern "C"
