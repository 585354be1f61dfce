// Host-pointer entry points of the C ABI: own the device mirrors, marshal H2D/D2H, block until done.
// These are what the C++ header shim (include/block_diag_ilu_b200.hpp) and the Python wrapper bind
// when the caller's data lives in host memory, as it always does in the reference.
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/bdilu.h"
#include "bdilu_kernels.h"

using namespace bdilu;

struct bdilu_ilu {
    int nsys, nblocks, blockw, ndiag, nsat, ld;
    int64_t sys_stride;  // device stride == host stride
    double *d_data = nullptr;
    int *d_ipiv = nullptr, *d_rowbycol = nullptr, *d_colbyrow = nullptr, *d_info = nullptr;
    double *d_b = nullptr, *d_x = nullptr;  // grown on demand by solve
    int *d_flags = nullptr;
    size_t vec_cap = 0, flag_cap = 0;
    cudaStream_t st = nullptr;
    int device = 0;  // the device the mirrors live on: every later call runs there, whatever the caller's current device
};

namespace {

// device entry points report through bdilu_last_error(); host ones reuse that channel
int set_err(int code, const std::string &msg);

#define CUH(call)                                                                        \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess) {                                                         \
            cudaGetLastError();                                                          \
            return set_err(BDILU_ERR_CUDA, std::string(#call ": ") + cudaGetErrorString(e_)); \
        }                                                                                \
    } while (0)

thread_local std::string t_host_err;
int set_err(int code, const std::string &msg)
{
    t_host_err = msg;
    bdilu::set_last_error(t_host_err.c_str());
    return code;
}

// current device := dev for the lifetime of the guard (a handle / workspace may be used from a thread whose current
// device is another one)
struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev)
    {
        if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; ok = false; return; }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~DeviceGuard()
    {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

void free_handle(bdilu_ilu *h)
{
    if (!h) return;
    DeviceGuard g(h->device);
    cudaFree(h->d_data); cudaFree(h->d_ipiv); cudaFree(h->d_rowbycol); cudaFree(h->d_colbyrow);
    cudaFree(h->d_info); cudaFree(h->d_b); cudaFree(h->d_x); cudaFree(h->d_flags);
    if (h->st) cudaStreamDestroy(h->st);
    delete h;
}

}  // namespace

extern "C" {

int bdilu_ilu_create(bdilu_ilu **out, double *host_data, int64_t sys_stride, int nsys, int nblocks, int blockw,
                     int ndiag, int nsat, int ld, int copy_back, int *ipiv, int *rowbycol, int *colbyrow, int *info)
{
    if (!out) return set_err(BDILU_ERR_ARG, "out is null");
    *out = nullptr;
    if (nsys < 1 || !host_data) return set_err(BDILU_ERR_ARG, "nsys < 1 or null data");
    if (ld == 0) ld = bdilu_default_ld(blockw);
    const int64_t nd = bdilu_ndata(nblocks, blockw, ndiag, nsat, ld);
    if (nsys == 1) sys_stride = nd;
    if (bdilu_device_count() < 1) return set_err(BDILU_ERR_CUDA, "no CUDA device: libbdilu_b200 has no CPU fallback");
    bdilu_ilu *h = new (std::nothrow) bdilu_ilu;
    if (!h) return set_err(BDILU_ERR_ARG, "out of host memory");
    h->nsys = nsys; h->nblocks = nblocks; h->blockw = blockw; h->ndiag = ndiag; h->nsat = nsat; h->ld = ld;
    h->sys_stride = sys_stride;
    if (cudaGetDevice(&h->device) != cudaSuccess) {
        delete h;
        cudaGetLastError();
        return set_err(BDILU_ERR_CUDA, "cudaGetDevice failed");
    }
    const size_t nbytes = sizeof(double) * (size_t)((int64_t)(nsys - 1) * sys_stride + nd);
    const size_t nidx = (size_t)nsys * nblocks * blockw;
#define CUC(call)                                                                                      \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess) {                                                                       \
            if (h->st) cudaStreamSynchronize(h->st); /* nothing may still write the caller's memory */ \
            free_handle(h);                                                                            \
            cudaGetLastError();                                                                        \
            return set_err(BDILU_ERR_CUDA, std::string(#call ": ") + cudaGetErrorString(e_));          \
        }                                                                                              \
    } while (0)
    CUC(cudaStreamCreateWithFlags(&h->st, cudaStreamNonBlocking));
    CUC(cudaMalloc(&h->d_data, nbytes ? nbytes : 8));
    CUC(cudaMalloc(&h->d_ipiv, sizeof(int) * (nidx ? nidx : 1)));
    CUC(cudaMalloc(&h->d_rowbycol, sizeof(int) * (nidx ? nidx : 1)));
    CUC(cudaMalloc(&h->d_colbyrow, sizeof(int) * (nidx ? nidx : 1)));
    CUC(cudaMalloc(&h->d_info, sizeof(int) * (size_t)nsys));
    CUC(cudaMemcpyAsync(h->d_data, host_data, nbytes, cudaMemcpyHostToDevice, h->st));
    int rc = bdilu_factor(h->d_data, sys_stride, nsys, nblocks, blockw, ndiag, nsat, ld, h->d_ipiv, h->d_rowbycol,
                          h->d_colbyrow, h->d_info, h->st);
    if (rc) { cudaStreamSynchronize(h->st); free_handle(h); return rc; }
    if (copy_back) CUC(cudaMemcpyAsync(host_data, h->d_data, nbytes, cudaMemcpyDeviceToHost, h->st));
    if (ipiv) CUC(cudaMemcpyAsync(ipiv, h->d_ipiv, sizeof(int) * nidx, cudaMemcpyDeviceToHost, h->st));
    if (rowbycol) CUC(cudaMemcpyAsync(rowbycol, h->d_rowbycol, sizeof(int) * nidx, cudaMemcpyDeviceToHost, h->st));
    if (colbyrow) CUC(cudaMemcpyAsync(colbyrow, h->d_colbyrow, sizeof(int) * nidx, cudaMemcpyDeviceToHost, h->st));
    if (info) CUC(cudaMemcpyAsync(info, h->d_info, sizeof(int) * (size_t)nsys, cudaMemcpyDeviceToHost, h->st));
    CUC(cudaStreamSynchronize(h->st));
#undef CUC
    *out = h;
    return BDILU_OK;
}

int bdilu_ilu_solve(bdilu_ilu *h, const double *host_b, double *host_x, int nrhs, int *flags)
{
    if (!h || !host_b || !host_x || nrhs < 1) return set_err(BDILU_ERR_ARG, "null handle/pointer or nrhs < 1");
    DeviceGuard guard(h->device);
    if (!guard.ok) return set_err(BDILU_ERR_CUDA, "cannot select the handle's device");
    const int64_t n = (int64_t)h->nblocks * h->blockw;
    const size_t nvec = (size_t)h->nsys * nrhs;
    const size_t vbytes = sizeof(double) * nvec * (size_t)n;
    if (vbytes > h->vec_cap) {
        cudaFree(h->d_b); cudaFree(h->d_x);
        h->d_b = h->d_x = nullptr; h->vec_cap = 0;
        CUH(cudaMalloc(&h->d_b, vbytes ? vbytes : 8));
        CUH(cudaMalloc(&h->d_x, vbytes ? vbytes : 8));
        h->vec_cap = vbytes;
    }
    if (nvec > h->flag_cap) {
        cudaFree(h->d_flags); h->d_flags = nullptr; h->flag_cap = 0;
        CUH(cudaMalloc(&h->d_flags, sizeof(int) * nvec));
        h->flag_cap = nvec;
    }
    std::vector<int> own;  // flags when the caller does not want them (freed on every path)
    int *hf = flags;
    if (!hf) {
        try {
            own.resize(nvec);
        } catch (const std::bad_alloc &) {
            return set_err(BDILU_ERR_ARG, "out of host memory");
        }
        hf = own.data();
    }
    // from here on copies into the caller's memory may be in flight: every exit goes through the synchronisation
    int rc = BDILU_OK;
    cudaError_t e = cudaMemcpyAsync(h->d_b, host_b, vbytes, cudaMemcpyHostToDevice, h->st);
    if (e == cudaSuccess) {
        rc = bdilu_solve(h->d_data, h->sys_stride, h->nsys, h->nblocks, h->blockw, h->ndiag, h->nsat, h->ld, h->d_rowbycol,
                         h->d_colbyrow, h->d_b, h->d_x, nrhs, n, n * nrhs, nullptr, h->d_flags, h->st);
        if (rc == BDILU_OK) e = cudaMemcpyAsync(host_x, h->d_x, vbytes, cudaMemcpyDeviceToHost, h->st);
        if (rc == BDILU_OK && e == cudaSuccess)
            e = cudaMemcpyAsync(hf, h->d_flags, sizeof(int) * nvec, cudaMemcpyDeviceToHost, h->st);
    }
    const cudaError_t es = cudaStreamSynchronize(h->st);
    if (rc != BDILU_OK) return rc;
    if (e != cudaSuccess || es != cudaSuccess) {
        cudaGetLastError();
        return set_err(BDILU_ERR_CUDA, std::string("bdilu_ilu_solve: ") + cudaGetErrorString(e != cudaSuccess ? e : es));
    }
    int first = 0;
    for (size_t i = 0; i < nvec && !first; ++i) first = hf[i];
    return first;
}

void bdilu_ilu_destroy(bdilu_ilu *h) { free_handle(h); }

int bdilu_host_dot_vec(const double *host_data, int nblocks, int blockw, int ndiag, int nsat, int ld,
                       const double *host_vec, double *host_out, int nrhs)
{
    if (!host_data || !host_vec || !host_out || nrhs < 1) return set_err(BDILU_ERR_ARG, "null pointer or nrhs < 1");
    if (bdilu_device_count() < 1) return set_err(BDILU_ERR_CUDA, "no CUDA device: libbdilu_b200 has no CPU fallback");
    if (ld == 0) ld = bdilu_default_ld(blockw);
    const int64_t nd = bdilu_ndata(nblocks, blockw, ndiag, nsat, ld), n = (int64_t)nblocks * blockw;
    double *d_data = nullptr, *d_v = nullptr, *d_o = nullptr;
    int rc = BDILU_OK;
    cudaError_t e;
    if ((e = cudaMalloc(&d_data, sizeof(double) * (nd ? nd : 1))) != cudaSuccess ||
        (e = cudaMalloc(&d_v, sizeof(double) * (n * nrhs ? n * nrhs : 1))) != cudaSuccess ||
        (e = cudaMalloc(&d_o, sizeof(double) * (n * nrhs ? n * nrhs : 1))) != cudaSuccess ||
        (e = cudaMemcpy(d_data, host_data, sizeof(double) * nd, cudaMemcpyHostToDevice)) != cudaSuccess ||
        (e = cudaMemcpy(d_v, host_vec, sizeof(double) * n * nrhs, cudaMemcpyHostToDevice)) != cudaSuccess) {
        rc = set_err(BDILU_ERR_CUDA, cudaGetErrorString(e));
    } else {
        rc = bdilu_dot_vec(d_data, nd, 1, nblocks, blockw, ndiag, nsat, ld, d_v, d_o, nrhs, n, n * nrhs, nullptr);
        if (!rc && (e = cudaMemcpy(host_out, d_o, sizeof(double) * n * nrhs, cudaMemcpyDeviceToHost)) != cudaSuccess)
            rc = set_err(BDILU_ERR_CUDA, cudaGetErrorString(e));
    }
    cudaFree(d_data); cudaFree(d_v); cudaFree(d_o);
    return rc;
}

int bdilu_host_scale_diag_add(double *host_dst, const double *host_src, int nblocks, int blockw, int ndiag, int nsat,
                              int ld, double scale, double diag_add)
{
    if (!host_dst || !host_src) return set_err(BDILU_ERR_ARG, "null pointer");
    if (bdilu_device_count() < 1) return set_err(BDILU_ERR_CUDA, "no CUDA device: libbdilu_b200 has no CPU fallback");
    if (ld == 0) ld = bdilu_default_ld(blockw);
    const int64_t nd = bdilu_ndata(nblocks, blockw, ndiag, nsat, ld);
    double *d = nullptr;
    int rc = BDILU_OK;
    cudaError_t e;
    if ((e = cudaMalloc(&d, sizeof(double) * (nd ? nd : 1))) != cudaSuccess ||
        (e = cudaMemcpy(d, host_src, sizeof(double) * nd, cudaMemcpyHostToDevice)) != cudaSuccess) {
        rc = set_err(BDILU_ERR_CUDA, cudaGetErrorString(e));
    } else {
        rc = bdilu_scale_diag_add(d, d, nd, 1, nblocks, blockw, ndiag, nsat, ld, scale, diag_add, nullptr);
        if (!rc && (e = cudaMemcpy(host_dst, d, sizeof(double) * nd, cudaMemcpyDeviceToHost)) != cudaSuccess)
            rc = set_err(BDILU_ERR_CUDA, cudaGetErrorString(e));
    }
    cudaFree(d);
    return rc;
}

}  // extern "C"

// ---- one-shot ILU(A) + solve(b, x) on host memory, pipelined over slabs of systems ---------------
namespace {
struct Workspace {
    std::mutex mu;
    double *d_data = nullptr, *d_b = nullptr, *d_x = nullptr;
    int *d_idx = nullptr, *d_info = nullptr, *d_flags = nullptr;
    int *h_codes = nullptr;  // pinned staging for info + flags (see bdilu_host_factor_solve)
    size_t cap_data = 0, cap_vec = 0, cap_x = 0, cap_idx = 0, cap_info = 0, cap_flags = 0, cap_codes = 0;
    cudaStream_t st[3] = {nullptr, nullptr, nullptr};
    int device = -1;
    void release()
    {
        if (device < 0) return;
        DeviceGuard g(device);
        cudaFree(d_data); cudaFree(d_b); cudaFree(d_x); cudaFree(d_idx); cudaFree(d_info); cudaFree(d_flags);
        if (h_codes) cudaFreeHost(h_codes);
        for (auto &s : st)
            if (s) cudaStreamDestroy(s);
        d_data = d_b = d_x = nullptr; d_idx = d_info = d_flags = nullptr; h_codes = nullptr;
        cap_data = cap_vec = cap_x = cap_idx = cap_info = cap_flags = cap_codes = 0;
        st[0] = st[1] = st[2] = nullptr;
        cudaGetLastError();
    }
};
// one workspace per device (the mirrors of one device are useless on another), released when the library is unloaded
constexpr int MAX_DEVICES = 64;
struct Workspaces {
    Workspace ws[MAX_DEVICES];
    ~Workspaces()
    {
        for (auto &w : ws) w.release();
    }
};
Workspaces g_wss;

template <class T>
cudaError_t grow(T *&p, size_t &cap, size_t want)
{
    if (want <= cap) return cudaSuccess;
    cudaFree(p);
    p = nullptr; cap = 0;
    cudaError_t e = cudaMalloc(&p, want * sizeof(T));
    if (e == cudaSuccess) cap = want;
    return e;
}
}  // namespace

extern "C" int bdilu_host_factor_solve(const double *host_data, int64_t sys_stride, int nsys, int nblocks,
                                       int blockw, int ndiag, int nsat, int ld, const double *host_b,
                                       double *host_x, int nrhs, int *info, int *flags)
{
    if (!host_data || !host_b || !host_x || nsys < 1 || nrhs < 1)
        return set_err(BDILU_ERR_ARG, "null pointer, nsys < 1 or nrhs < 1");
    if (bdilu_device_count() < 1) return set_err(BDILU_ERR_CUDA, "no CUDA device: libbdilu_b200 has no CPU fallback");
    if (ld == 0) ld = bdilu_default_ld(blockw);
    const int64_t nd = bdilu_ndata(nblocks, blockw, ndiag, nsat, ld), n = (int64_t)nblocks * blockw;
    if (nsys == 1) sys_stride = nd;
    int device = 0;
    CUH(cudaGetDevice(&device));
    if (device < 0 || device >= MAX_DEVICES) return set_err(BDILU_ERR_UNSUPPORTED, "device index out of range");
    Workspace &W = g_wss.ws[device];
    std::lock_guard<std::mutex> lock(W.mu);
    W.device = device;
    // Copies into the caller's memory run asynchronously on three streams: whatever happens below, nothing returns
    // before those streams are drained (the workspace is reused by the next call, host_x belongs to the caller).
    struct Drain {
        Workspace &W;
        ~Drain()
        {
            for (auto &s : W.st)
                if (s) cudaStreamSynchronize(s);
        }
    } drain{W};
    for (auto &s : W.st)
        if (!s) CUH(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    // device copy is packed tight (stride nd) whatever the host stride
    CUH(grow(W.d_data, W.cap_data, (size_t)nsys * nd));
    CUH(grow(W.d_b, W.cap_vec, (size_t)nsys * nrhs * n));
    CUH(grow(W.d_x, W.cap_x, (size_t)nsys * nrhs * n));
    CUH(grow(W.d_idx, W.cap_idx, (size_t)3 * nsys * n));
    CUH(grow(W.d_info, W.cap_info, (size_t)nsys));
    CUH(grow(W.d_flags, W.cap_flags, (size_t)nsys * nrhs));
    // info/flags land in a pinned staging buffer and are handed over after the last slab: an asynchronous copy into
    // the caller's (usually pageable) int arrays would block the host at every slab and serialise the pipeline
    // (measured: 124 ms instead of ~93 ms per cfg3 batch).
    const size_t ncodes = (size_t)nsys + (size_t)nsys * nrhs;
    if (ncodes > W.cap_codes) {
        if (W.h_codes) cudaFreeHost(W.h_codes);
        W.h_codes = nullptr; W.cap_codes = 0;
        CUH(cudaHostAlloc((void **)&W.h_codes, ncodes * sizeof(int), cudaHostAllocDefault));
        W.cap_codes = ncodes;
    }
    int *h_info = W.h_codes, *h_flags = W.h_codes + nsys;
    int *d_ipiv = W.d_idx, *d_rbc = W.d_idx + (size_t)nsys * n, *d_cbr = W.d_idx + (size_t)2 * nsys * n;

    // slabs of systems round-robin over 3 streams: H2D(slab k+1) overlaps factor+solve(slab k) and D2H(slab k-1)
    int slab = (nsys + 11) / 12;
    if (slab < 1) slab = 1;
    int k = 0;
    for (int s0 = 0; s0 < nsys; s0 += slab, ++k) {
        const int ns = (s0 + slab <= nsys) ? slab : nsys - s0;
        cudaStream_t st = W.st[k % 3];
        if (sys_stride == nd) {
            CUH(cudaMemcpyAsync(W.d_data + (size_t)s0 * nd, host_data + (size_t)s0 * nd, sizeof(double) * (size_t)ns * nd,
                                cudaMemcpyHostToDevice, st));
        } else {
            CUH(cudaMemcpy2DAsync(W.d_data + (size_t)s0 * nd, sizeof(double) * nd, host_data + (size_t)s0 * sys_stride,
                                  sizeof(double) * sys_stride, sizeof(double) * nd, ns, cudaMemcpyHostToDevice, st));
        }
        CUH(cudaMemcpyAsync(W.d_b + (size_t)s0 * nrhs * n, host_b + (size_t)s0 * nrhs * n,
                            sizeof(double) * (size_t)ns * nrhs * n, cudaMemcpyHostToDevice, st));
        int rc = bdilu_factor_solve(W.d_data + (size_t)s0 * nd, nd, ns, nblocks, blockw, ndiag, nsat, ld,
                                    d_ipiv + (size_t)s0 * n, d_rbc + (size_t)s0 * n, d_cbr + (size_t)s0 * n,
                                    W.d_info + s0, W.d_b + (size_t)s0 * nrhs * n, W.d_x + (size_t)s0 * nrhs * n, nrhs, n,
                                    n * nrhs, W.d_flags + (size_t)s0 * nrhs, st);
        if (rc) return rc;
        CUH(cudaMemcpyAsync(host_x + (size_t)s0 * nrhs * n, W.d_x + (size_t)s0 * nrhs * n,
                            sizeof(double) * (size_t)ns * nrhs * n, cudaMemcpyDeviceToHost, st));
        if (info) CUH(cudaMemcpyAsync(h_info + s0, W.d_info + s0, sizeof(int) * ns, cudaMemcpyDeviceToHost, st));
        if (flags)
            CUH(cudaMemcpyAsync(h_flags + (size_t)s0 * nrhs, W.d_flags + (size_t)s0 * nrhs, sizeof(int) * (size_t)ns * nrhs,
                                cudaMemcpyDeviceToHost, st));
    }
    for (auto &s : W.st) CUH(cudaStreamSynchronize(s));
    if (info) memcpy(info, h_info, sizeof(int) * (size_t)nsys);
    if (flags) memcpy(flags, h_flags, sizeof(int) * (size_t)nsys * nrhs);
    return BDILU_OK;
}
