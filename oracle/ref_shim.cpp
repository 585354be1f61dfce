// ref_shim.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// A C-ABI window onto the UNMODIFIED reference header.  It is compiled by oracle/Makefile with
//   -I/root/reference/block_diag_ilu/include -I/root/reference/external/anyode/include
// i.e. the reference sources are used where they lie; nothing of them is copied into this repo.
// The resulting oracle/_ref/libbdilu_ref.so is git-ignored, travels to the GPU box with the
// snapshot, and is used only by tests/, __graft_entry__.smoke() and bench.py (cpu_baseline /
// --impl reference) as checker and CPU baseline.
//
// Everything numerical below is the reference's own code: BlockDiagMatrix (view ctor,
// block_diag_ilu.hpp:110-123), ILU_inplace (ctor = factor, :430-485; solve, :486-538),
// BlockDiagMatrix::dot_vec (:295-326), scale_diag_add (:276-287), and LAPACK dgetrf through
// AnyODE::getrf_callback (anyode_blas_lapack.hpp:55-57).
#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "block_diag_ilu.hpp"

#define EXPORT extern "C" __attribute__((visibility("default")))

using block_diag_ilu::BlockDiagMatrix;
using block_diag_ilu::ILU_inplace;

namespace {
struct RefILU {
    std::unique_ptr<BlockDiagMatrix<double>> view;
    std::unique_ptr<ILU_inplace<double>> ilu;
};
double now() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
}  // namespace

EXPORT int ref_built_with_openmp() {
#if defined(_OPENMP) && defined(BLOCK_DIAG_ILU_WITH_OPENMP)
    return 1;
#else
    return 0;
#endif
}

EXPORT int ref_max_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

// thread count of the header's own OpenMP loop over blocks (block_diag_ilu.hpp:442-448)
EXPORT void ref_set_factor_threads(int n) {
    char buf[32];
    snprintf(buf, sizeof buf, "%d", n < 1 ? 1 : n);
    setenv("BLOCK_DIAG_ILU_NUM_THREADS", buf, 1);
}

EXPORT int64_t ref_ndata(int nblocks, int blockw, int ndiag, int nsat, int ld) {
    // data=nullptr lets the header pick its default ld when ld==0 (hpp:98)
    BlockDiagMatrix<double> m(nullptr, nblocks, blockw, ndiag, nsat, ld);
    return m.m_ndata;
}

EXPORT int ref_default_ld(int blockw) {
    BlockDiagMatrix<double> m(nullptr, 1, blockw, 0, 0, 0);
    return m.m_ld;
}

// Factor IN PLACE in the caller's buffer (view ctor + ILU_inplace ctor).  Returns a handle, or
// nullptr when the reference throws ("ILU failed!"); *threw is set accordingly.
EXPORT void *ref_ilu_create(double *data, int nblocks, int blockw, int ndiag, int nsat, int ld,
                            int *ipiv, int *rowbycol, int *colbyrow, int *threw) {
    auto *h = new RefILU;
    *threw = 0;
    try {
        h->view.reset(new BlockDiagMatrix<double>(data, nblocks, blockw, ndiag, nsat, ld));
        h->ilu.reset(new ILU_inplace<double>(h->view.get()));
    } catch (const std::exception &) {
        *threw = 1;
        delete h;
        return nullptr;
    }
    const int n = nblocks * blockw;
    for (int i = 0; i < n; ++i) {
        if (ipiv) ipiv[i] = h->ilu->m_ipiv[i];
        if (rowbycol) rowbycol[i] = h->ilu->m_rowbycol[i];
        if (colbyrow) colbyrow[i] = h->ilu->m_colbyrow[i];
    }
    return h;
}

EXPORT int ref_ilu_solve(void *handle, const double *b, double *x) {
    return static_cast<RefILU *>(handle)->ilu->solve(b, x);
}

EXPORT void ref_ilu_destroy(void *handle) { delete static_cast<RefILU *>(handle); }

EXPORT void ref_dot_vec(double *data, int nblocks, int blockw, int ndiag, int nsat, int ld,
                        const double *vec, double *out) {
    BlockDiagMatrix<double> m(data, nblocks, blockw, ndiag, nsat, ld);
    m.dot_vec(vec, out);
}

EXPORT void ref_scale_diag_add(double *dst, double *src, int nblocks, int blockw, int ndiag, int nsat,
                               int ld, double scale, double diag_add) {
    BlockDiagMatrix<double> d(dst, nblocks, blockw, ndiag, nsat, ld);
    BlockDiagMatrix<double> s(src, nblocks, blockw, ndiag, nsat, ld);
    d.scale_diag_add(s, scale, diag_add);
}

EXPORT void ref_rowpiv2rowbycol(int n, const int *piv, int *rowbycol) {
    block_diag_ilu::rowpiv2rowbycol(n, piv, rowbycol);
}
EXPORT void ref_rowbycol2colbyrow(int n, const int *rowbycol, int *colbyrow) {
    block_diag_ilu::rowbycol2colbyrow(n, rowbycol, colbyrow);
}

// Batched factor + nrhs solves over nsys systems; data is factored in place.
//   mode 0 ("as shipped"): systems one after another; the header's OpenMP loop over blocks
//           uses `nthreads`; solve is serial (it always is in the reference).
//   mode 1 ("system-parallel harness"): omp-parallel over systems, single-threaded header inside.
// Returns wall seconds.  infos[s]: 0 ok, -1 reference threw, else first nonzero solve flag.
EXPORT double ref_factor_solve_batch(double *data, int64_t sys_stride, int nsys, int nblocks, int blockw,
                                     int ndiag, int nsat, int ld, int *ipiv, int *rowbycol, int *colbyrow,
                                     const double *b, double *x, int nrhs, int *infos, int mode, int nthreads) {
    const int64_t n = (int64_t)nblocks * blockw;
    if (nthreads < 1) nthreads = ref_max_threads();
    auto one = [&](int s) {
        int threw = 0;
        void *h = ref_ilu_create(data + s * sys_stride, nblocks, blockw, ndiag, nsat, ld,
                                 ipiv ? ipiv + s * n : nullptr, rowbycol ? rowbycol + s * n : nullptr,
                                 colbyrow ? colbyrow + s * n : nullptr, &threw);
        int info = threw ? -1 : 0;
        if (h) {
            for (int r = 0; r < nrhs; ++r) {
                int flag = ref_ilu_solve(h, b + (s * (int64_t)nrhs + r) * n, x + (s * (int64_t)nrhs + r) * n);
                if (flag && !info) info = flag;
            }
            ref_ilu_destroy(h);
        }
        if (infos) infos[s] = info;
    };
    double t0;
    if (mode == 0) {
        ref_set_factor_threads(nthreads);
        t0 = now();
        for (int s = 0; s < nsys; ++s) one(s);
    } else {
        ref_set_factor_threads(1);
        t0 = now();
#ifdef _OPENMP
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 1)
#endif
        for (int s = 0; s < nsys; ++s) one(s);
    }
    return now() - t0;
}

// ---- Real_t = float: the same header instantiated with float (sgetrf through AnyODE::getrf_callback<float>,
// anyode_blas_lapack.hpp:56) ----------------------------------------------------------------------------------------
namespace {
struct RefILUf {
    std::unique_ptr<BlockDiagMatrix<float>> view;
    std::unique_ptr<ILU_inplace<float>> ilu;
};
}  // namespace

EXPORT int64_t ref_ndata_f32(int nblocks, int blockw, int ndiag, int nsat, int ld) {
    BlockDiagMatrix<float> m(nullptr, nblocks, blockw, ndiag, nsat, ld);
    return m.m_ndata;
}

EXPORT int ref_default_ld_f32(int blockw) {
    BlockDiagMatrix<float> m(nullptr, 1, blockw, 0, 0, 0);
    return m.m_ld;
}

EXPORT void *ref_ilu_create_f32(float *data, int nblocks, int blockw, int ndiag, int nsat, int ld,
                                int *ipiv, int *rowbycol, int *colbyrow, int *threw) {
    auto *h = new RefILUf;
    *threw = 0;
    try {
        h->view.reset(new BlockDiagMatrix<float>(data, nblocks, blockw, ndiag, nsat, ld));
        h->ilu.reset(new ILU_inplace<float>(h->view.get()));
    } catch (const std::exception &) {
        *threw = 1;
        delete h;
        return nullptr;
    }
    const int n = nblocks * blockw;
    for (int i = 0; i < n; ++i) {
        if (ipiv) ipiv[i] = h->ilu->m_ipiv[i];
        if (rowbycol) rowbycol[i] = h->ilu->m_rowbycol[i];
        if (colbyrow) colbyrow[i] = h->ilu->m_colbyrow[i];
    }
    return h;
}

EXPORT int ref_ilu_solve_f32(void *handle, const float *b, float *x) {
    return static_cast<RefILUf *>(handle)->ilu->solve(b, x);
}

EXPORT void ref_ilu_destroy_f32(void *handle) { delete static_cast<RefILUf *>(handle); }

EXPORT void ref_dot_vec_f32(float *data, int nblocks, int blockw, int ndiag, int nsat, int ld,
                            const float *vec, float *out) {
    BlockDiagMatrix<float> m(data, nblocks, blockw, ndiag, nsat, ld);
    m.dot_vec(vec, out);
}

EXPORT void ref_scale_diag_add_f32(float *dst, float *src, int nblocks, int blockw, int ndiag, int nsat,
                                   int ld, float scale, float diag_add) {
    BlockDiagMatrix<float> d(dst, nblocks, blockw, ndiag, nsat, ld);
    BlockDiagMatrix<float> s(src, nblocks, blockw, ndiag, nsat, ld);
    d.scale_diag_add(s, scale, diag_add);
}
