// block_diag_ilu_b200.hpp -- C++ drop-in for the hot path of block_diag_ilu/include/block_diag_ilu.hpp.
//
// Same namespace, class names, constructor signatures, public fields and error behaviour as the reference's
// BlockDiagMatrix<Real_t> (hpp:104-424), ILU_inplace<Real_t> (:426-539) and ILU<Real_t> (:541-551), for
// Real_t = double and Real_t = float (the *_f32 entry points of bdilu.h).  Storage and accessors stay on the host exactly as in the reference (callers fill the
// matrix element by element); everything that computes -- the ILU_inplace constructor (factor), solve(),
// dot_vec(), scale_diag_add() -- is one call into libbdilu_b200.so (bdilu.h) and runs on the GPU.  There is no
// CPU fallback: without a CUDA device those calls throw std::runtime_error carrying bdilu_last_error().
//
// Not provided (off the hot path): as_banded_padded / BandedLU / DenseLU / SVD views, rms_diag,
// average_diag_weight, the long double instantiation.  The MatrixBase source constructor (hpp:124-159) is a template
// over any source with operator()(row, col).
//
// Link with -lbdilu_b200 (block_diag_ilu_b200/libbdilu_b200.so); see INTEGRATION.md.
#pragma once
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

#include "bdilu.h"

namespace block_diag_ilu {

namespace detail {
inline void check(int rc, const char *what)
{
    if (rc < 0) throw std::runtime_error(std::string(what) + ": " + bdilu_last_error());
}
// the C entry points behind each Real_t
template <typename Real_t> struct capi;
template <> struct capi<double> {
    typedef bdilu_ilu handle;
    static int default_ld(int w) { return bdilu_default_ld(w); }
    static int create(handle **o, double *d, int64_t ss, int ns, int N, int w, int nd, int nsat, int ld, int cb, int *ip,
                      int *rbc, int *cbr, int *info)
    {
        return bdilu_ilu_create(o, d, ss, ns, N, w, nd, nsat, ld, cb, ip, rbc, cbr, info);
    }
    static int solve(handle *h, const double *b, double *x, int nrhs, int *flags) { return bdilu_ilu_solve(h, b, x, nrhs, flags); }
    static void destroy(handle *h) { bdilu_ilu_destroy(h); }
    static int dot_vec(const double *d, int N, int w, int nd, int nsat, int ld, const double *v, double *o, int nrhs)
    {
        return bdilu_host_dot_vec(d, N, w, nd, nsat, ld, v, o, nrhs);
    }
    static int scale_diag_add(double *dst, const double *src, int N, int w, int nd, int nsat, int ld, double s, double a)
    {
        return bdilu_host_scale_diag_add(dst, src, N, w, nd, nsat, ld, s, a);
    }
};
template <> struct capi<float> {
    typedef bdilu_ilu_f32 handle;
    static int default_ld(int w) { return bdilu_default_ld_f32(w); }
    static int create(handle **o, float *d, int64_t ss, int ns, int N, int w, int nd, int nsat, int ld, int cb, int *ip,
                      int *rbc, int *cbr, int *info)
    {
        return bdilu_ilu_create_f32(o, d, ss, ns, N, w, nd, nsat, ld, cb, ip, rbc, cbr, info);
    }
    static int solve(handle *h, const float *b, float *x, int nrhs, int *flags) { return bdilu_ilu_solve_f32(h, b, x, nrhs, flags); }
    static void destroy(handle *h) { bdilu_ilu_destroy_f32(h); }
    static int dot_vec(const float *d, int N, int w, int nd, int nsat, int ld, const float *v, float *o, int nrhs)
    {
        return bdilu_host_dot_vec_f32(d, N, w, nd, nsat, ld, v, o, nrhs);
    }
    static int scale_diag_add(float *dst, const float *src, int N, int w, int nd, int nsat, int ld, float s, float a)
    {
        return bdilu_host_scale_diag_add_f32(dst, src, N, w, nd, nsat, ld, s, a);
    }
};
}  // namespace detail

// rowpiv2rowbycol / rowbycol2colbyrow (hpp:34-52): kept for callers that use them directly (host, O(w^2)).
inline void rowpiv2rowbycol(int n, const int *piv, int *rowbycol)
{
    for (int i = 0; i < n; ++i) rowbycol[i] = i;
    for (int i = 0; i < n; ++i) std::swap(rowbycol[i], rowbycol[piv[i] - 1]);
}
inline void rowbycol2colbyrow(int n, const int *rowbycol, int *colbyrow)
{
    for (int j = 0; j < n; ++j) colbyrow[rowbycol[j]] = j;
}

template <typename Real_t = double>
class BlockDiagMatrix {
    static_assert(std::is_same<Real_t, double>::value || std::is_same<Real_t, float>::value,
                  "block_diag_ilu_b200 implements Real_t = double and Real_t = float");
    typedef detail::capi<Real_t> api;
    bool m_own_data;

    static Real_t *alloc_(int ndata)
    {
        void *p = nullptr;
        if (posix_memalign(&p, 64, sizeof(Real_t) * (size_t)(ndata > 0 ? ndata : 1))) throw std::bad_alloc();
        std::memset(p, 0, sizeof(Real_t) * (size_t)(ndata > 0 ? ndata : 1));
        return static_cast<Real_t *>(p);
    }
    int diag_idx_(int di, int bi, int ci) const { return (di * m_nblocks - (di * di + di) / 2 + bi) * m_ld + ci; }
    int sat_idx_(int si, int bi, int ci) const { return ((si * si + si) / 2 + bi) * m_ld + ci; }

public:
    // public fields of MatrixBase (anyode_matrix.hpp:34-37) and BlockDiagMatrix (hpp:106-108)
    Real_t *m_data;
    int m_nr, m_nc, m_ld, m_ndata;
    const int m_nblocks, m_blockw, m_ndiag, m_nsat;
    const int m_blk_ndata, m_diag_hlf_ndata, m_sat_hlf_ndata;

    // hpp:110-123: data == nullptr -> owned, zero-initialised, 64-byte aligned, ld defaults to blockw rounded
    // up to 64 bytes; data != nullptr -> non-owning view and ld is mandatory.
    BlockDiagMatrix(Real_t *const data, const int nblocks, const int blockw, const int ndiag = 0, const int nsat = 0,
                    const int ld = 0)
        : m_own_data(data == nullptr), m_data(data), m_nr(nblocks * blockw), m_nc(nblocks * blockw),
          m_ld(ld ? ld : api::default_ld(blockw)),
          m_ndata((int)bdilu_ndata(nblocks, blockw, ndiag, nsat, ld ? ld : api::default_ld(blockw))),
          m_nblocks(nblocks), m_blockw(blockw), m_ndiag(ndiag), m_nsat(nsat),
          m_blk_ndata(nblocks * blockw * (ld ? ld : api::default_ld(blockw))),
          m_diag_hlf_ndata((ndiag * nblocks - (ndiag * ndiag + ndiag) / 2) * (ld ? ld : api::default_ld(blockw))),
          m_sat_hlf_ndata(((nsat * nsat + nsat) / 2) * (ld ? ld : api::default_ld(blockw)))
    {
        if (data != nullptr && ld == 0) throw std::runtime_error("give ld when providing data pointer.");
        if (m_own_data) m_data = alloc_(m_ndata);
    }
    // hpp:124-159: owned copy of the entries of `source` that fall inside the sparsity pattern.  The reference takes a
    // MatrixBase<Real_t>; any type with `operator()(int row, int col) const` works here (dense, banded, another
    // BlockDiagMatrix).
    template <typename Source, typename = decltype(std::declval<const Source &>()(0, 0))>
    BlockDiagMatrix(const Source &source, const int nblocks, const int blockw, const int ndiag = 0, const int nsat = 0,
                    const int ld = 0)
        : BlockDiagMatrix(static_cast<Real_t *>(nullptr), nblocks, blockw, ndiag, nsat, ld)
    {
        for (int bi = 0; bi < nblocks; ++bi)
            for (int ci = 0; ci < blockw; ++ci)
                for (int ri = 0; ri < blockw; ++ri) block(bi, ri, ci) = source(bi * blockw + ri, bi * blockw + ci);
        for (int di = 0; di < ndiag; ++di)
            for (int bi = 0; bi < nblocks - di - 1; ++bi)
                for (int ci = 0; ci < blockw; ++ci) {
                    sub(di, bi, ci) = source((bi + di + 1) * blockw + ci, bi * blockw + ci);
                    sup(di, bi, ci) = source(bi * blockw + ci, (bi + di + 1) * blockw + ci);
                }
        for (int si = 0; si < nsat; ++si)
            for (int bi = 0; bi <= si; ++bi)
                for (int ci = 0; ci < blockw; ++ci) {
                    bot(si, bi, ci) = source((nblocks - 1 - si + bi) * blockw + ci, bi * blockw + ci);
                    top(si, bi, ci) = source(bi * blockw + ci, (nblocks - 1 - si + bi) * blockw + ci);
                }
    }
    BlockDiagMatrix(const BlockDiagMatrix<Real_t> &ori)  // deep copy (anyode_matrix.hpp:45-47)
        : m_own_data(true), m_data(alloc_(ori.m_ndata)), m_nr(ori.m_nr), m_nc(ori.m_nc), m_ld(ori.m_ld),
          m_ndata(ori.m_ndata), m_nblocks(ori.m_nblocks), m_blockw(ori.m_blockw), m_ndiag(ori.m_ndiag),
          m_nsat(ori.m_nsat), m_blk_ndata(ori.m_blk_ndata), m_diag_hlf_ndata(ori.m_diag_hlf_ndata),
          m_sat_hlf_ndata(ori.m_sat_hlf_ndata)
    {
        std::copy(ori.m_data, ori.m_data + ori.m_ndata, m_data);
    }
    BlockDiagMatrix &operator=(const BlockDiagMatrix &) = delete;
    virtual ~BlockDiagMatrix()
    {
        if (m_own_data && m_data) std::free(m_data);
    }

    // element accessors (hpp:203-221)
    Real_t &block(int bi, int ri, int ci) const noexcept { return m_data[(bi * m_blockw + ci) * m_ld + ri]; }
    Real_t &sub(int di, int bi, int ci) const noexcept { return m_data[m_blk_ndata + diag_idx_(di, bi, ci)]; }
    Real_t &sup(int di, int bi, int ci) const noexcept
    {
        return m_data[m_blk_ndata + m_diag_hlf_ndata + diag_idx_(di, bi, ci)];
    }
    Real_t &bot(int si, int bi, int ci) const noexcept
    {
        return m_data[m_blk_ndata + 2 * m_diag_hlf_ndata + sat_idx_(si, bi, ci)];
    }
    Real_t &top(int si, int bi, int ci) const noexcept
    {
        return m_data[m_blk_ndata + 2 * m_diag_hlf_ndata + m_sat_hlf_ndata + sat_idx_(si, bi, ci)];
    }
    void set_block(int bi, int ri, int ci, Real_t v) const noexcept { block(bi, ri, ci) = v; }
    void set_sub(int di, int bi, int ci, Real_t v) const noexcept { sub(di, bi, ci) = v; }
    void set_sup(int di, int bi, int ci, Real_t v) const noexcept { sup(di, bi, ci) = v; }
    void set_bot(int si, int bi, int ci, Real_t v) const noexcept { bot(si, bi, ci) = v; }
    void set_top(int si, int bi, int ci, Real_t v) const noexcept { top(si, bi, ci) = v; }

    // structural-sparsity test and dense indexing (hpp:222-275): "Illegal index" outside the pattern
    bool valid_index(int ri, int ci) const
    {
        const int bri = ri / m_blockw, bci = ci / m_blockw;
        if (bri == bci) return true;
        if (ri - bri * m_blockw != ci - bci * m_blockw) return false;
        const int dist = bri > bci ? bri - bci : bci - bri;
        return !(dist > m_ndiag && m_nblocks - dist > m_nsat);
    }
    bool guaranteed_zero_index(int ri, int ci) const { return !valid_index(ri, ci); }
    Real_t &operator()(int ri, int ci)
    {
        const int bri = ri / m_blockw, bci = ci / m_blockw;
        const int lri = ri - bri * m_blockw, lci = ci - bci * m_blockw;
        if (bri == bci) return block(bri, lri, lci);
        if (lri != lci) throw std::runtime_error("Illegal index");
        const int dist = bri > bci ? bri - bci : bci - bri;
        const int lo = bri > bci ? bci : bri;  // block index of the upper-left end of the coupling
        if (dist > m_ndiag) {
            if (m_nblocks - dist > m_nsat) throw std::runtime_error("Illegal index");
            return bri > bci ? bot(m_nblocks - 1 - dist, lo, lci) : top(m_nblocks - 1 - dist, lo, lci);
        }
        return bri > bci ? sub(dist - 1, lo, lci) : sup(dist - 1, lo, lci);
    }
    const Real_t &operator()(int ri, int ci) const { return (*const_cast<BlockDiagMatrix *>(this))(ri, ci); }

    // hpp:276-290 -> bdilu_host_scale_diag_add (GPU)
    void scale_diag_add(const BlockDiagMatrix<Real_t> &source, Real_t scale = 1, Real_t diag_add = 0)
    {
        if (source.m_ndata != m_ndata) throw std::runtime_error("scale_diag_add: incompatible sizes");
        detail::check(api::scale_diag_add(m_data, source.m_data, m_nblocks, m_blockw, m_ndiag, m_nsat, m_ld, scale, diag_add),
                      "scale_diag_add");
    }
    void set_to_eye_plus_scaled_mtx(Real_t scale, const BlockDiagMatrix<Real_t> &source)
    {
        scale_diag_add(source, scale, 1);
    }
    // hpp:295-326 -> bdilu_host_dot_vec (GPU)
    void dot_vec(const Real_t *const vec, Real_t *const out)
    {
        detail::check(api::dot_vec(m_data, m_nblocks, m_blockw, m_ndiag, m_nsat, m_ld, vec, out, 1), "dot_vec");
    }
};

// ILU_inplace (hpp:426-539): the constructor factors the caller's view IN PLACE (on the GPU, results copied back
// into view->m_data so callers that peek at the factors keep working) and keeps only a raw pointer to it; the
// factors also stay device-resident inside the handle for the solves.
template <typename Real_t = double>
class ILU_inplace {
    typedef detail::capi<Real_t> api;
    typename api::handle *m_handle = nullptr;

public:
    BlockDiagMatrix<Real_t> *m_view;
    std::vector<int> m_ipiv, m_rowbycol, m_colbyrow;  // hpp:428-429 (buffer_t there; .data()/[] work alike)

    explicit ILU_inplace(BlockDiagMatrix<Real_t> *view)
        : m_view(view), m_ipiv((size_t)view->m_nblocks * view->m_blockw),
          m_rowbycol((size_t)view->m_nblocks * view->m_blockw), m_colbyrow((size_t)view->m_nblocks * view->m_blockw)
    {
        int info = 0;
        detail::check(api::create(&m_handle, view->m_data, view->m_ndata, 1, view->m_nblocks, view->m_blockw,
                                       view->m_ndiag, view->m_nsat, view->m_ld, /*copy_back=*/1, m_ipiv.data(),
                                       m_rowbycol.data(), m_colbyrow.data(), &info),
                      "ILU_inplace");
        if (info) {
            api::destroy(m_handle);
            m_handle = nullptr;
            throw std::runtime_error("ILU failed!");  // hpp:483-484
        }
    }
    ILU_inplace(const ILU_inplace &) = delete;
    ILU_inplace &operator=(const ILU_inplace &) = delete;
    ~ILU_inplace() { api::destroy(m_handle); }

    // hpp:486-538: 0 ok; 1..n first NaN in b (1-based); n+k zero U diagonal at row k.  x is always written.
    int solve(const Real_t *__restrict__ const b, Real_t *__restrict__ const x) const
    {
        const int rc = api::solve(m_handle, b, x, 1, nullptr);
        detail::check(rc, "ILU_inplace::solve");
        return rc;
    }
    // Beyond the reference: nrhs right-hand sides ([nrhs][n] contiguous) in one launch; flags[nrhs] nullable.
    int solve_many(const Real_t *b, Real_t *x, int nrhs, int *flags = nullptr) const
    {
        const int rc = api::solve(m_handle, b, x, nrhs, flags);
        detail::check(rc, "ILU_inplace::solve_many");
        return rc;
    }
};

// ILU (hpp:541-551): owns a deep copy of the source and factors the copy.
template <typename Real_t = double>
class ILU {
    BlockDiagMatrix<Real_t> m_mat;

public:
    ILU_inplace<Real_t> m_ilu_inplace;
    ILU(const BlockDiagMatrix<Real_t> &mat) : m_mat(mat), m_ilu_inplace(&m_mat) {}
    int solve(const Real_t *__restrict__ const b, Real_t *__restrict__ const x) { return m_ilu_inplace.solve(b, x); }
};

}  // namespace block_diag_ilu
