/*
 * bdilu_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C CPU restatement of the block_diag_ilu hot path.  It exists only so that
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs can
 * CHECK the CUDA path.  Nothing under block_diag_ilu_b200/ or include/ may link or call it.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this file against
 *   (1) every known-answer vector the reference's own tests hold for the path
 *       (tests/golden/reference_kats.py, transcribed with file:line citations), and
 *   (2) outputs of the untouched reference header compiled in place (oracle/_ref/, see
 *       oracle/build_oracle.py) on seeded random inputs, committed as tests/golden/ref_vectors.npz.
 *
 * Each function cites the reference lines it restates (paths relative to /root/reference).
 * The per-block LU in the reference is LAPACK dgetrf (third-party, not vendored, not version
 * pinned: block_diag_ilu/_config.py:2, setup.py:102).  Its published algorithm -- partial
 * pivoting, pivot = first entry of maximum modulus (idamax), unit-lower L, 1-based ipiv,
 * info = index of the first exactly-zero pivot, the zero column is left unscaled -- is restated
 * here in the unblocked right-looking form the reference itself ships as its LAPACK-free
 * fallback (external/anyode/include/anyode/anyode_blasless.hpp:11-60).
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define EXPORT __attribute__((visibility("default")))

/* ---- layout: block_diag_ilu/include/block_diag_ilu.hpp:98-102, 203-221 ------------------ */
typedef struct {
    int nblocks, blockw, ndiag, nsat, ld;
    int64_t blk_ndata, diag_hlf_ndata, sat_hlf_ndata;
} dims_t;

static dims_t mkdims(int nblocks, int blockw, int ndiag, int nsat, int ld)
{
    dims_t d;
    d.nblocks = nblocks; d.blockw = blockw; d.ndiag = ndiag; d.nsat = nsat; d.ld = ld;
    d.blk_ndata = (int64_t)nblocks * blockw * ld;                                  /* :99  */
    d.diag_hlf_ndata = ((int64_t)ndiag * nblocks - (ndiag * ndiag + ndiag) / 2) * ld; /* :100 */
    d.sat_hlf_ndata = (int64_t)((nsat * nsat + nsat) / 2) * ld;                    /* :101 */
    return d;
}

EXPORT int64_t oracle_ndata(int nblocks, int blockw, int ndiag, int nsat, int ld)
{
    dims_t d = mkdims(nblocks, blockw, ndiag, nsat, ld);
    return d.blk_ndata + 2 * (d.diag_hlf_ndata + d.sat_hlf_ndata);                 /* :102 */
}

/* default ld: blockw rounded up to 64 bytes (hpp:98, anyode_matrix.hpp:18-21) */
EXPORT int oracle_default_ld(int blockw)
{
    return (int)((((size_t)blockw * sizeof(double) + 63) & ~(size_t)63) / sizeof(double));
}

#define BLOCK(D, b, r, c) data[((int64_t)(b) * (D).blockw + (c)) * (D).ld + (r)]                /* :203-205 */
#define DIAGIDX(D, di, b, c) (((int64_t)(di) * (D).nblocks - ((di) * (di) + (di)) / 2 + (b)) * (D).ld + (c))
#define SUB(D, di, b, c) data[DIAGIDX(D, di, b, c) + (D).blk_ndata]                             /* :206-209 */
#define SUP(D, di, b, c) data[DIAGIDX(D, di, b, c) + (D).blk_ndata + (D).diag_hlf_ndata]        /* :210-212 */
#define SATIDX(D, si, b, c) ((int64_t)(((si) * (si) + (si)) / 2 + (b)) * (D).ld + (c))
#define BOT(D, si, b, c) data[SATIDX(D, si, b, c) + (D).blk_ndata + 2 * (D).diag_hlf_ndata]     /* :214-217 */
#define TOP(D, si, b, c) data[SATIDX(D, si, b, c) + (D).blk_ndata + 2 * (D).diag_hlf_ndata + (D).sat_hlf_ndata] /* :218-220 */

/* ---- LU of one block: LAPACK dgetrf semantics (call site hpp:460-467) ------------------- */
EXPORT int oracle_getrf(int n, double *a, int lda, int *ipiv)
{
    int info = 0;
    for (int i = 0; i < n; ++i) {
        int pivrow = i;
        double absmax = fabs(a[(size_t)i * lda + i]);
        for (int j = i + 1; j < n; ++j) {            /* first strict maximum == idamax */
            double cur = fabs(a[(size_t)i * lda + j]);
            if (cur > absmax) { absmax = cur; pivrow = j; }
        }
        ipiv[i] = pivrow + 1;                        /* Fortran indexing */
        if (absmax == 0.0) {                         /* dgetf2: record, skip the column */
            if (info == 0) info = i + 1;
            continue;
        }
        if (pivrow != i)
            for (int c = 0; c < n; ++c) {
                double t = a[(size_t)c * lda + i];
                a[(size_t)c * lda + i] = a[(size_t)c * lda + pivrow];
                a[(size_t)c * lda + pivrow] = t;
            }
        const double piv = a[(size_t)i * lda + i];
        for (int r = i + 1; r < n; ++r)
            a[(size_t)i * lda + r] /= piv;           /* true division (tests check exact quotients) */
        for (int c = i + 1; c < n; ++c) {
            const double u = a[(size_t)c * lda + i];
            for (int r = i + 1; r < n; ++r)
                a[(size_t)c * lda + r] -= a[(size_t)i * lda + r] * u;
        }
    }
    return info;
}

/* hpp:34-41 */
EXPORT void oracle_rowpiv2rowbycol(int n, const int *piv, int *rowbycol)
{
    for (int i = 0; i < n; ++i) rowbycol[i] = i;
    for (int i = 0; i < n; ++i) {
        int j = piv[i] - 1, t = rowbycol[i];
        rowbycol[i] = rowbycol[j];
        rowbycol[j] = t;
    }
}

/* hpp:43-52 (inverse permutation) */
EXPORT void oracle_rowbycol2colbyrow(int n, const int *rowbycol, int *colbyrow)
{
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j)
            if (rowbycol[j] == i) { colbyrow[i] = j; break; }
}

/* ILU_inplace ctor, hpp:430-485.  Returns the first nonzero per-block getrf info in block
 * order (the reference throws "ILU failed!" when this is nonzero, hpp:483-484). */
EXPORT int oracle_factor(double *data, int nblocks, int blockw, int ndiag, int nsat, int ld,
                         int *ipiv, int *rowbycol, int *colbyrow)
{
    dims_t D = mkdims(nblocks, blockw, ndiag, nsat, ld);
    int info_ = 0;
    for (int bi = 0; bi < nblocks; ++bi) {
        int info = oracle_getrf(blockw, &BLOCK(D, bi, 0, 0), ld, ipiv + (size_t)bi * blockw);
        if (info != 0 && info_ == 0) info_ = info;                                /* :469-470 */
        for (int ci = 0; ci < blockw; ++ci)
            for (int di = 0; di < ndiag && bi + di < nblocks - 1; ++di)
                SUB(D, di, bi, ci) /= BLOCK(D, bi, ci, ci);                        /* :471-475 */
        if (bi < nsat)
            for (int sati = bi; sati < nsat; ++sati)
                for (int ci = 0; ci < blockw; ++ci)
                    BOT(D, sati, bi, ci) /= BLOCK(D, bi, ci, ci);                  /* :476-479 */
        oracle_rowpiv2rowbycol(blockw, ipiv + (size_t)bi * blockw, rowbycol + (size_t)bi * blockw);
        oracle_rowbycol2colbyrow(blockw, rowbycol + (size_t)bi * blockw, colbyrow + (size_t)bi * blockw);
    }
    return info_;
}

/* ILU_inplace::solve, hpp:486-538.  y is caller scratch of nblocks*blockw doubles. */
EXPORT int oracle_solve(const double *data, int nblocks, int blockw, int ndiag, int nsat, int ld,
                        const int *rowbycol, const int *colbyrow,
                        const double *b, double *x, double *y)
{
    dims_t D = mkdims(nblocks, blockw, ndiag, nsat, ld);
    const int nblk = nblocks, blkw = blockw, n = nblocks * blockw;
    int info = 0;
    for (int i = 0; i < n; ++i)
        if (isnan(b[i])) { info = i + 1; break; }                                 /* :500-501 */
    for (int bri = 0; bri < nblk; ++bri) {                                        /* :502-517 */
        for (int li = 0; li < blkw; ++li) {
            double s = 0.0;
            for (int lci = 0; lci < li; ++lci)
                s += BLOCK(D, bri, li, lci) * y[bri * blkw + lci];
            const int ci = colbyrow[bri * blkw + li];
            const int dmax = ndiag < bri ? ndiag : bri;
            for (int di = 1; di <= dmax; ++di)
                s += SUB(D, di - 1, bri - di, ci) * y[(bri - di) * blkw + ci];
            for (int bci = 0; bci <= nsat + bri - nblk; ++bci)
                s += BOT(D, nblk + bci - bri - 1, bci, ci) * y[bci * blkw + ci];
            y[bri * blkw + li] = b[bri * blkw + rowbycol[bri * blkw + li]] - s;
        }
    }
    for (int bri = nblk - 1; bri >= 0; --bri) {                                   /* :518-536 */
        for (int li = blkw; li > 0; --li) {
            double s = 0.0;
            for (int ci = li; ci < blkw; ++ci)
                s += BLOCK(D, bri, li - 1, ci) * x[bri * blkw + ci];
            const int ci = colbyrow[bri * blkw + li - 1];
            const int dmax = (nblk - bri - 1) < ndiag ? (nblk - bri - 1) : ndiag;
            for (int di = 1; di <= dmax; ++di)
                s += SUP(D, di - 1, bri, ci) * x[(bri + di) * blkw + ci];
            for (int sati = nsat; sati > bri; --sati)
                s += TOP(D, sati - 1, bri, ci) * x[(nblk - sati + bri) * blkw + ci];
            x[bri * blkw + li - 1] = (y[bri * blkw + li - 1] - s) / BLOCK(D, bri, li - 1, li - 1);
            if (BLOCK(D, bri, li - 1, li - 1) == 0 && info == 0)
                info = nblk * blkw + bri * blkw + (li - 1);
        }
    }
    return info;
}

/* BlockDiagMatrix::dot_vec, hpp:295-326 */
EXPORT void oracle_dot_vec(const double *data, int nblocks, int blockw, int ndiag, int nsat, int ld,
                           const double *vec, double *out)
{
    dims_t D = mkdims(nblocks, blockw, ndiag, nsat, ld);
    const int nblk = nblocks, blkw = blockw;
    for (int i = 0; i < nblk * blkw; ++i) out[i] = 0.0;
    for (int bri = 0; bri < nblk; ++bri)
        for (int lci = 0; lci < blkw; ++lci)
            for (int lri = 0; lri < blkw; ++lri)
                out[bri * blkw + lri] += vec[bri * blkw + lci] * BLOCK(D, bri, lri, lci);
    for (int di = 0; di < ndiag; ++di)
        for (int bi = 0; bi < nblk - di - 1; ++bi)
            for (int ci = 0; ci < blkw; ++ci) {
                out[bi * blkw + ci] += SUP(D, di, bi, ci) * vec[(bi + di + 1) * blkw + ci];
                out[(bi + di + 1) * blkw + ci] += SUB(D, di, bi, ci) * vec[bi * blkw + ci];
            }
    for (int sati = 0; sati < nsat; ++sati)
        for (int bi = 0; bi <= sati; ++bi)
            for (int ci = 0; ci < blkw; ++ci) {
                out[bi * blkw + ci] += TOP(D, sati, bi, ci) * vec[(nblk - 1 - sati + bi) * blkw + ci];
                out[(nblk - 1 - sati + bi) * blkw + ci] += BOT(D, sati, bi, ci) * vec[bi * blkw + ci];
            }
}

/* BlockDiagMatrix::scale_diag_add, hpp:276-287 (dst and src share dims) */
EXPORT void oracle_scale_diag_add(double *data, const double *src, int nblocks, int blockw, int ndiag,
                                  int nsat, int ld, double scale, double diag_add)
{
    dims_t D = mkdims(nblocks, blockw, ndiag, nsat, ld);
    const int64_t nd = oracle_ndata(nblocks, blockw, ndiag, nsat, ld);
    for (int64_t i = 0; i < nd; ++i) data[i] = scale * src[i];
    for (int bi = 0; bi < nblocks; ++bi)
        for (int ci = 0; ci < blockw; ++ci)
            BLOCK(D, bi, ci, ci) += diag_add;
}

/* ---- batched driver used for the "port" CPU baseline and batched parity ------------------
 * Runs factor + nrhs solves over nsys independent systems laid out like the C ABI's batches:
 * data[s*sys_stride ...], b/x[(s*nrhs + r)*n ...], int arrays [s*n ...].  Systems are the
 * OpenMP loop (the reference's own solve is serial, hpp:486-538).  Returns wall seconds;
 * infos[s] = factor info (or first nonzero solve flag + 1000000 if the factor was clean). */
EXPORT double oracle_factor_solve_batch(double *data, int64_t sys_stride, int nsys,
                                        int nblocks, int blockw, int ndiag, int nsat, int ld,
                                        int *ipiv, int *rowbycol, int *colbyrow,
                                        const double *b, double *x, int nrhs, int *infos, int nthreads)
{
    const int n = nblocks * blockw;
    double t0 = 0, t1 = 0;
#ifdef _OPENMP
    if (nthreads < 1) nthreads = omp_get_max_threads();
    t0 = omp_get_wtime();
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 1)
#endif
    for (int s = 0; s < nsys; ++s) {
        double *y = (double *)malloc(sizeof(double) * (size_t)n);
        int info = oracle_factor(data + s * sys_stride, nblocks, blockw, ndiag, nsat, ld,
                                 ipiv + (size_t)s * n, rowbycol + (size_t)s * n, colbyrow + (size_t)s * n);
        for (int r = 0; r < nrhs; ++r) {
            int flag = oracle_solve(data + s * sys_stride, nblocks, blockw, ndiag, nsat, ld,
                                    rowbycol + (size_t)s * n, colbyrow + (size_t)s * n,
                                    b + ((size_t)s * nrhs + r) * n, x + ((size_t)s * nrhs + r) * n, y);
            if (info == 0 && flag != 0) info = 1000000 + flag;
        }
        if (infos) infos[s] = info;
        free(y);
    }
#ifdef _OPENMP
    t1 = omp_get_wtime();
#endif
    return t1 - t0;
}
