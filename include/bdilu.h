/*
 * bdilu.h -- C ABI of the B200-native block_diag_ilu hot path (libbdilu_b200.so).
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++ or torch types.  Every entry point
 * names the reference interface it replaces (paths relative to the chemreac/block_diag_ilu tree).
 * The reference has no FFI for this path today -- callers compile against the header-only C++
 * templates (block_diag_ilu/include/block_diag_ilu.hpp) or the Cython wrapper
 * (block_diag_ilu/_block_diag_ilu.pyx) -- so these are exactly the calls a maintainer would bind
 * from those two places; INTEGRATION.md shows both bindings.
 *
 * Data layout (identical to BlockDiagMatrix<double>::m_data, block_diag_ilu.hpp:98-102, 203-221):
 *   one system = [blocks: nblocks*blockw*ld][sub: D*ld][sup: D*ld][bot: S*ld][top: S*ld] doubles,
 *   D = ndiag*nblocks - (ndiag*ndiag+ndiag)/2, S = (nsat*nsat+nsat)/2,
 *   block(b,r,c) = data[(b*blockw + c)*ld + r]   (column-major blocks, ld >= blockw)
 *   sub(di,b,c)  = data[blk + (di*nblocks - (di*di+di)/2 + b)*ld + c];  sup = sub + D*ld.
 * A batch is nsys such systems `sys_stride` doubles apart (sys_stride >= bdilu_ndata(...)).
 * Index arrays are int32, laid out [nsys][nblocks*blockw], block-local values as in the reference
 * (ipiv 1-based LAPACK convention; rowbycol/colbyrow 0-based, block_diag_ilu.hpp:34-52).
 * Right-hand sides / solutions: vector r of system s starts at
 *   b + s*sys_rhs_stride + r*rhs_stride   (each vector contiguous, nblocks*blockw doubles).
 *
 * Device entry points (bdilu_factor ... bdilu_residual_norm2) take DEVICE pointers and a
 * cudaStream_t passed as void*; they are asynchronous, reentrant and never synchronise; the only
 * allocation is stream-ordered scratch of one int per vector inside bdilu_solve at blockw 32 (cudaMallocAsync /
 * cudaFreeAsync on the caller's stream).  Host entry points (bdilu_ilu_*, bdilu_host_*) take HOST pointers, own the device
 * mirrors, and block until the result is in host memory (the reference's calls are blocking).
 *
 * Limits: blockw <= BDILU_MAX_BLOCKW; nblocks < 2^24; any number of systems; in bdilu_dot_vec / bdilu_residual_norm2
 * at most 65535 groups of 8 vectors per system.  bdilu_residual_norm2 allocates stream-ordered scratch too.
 *
 * Return value of every int function: 0 = ok, <0 = BDILU_ERR_* (see bdilu_last_error()).
 * There is no CPU fallback: without a CUDA device every compute entry point fails with
 * BDILU_ERR_CUDA.
 */
#ifndef BDILU_H
#define BDILU_H

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define BDILU_API __attribute__((visibility("default")))
#else
#define BDILU_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define BDILU_VERSION 100            /* 0.1.0 */
#define BDILU_MAX_BLOCKW 64

#define BDILU_OK 0
#define BDILU_ERR_ARG (-1)           /* bad dimension / null pointer / stride too small */
#define BDILU_ERR_UNSUPPORTED (-2)   /* blockw > BDILU_MAX_BLOCKW */
#define BDILU_ERR_CUDA (-3)          /* CUDA runtime error (no device, launch failure, OOM) */

/* Factor info word, one per system.  0 = every block factored.  Otherwise the FIRST failing block
 * in block order (what the serial reference loop records in info_, block_diag_ilu.hpp:469-470):
 *   BDILU_INFO_PIVOT(info) = LAPACK dgetrf info of that block (1-based index of the zero pivot)
 *   BDILU_INFO_BLOCK(info) = index of that block.
 * The reference throws std::runtime_error("ILU failed!") when info != 0 (:483-484). */
#define BDILU_INFO_PIVOT(info) ((info) & 127)
#define BDILU_INFO_BLOCK(info) ((info) >> 7)

BDILU_API int bdilu_version(void);
/* Message of the last error on the calling thread ("" if none). */
BDILU_API const char *bdilu_last_error(void);
/* Number of CUDA devices visible (0 when there is no usable device). */
BDILU_API int bdilu_device_count(void);

/* ---- layout helpers: BlockDiagMatrix ctor arithmetic, block_diag_ilu.hpp:98-102 ------------- */
/* default leading dimension: blockw rounded up to 64 bytes (hpp:98; anyode_matrix.hpp:18-21) */
BDILU_API int bdilu_default_ld(int blockw);
/* m_ndata for the given shape (ld == 0 -> default ld) */
BDILU_API int64_t bdilu_ndata(int nblocks, int blockw, int ndiag, int nsat, int ld);

/* ---- device entry points ------------------------------------------------------------------- */

/* ILU_inplace<double>::ILU_inplace(view)  (block_diag_ilu.hpp:430-485) for a batch of systems:
 * per diagonal block LU with partial pivoting (replaces dgetrf_, anyode_blas_lapack.hpp:55-57),
 * sub(di,b,c) /= U_b[c,c] (:471-475), bot(sati,b,c) /= U_b[c,c] (:476-479), ipiv -> rowbycol ->
 * colbyrow (:480-481).  `data` is overwritten with L\U and the scaled sub/bot, as in the reference.
 * info: int32[nsys], written by the call (see BDILU_INFO_*). */
BDILU_API int bdilu_factor(double *data, int64_t sys_stride, int nsys,
                 int nblocks, int blockw, int ndiag, int nsat, int ld,
                 int *ipiv, int *rowbycol, int *colbyrow, int *info, void *stream);

/* ILU_inplace<double>::solve(b, x)  (block_diag_ilu.hpp:486-538) for nsys systems x nrhs vectors.
 * flags: int32[nsys*nrhs] with exactly the reference's return value per (system, vector):
 *   0 ok; 1..n = 1-based index of the first NaN in b; n + k = zero U diagonal at 0-based row k
 *   (first error wins in the reference's loop order, :500-501, 533-534).  x is always written.
 * b and x must not alias (they are __restrict__ in the reference).
 * The nrhs vectors of a system share its factor: batches of them (any number at blockw > 32, >= 8 otherwise, size permitting)
 * run on kernels that read every block once per group of 2-8 vectors instead of once per vector.
 * y_scratch: nullable; unused by the current kernels (y stays on chip), kept for ABI stability. */
BDILU_API int bdilu_solve(const double *data, int64_t sys_stride, int nsys,
                int nblocks, int blockw, int ndiag, int nsat, int ld,
                const int *rowbycol, const int *colbyrow,
                const double *b, double *x, int nrhs, int64_t rhs_stride, int64_t sys_rhs_stride,
                double *y_scratch, int *flags, void *stream);

/* ILU<double>(A) followed by solve(b, x)  (block_diag_ilu.hpp:541-551) without the deep copy:
 * factor in place, then solve, enqueued back to back on `stream`. */
BDILU_API int bdilu_factor_solve(double *data, int64_t sys_stride, int nsys,
                       int nblocks, int blockw, int ndiag, int nsat, int ld,
                       int *ipiv, int *rowbycol, int *colbyrow, int *info,
                       const double *b, double *x, int nrhs, int64_t rhs_stride, int64_t sys_rhs_stride,
                       int *flags, void *stream);

/* BlockDiagMatrix<double>::dot_vec(vec, out)  (block_diag_ilu.hpp:295-326): out = A*vec on the
 * UNFACTORED matrix, nsys systems x nrhs vectors (same vector strides as bdilu_solve). */
BDILU_API int bdilu_dot_vec(const double *data, int64_t sys_stride, int nsys,
                  int nblocks, int blockw, int ndiag, int nsat, int ld,
                  const double *vec, double *out, int nrhs, int64_t rhs_stride, int64_t sys_rhs_stride,
                  void *stream);

/* New (not in the reference, built on dot_vec for validation): norm2[s*nrhs + r] = ||b - A x||_2^2
 * of the UNFACTORED matrix.  norm2 is overwritten; the sum is formed in a fixed order (bit-reproducible). */
BDILU_API int bdilu_residual_norm2(const double *data, int64_t sys_stride, int nsys,
                         int nblocks, int blockw, int ndiag, int nsat, int ld,
                         const double *b, const double *x, int nrhs, int64_t rhs_stride,
                         int64_t sys_rhs_stride, double *norm2, void *stream);

/* bdilu_residual_norm2 without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments and results (bit-identical) as bdilu_residual_norm2, but the per-CTA partial sums
 * live in a file-scope device array of BDILU_STATIC_NORM2_PARTIALS doubles (1 MiB, held by every GPU that loads the
 * library until exit) instead of stream-ordered scratch, and the call makes no cudaMalloc / cudaFree / cudaMallocAsync
 * of any kind.  Limit: nsys * nrhs * ceil(nblocks*blockw / 128) <= BDILU_STATIC_NORM2_PARTIALS, otherwise
 * BDILU_ERR_UNSUPPORTED with a message naming both numbers.  Every call on a device shares the one array: calls that
 * overlap on the GPU (different streams or threads without ordering between them) race; calls ordered on one stream are
 * safe.  Nothing in the array needs to start at zero (every slot read is written by the same call first). */
#define BDILU_STATIC_NORM2_PARTIALS 131072
BDILU_API int bdilu_residual_norm2_static(const double *data, int64_t sys_stride, int nsys,
                                int nblocks, int blockw, int ndiag, int nsat, int ld,
                                const double *b, const double *x, int nrhs, int64_t rhs_stride,
                                int64_t sys_rhs_stride, double *norm2, void *stream);

/* BlockDiagMatrix<double>::scale_diag_add(source, scale, diag_add)  (block_diag_ilu.hpp:276-287;
 * set_to_eye_plus_scaled_mtx(scale, source) == scale_diag_add(source, scale, 1), :288-290):
 * dst = scale*src, then diag_add added to every diagonal element of every block.  dst may equal src. */
BDILU_API int bdilu_scale_diag_add(double *dst, const double *src, int64_t sys_stride, int nsys,
                         int nblocks, int blockw, int ndiag, int nsat, int ld,
                         double scale, double diag_add, void *stream);

/* Fused assembly + factorisation, the pair an implicit integrator runs at every Jacobian refresh
 * (`m.set_to_eye_plus_scaled_mtx(-gamma, J); ILU_inplace ilu(&m);`, block_diag_ilu.hpp:276-290 + 430-485;
 * caller external/anyode/include/anyode/anyode_iterative.hpp:58-62): dst receives the ILU factors of
 * scale*src + diag_add*I.  src (the Jacobian, nsys systems src_stride doubles apart) is only read and stays
 * valid for the next gamma; it is scaled inside the factor kernel's load, so the separate pass of
 * bdilu_scale_diag_add over the matrix (one full read + write) disappears.  dst may equal src.  Results are
 * identical to bdilu_scale_diag_add followed by bdilu_factor, except that padding elements of dst
 * (rows >= blockw when ld > blockw) are left untouched.  Outputs as bdilu_factor. */
BDILU_API int bdilu_factor_scaled(double *dst, int64_t dst_stride, const double *src, int64_t src_stride, int nsys,
                        int nblocks, int blockw, int ndiag, int nsat, int ld,
                        double scale, double diag_add,
                        int *ipiv, int *rowbycol, int *colbyrow, int *info, void *stream);

/* Scratch bytes the device entry points need from the caller: always 0 today (y stays on chip). */
BDILU_API size_t bdilu_workspace_bytes(int nsys, int nblocks, int blockw, int ndiag, int nsat, int ld, int nrhs);

/* ---- host entry points (what the C++ header shim and the Python wrapper bind) --------------- */

typedef struct bdilu_ilu bdilu_ilu;  /* opaque: device mirrors of one factored batch */

/* ILU_inplace ctor on HOST memory: copies the batch to the device, factors there, and keeps the
 * factors device-resident for later solves.  If copy_back != 0, host_data is overwritten with the
 * factors (ILU_inplace semantics, the caller's view is factored in place); ipiv/rowbycol/colbyrow
 * /info (host, nullable) receive the index arrays (m_ipiv, m_rowbycol, m_colbyrow, :428-429).
 * Returns 0 and a handle even when some info[s] != 0 (the caller decides whether to throw). */
BDILU_API int bdilu_ilu_create(bdilu_ilu **out, double *host_data, int64_t sys_stride, int nsys,
                     int nblocks, int blockw, int ndiag, int nsat, int ld, int copy_back,
                     int *ipiv, int *rowbycol, int *colbyrow, int *info);
/* ILU_inplace::solve on HOST vectors: b, x are [nsys][nrhs][n] contiguous; flags int32[nsys*nrhs]
 * (nullable).  Returns the first nonzero flag (the reference's return value when nsys*nrhs == 1)
 * or a negative BDILU_ERR_*. */
BDILU_API int bdilu_ilu_solve(bdilu_ilu *h, const double *host_b, double *host_x, int nrhs, int *flags);
BDILU_API void bdilu_ilu_destroy(bdilu_ilu *h);

/* BlockDiagMatrix::dot_vec on HOST memory (one system x nrhs vectors). */
BDILU_API int bdilu_host_dot_vec(const double *host_data, int nblocks, int blockw, int ndiag, int nsat, int ld,
                       const double *host_vec, double *host_out, int nrhs);

/* BlockDiagMatrix::scale_diag_add / set_to_eye_plus_scaled_mtx on HOST memory (block_diag_ilu.hpp:276-290),
 * one system; host_dst may equal host_src. */
BDILU_API int bdilu_host_scale_diag_add(double *host_dst, const double *host_src, int nblocks, int blockw,
                              int ndiag, int nsat, int ld, double scale, double diag_add);

/* One-shot ILU(A) + solve(b, x) on HOST memory for a batch: H2D of data and b, fused factor +
 * solve, D2H of x (and flags / info).  host_data is NOT modified (ILU copies its source). */
BDILU_API int bdilu_host_factor_solve(const double *host_data, int64_t sys_stride, int nsys,
                            int nblocks, int blockw, int ndiag, int nsat, int ld,
                            const double *host_b, double *host_x, int nrhs, int *info, int *flags);

/* ---- Real_t = float ---------------------------------------------------------------------------------------------
 * The reference's classes are templates over Real_t (block_diag_ilu.hpp:63, 104, 426; float build of its tests:
 * tests/test_block_diag.cpp:649-712; LAPACK side sgetrf_, anyode_blas_lapack.hpp:56).  The *_f32 entry points are the
 * float instantiation: the same packed layout with float elements (default ld = blockw rounded up to 64 BYTES, i.e. to
 * a multiple of 16 floats: bdilu_default_ld_f32; bdilu_ndata(..., ld) counts elements whatever their type), the same
 * int32 index arrays, info words and solve flags, the same argument meaning and error codes as their double
 * namesakes above.  Kernels: csrc/bdilu_f32.cu (one lane group per block / per vector; the wavefront and many-vector
 * solve kernels are double-only). */
BDILU_API int bdilu_default_ld_f32(int blockw);
/* ILU_inplace<float>::ILU_inplace(view)  (block_diag_ilu.hpp:430-485) */
BDILU_API int bdilu_factor_f32(float *data, int64_t sys_stride, int nsys,
                     int nblocks, int blockw, int ndiag, int nsat, int ld,
                     int *ipiv, int *rowbycol, int *colbyrow, int *info, void *stream);
/* ILU_inplace<float>::solve(b, x)  (block_diag_ilu.hpp:486-538) */
BDILU_API int bdilu_solve_f32(const float *data, int64_t sys_stride, int nsys,
                    int nblocks, int blockw, int ndiag, int nsat, int ld,
                    const int *rowbycol, const int *colbyrow,
                    const float *b, float *x, int nrhs, int64_t rhs_stride, int64_t sys_rhs_stride,
                    int *flags, void *stream);
/* ILU<float>(A) + solve(b, x)  (block_diag_ilu.hpp:541-551), in place, back to back on `stream` */
BDILU_API int bdilu_factor_solve_f32(float *data, int64_t sys_stride, int nsys,
                           int nblocks, int blockw, int ndiag, int nsat, int ld,
                           int *ipiv, int *rowbycol, int *colbyrow, int *info,
                           const float *b, float *x, int nrhs, int64_t rhs_stride, int64_t sys_rhs_stride,
                           int *flags, void *stream);
/* BlockDiagMatrix<float>::dot_vec  (block_diag_ilu.hpp:295-326); at most 65535 vectors per system */
BDILU_API int bdilu_dot_vec_f32(const float *data, int64_t sys_stride, int nsys,
                      int nblocks, int blockw, int ndiag, int nsat, int ld,
                      const float *vec, float *out, int nrhs, int64_t rhs_stride, int64_t sys_rhs_stride,
                      void *stream);
/* BlockDiagMatrix<float>::scale_diag_add  (block_diag_ilu.hpp:276-287) */
BDILU_API int bdilu_scale_diag_add_f32(float *dst, const float *src, int64_t sys_stride, int nsys,
                             int nblocks, int blockw, int ndiag, int nsat, int ld,
                             float scale, float diag_add, void *stream);
/* host forms, as bdilu_ilu_create / bdilu_ilu_solve / bdilu_ilu_destroy / bdilu_host_dot_vec / bdilu_host_scale_diag_add */
typedef struct bdilu_ilu_f32 bdilu_ilu_f32;
BDILU_API int bdilu_ilu_create_f32(bdilu_ilu_f32 **out, float *host_data, int64_t sys_stride, int nsys,
                         int nblocks, int blockw, int ndiag, int nsat, int ld, int copy_back,
                         int *ipiv, int *rowbycol, int *colbyrow, int *info);
BDILU_API int bdilu_ilu_solve_f32(bdilu_ilu_f32 *h, const float *host_b, float *host_x, int nrhs, int *flags);
BDILU_API void bdilu_ilu_destroy_f32(bdilu_ilu_f32 *h);
BDILU_API int bdilu_host_dot_vec_f32(const float *host_data, int nblocks, int blockw, int ndiag, int nsat, int ld,
                           const float *host_vec, float *host_out, int nrhs);
BDILU_API int bdilu_host_scale_diag_add_f32(float *host_dst, const float *host_src, int nblocks, int blockw,
                                  int ndiag, int nsat, int ld, float scale, float diag_add);

/* Kernels launched by this library since load / last reset (bench.py's gpu_launches claim). */
BDILU_API int64_t bdilu_launch_count(void);
BDILU_API void bdilu_launch_count_reset(void);

#ifdef __cplusplus
}
#endif
#endif /* BDILU_H */
