// Internal launch interface between the C ABI (bdilu_capi.cu) and the kernel translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "bdilu_common.cuh"

namespace bdilu {

void count_launch(int n);
// SMs of the current device (148 on a full B200), cached per device: grid sizes and the fill thresholds of the solve
// dispatch are expressed in it
int sm_count();
void set_last_error(const char *msg);  // feeds bdilu_last_error() on the calling thread

// argument checks / error channel of the C ABI, shared by the translation units that hold entry points
struct Dims;
int api_check_dims(int nsys, int64_t sys_stride, int nblocks, int blockw, int ndiag, int nsat, int &ld, Dims &D, int esize);
int api_fail(int code, const char *msg);
int api_cuda_fail(cudaError_t e, const char *where);

// strides of right-hand-side / solution vectors (in doubles)
struct VecLayout {
    int nrhs;
    int64_t rhs_stride, sys_stride;
};

int api_check_vecs(const Dims &D, int nsys, int nrhs, int64_t rhs_stride, int64_t sys_rhs_stride);

cudaError_t launch_factor(double *data, int64_t sys_stride, int nsys, const Dims &D, int *ipiv,
                          int *rowbycol, int *colbyrow, int *info, cudaStream_t st);

cudaError_t launch_solve(const double *data, int64_t sys_stride, int nsys, const Dims &D,
                         const int *rowbycol, const int *colbyrow, const double *b, double *x,
                         const VecLayout &V, int *flags, cudaStream_t st);

extern int g_solve_path;  // test hook (7 = quad32; 8 = experimental systolic blockw-32 kernel first, quad32 behind it): 0 = automatic, 1 = chain kernel, 2 = generic wavefront kernel, 3 = one-row wave32 kernel,
                          // 4 = many-vector kernel (one warp per group) whenever the shape allows it, 5 = automatic without the many-vector
                          // kernels, 6 = many-vector kernels (wavefront form where it applies) whenever the shape allows it,
                          // 9 = prefetched narrow-block chain kernel (bdilu_solve_pf.cu) whenever the shape allows it (11 / 12: the same with four rows /
                          // one row per shuffle round forced),
                          // 10 = two-SM blockw-32 kernel (bdilu_solve_half32.cu) first, quad32 behind it, whatever the batch size
// many right-hand sides per factor: a warp carries 8 vectors, blocks staged once per CTA (bdilu_solve_mrhs.cu)
bool try_launch_solve_mrhs(const double *data, int64_t sys_stride, int nsys, const Dims &D, const int *rowbycol,
                           const int *colbyrow, const double *b, double *x, const VecLayout &V, int *flags,
                           cudaStream_t st, cudaError_t *err);
// blockw <= 16 (bdilu_solve_pf.cu): one lane group per vector, every global load prefetched, the chain runs on shuffles
bool try_launch_solve_pf(const double *data, int64_t sys_stride, int nsys, const Dims &D, const int *rowbycol,
                         const int *colbyrow, const double *b, double *x, const VecLayout &V, int *flags, cudaStream_t st,
                         cudaError_t *err);
bool try_launch_solve_wave(const double *data, int64_t sys_stride, int nsys, const Dims &D, const int *rowbycol,
                           const int *colbyrow, const double *b, double *x, const VecLayout &V, int *flags,
                           cudaStream_t st, cudaError_t *err);

cudaError_t launch_solve_wave32(unsigned nvec, size_t smem, int NS, const double *data, int64_t sys_stride, const Dims &D,
                                const int *rowbycol, const int *colbyrow, const double *b, double *x,
                                const VecLayout &V, int *flags, cudaStream_t st);

// status (nullable): int per vector; when given, only the vectors with status != 0 are solved (second pass behind
// the systolic kernel)
bool try_launch_solve_quad32(unsigned nvec, const double *data, int64_t sys_stride, const Dims &D, const int *rowbycol,
                             const int *colbyrow, const double *b, double *x, const VecLayout &V, int *flags,
                             cudaStream_t st, cudaError_t *err, const int *status = nullptr);

// blockw == 32 on two SMs per chain (bdilu_solve_half32.cu): rows 0..15 and 16..31 of every block on the two CTAs of a
// cluster, the dependency between them one-directional per sweep.  Solves the vectors whose systems never pivoted across the
// middle of a block and marks the others in `status` (int per vector) for quad32; `scratch`: nvec * half32_scratch_ints() ints.
int half32_scratch_ints();
bool try_launch_solve_half32(unsigned nvec, const double *data, int64_t sys_stride, int nsys, const Dims &D,
                             const int *rowbycol, const int *colbyrow, const double *b, double *x, const VecLayout &V,
                             int *flags, int *status, int *scratch, cudaStream_t st, cudaError_t *err);

// blockw == 32 systolic solve (bdilu_solve_skew.cu; experimental, opt-in): lane = block, consecutive blocks one row
// apart.  Handles the vectors whose permutations are the identity (status == 0 after the permutation scan) and hands
// the others back through status.  Returns false when the shape does not fit.
bool try_launch_solve_skew32(unsigned nvec, const double *data, int64_t sys_stride, const Dims &D, const int *rowbycol,
                             const int *colbyrow, const double *b, double *x, const VecLayout &V, int *flags,
                             int *status, cudaStream_t st, cudaError_t *err);

bool try_launch_solve_mrhs_wave(int Q, const double *data, int64_t sys_stride, int nsys, const Dims &D, const int *rowbycol,
                                const int *colbyrow, const double *b, double *x, const VecLayout &V, int *flags,
                                cudaStream_t st, cudaError_t *err);

// counters of the systolic solve (test hook): solved / handed back at the end / skipped at the start
cudaError_t skew_stats(unsigned long long out[3], bool reset);

cudaError_t launch_dot_vec(const double *data, int64_t sys_stride, int nsys, const Dims &D,
                           const double *vec, double *out, const VecLayout &V, cudaStream_t st);

cudaError_t launch_residual_norm2(const double *data, int64_t sys_stride, int nsys, const Dims &D,
                                  const double *b, const double *x, const VecLayout &V, double *norm2,
                                  cudaStream_t st);

// the same with the partial sums in a file-scope device array: no allocation of any kind (bdilu_residual_norm2_static)
int64_t residual_norm2_static_partials(int nsys, const Dims &D, const VecLayout &V);
cudaError_t launch_residual_norm2_static(const double *data, int64_t sys_stride, int nsys, const Dims &D,
                                         const double *b, const double *x, const VecLayout &V, double *norm2,
                                         cudaStream_t st);

cudaError_t launch_scale_diag_add(double *dst, const double *src, int64_t dst_stride, int64_t src_stride, int nsys,
                                  const Dims &D, double scale, double diag_add, cudaStream_t st);

// dst = ILU factors of scale*src + diag_add*I, assembled in the factor kernel's load (blockw <= 32)
cudaError_t launch_factor_scaled(double *dst, int64_t dst_stride, const double *src, int64_t src_stride, int nsys,
                                 const Dims &D, double scale, double diag_add, int *ipiv, int *rowbycol, int *colbyrow,
                                 int *info, cudaStream_t st);

}  // namespace bdilu
