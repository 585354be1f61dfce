/*
 * bdilu_internal.h -- test and measurement hooks of libbdilu_b200.so.  NOT part of the drop-in boundary
 * (include/bdilu.h): nothing here corresponds to a reference interface, callers of the reference never see it.
 * The hooks are process-global and not thread-safe; tests/ and bench.py are their only users.
 */
#ifndef BDILU_INTERNAL_H
#define BDILU_INTERNAL_H

#include <stdint.h>

#include "bdilu.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Solve-kernel selection: 0 = automatic; 1 = general chain kernel; 2 = generic wavefront kernel; 3 = one-row wave32
 * kernel; 4 / 6 = many-vector kernels (one warp per group / wavefront form) whenever the shape allows; 5 = automatic
 * without the many-vector kernels; 7 = same as 0 at blockw 32 (quad32); 8 = blockw 32: experimental systolic kernel first, quad32 behind it for the
 * vectors it hands back (csrc/bdilu_solve_skew.cu; slower than quad32 on B200, kept for the record); 9 = prefetched narrow-block
 * chain kernel (csrc/bdilu_solve_pf.cu) whenever the shape allows it (11 / 12: the same with four rows / one row per shuffle round
 * forced); 10 = blockw 32: experimental two-SMs-per-chain kernel first
 * (csrc/bdilu_solve_half32.cu), quad32 behind it for the systems that pivoted across the middle of a block.
 * Every path is parity-tested.  Returns the previous setting. */
BDILU_API int bdilu_debug_force_chain_solve(int on);

/* Vectors seen by the experimental systolic blockw-32 solve (hook 8) since load / last reset: out[0] = solved by it, out[1] = handed
 * to the general kernel at the end (NaN in b, U diagonal outside the fast reciprocal's range), out[2] = skipped at
 * the start (the system pivoted).  Synchronises the device. */
BDILU_API int bdilu_debug_skew_stats(int64_t *out3, int reset);

#ifdef __cplusplus
}
#endif
#endif /* BDILU_INTERNAL_H */
