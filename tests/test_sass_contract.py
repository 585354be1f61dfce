"""Build-time checks on the SASS of the shipped library (run on CPU: cuobjdump reads the object files nvcc left in
block_diag_ilu_b200/build/).

The lean chain of the blockw-32 solve kernel (csrc/bdilu_solve_quad32.cu) publishes a finished row with a PREDICATED
shared-memory store and reads it back with a broadcast load in the same converged warp, with no barrier in between: it
is correct only while the compiler predicates that store instead of branching around it.  This pins the
property: if a toolchain change turns the 64 chain stores into divergent branches, this test fails before any GPU run.
It also pins the instructions the design rests on: bulk-copy (TMA) staging and mbarriers in the wavefront kernels, FP64
FMAs and warp reductions in the factor kernel, cp.async in the narrow-block kernel, and the absence of tensor-core ops.
"""
import os
import re
import shutil
import subprocess

import pytest

from block_diag_ilu_b200 import _build

BUILD = os.path.join(os.path.dirname(_build.LIB), "build")


def sass(obj, function_substring):
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(exe):
        pytest.skip("cuobjdump not available")
    _build.build()
    path = os.path.join(BUILD, obj)
    if not os.path.exists(path):
        pytest.skip("%s not built here (prebuilt library without objects)" % obj)
    out = subprocess.run([exe, "-sass", path], capture_output=True, text=True, check=True).stdout
    chunks = re.split(r"\n\s*Function : ", out)
    mine = [c for c in chunks if function_substring in c.split("\n", 1)[0]]
    assert mine, "no function matching %r in %s" % (function_substring, obj)
    return "\n".join(mine)


def count(text, pattern):
    return len(re.findall(pattern, text))


def test_quad32_chain_stores_are_predicated():
    s = sass("bdilu_solve_quad32.o", "solve_quad32_kernelILi2ELi2ELb1")  # ND = 2, coupling merged every 2 rows, ld = 32
    predicated = count(s, r"@!?P\d\s+STS\.64")
    assert predicated >= 64, "the 2 x 32 row publications must be predicated stores, found %d" % predicated
    assert count(s, r"\bUBLKCP\b") >= 1 and count(s, r"\bSYNCS\b") >= 4      # TMA bulk copies + mbarriers
    assert count(s, r"\bDFMA\b") >= 64
    assert count(s, r"HMMA|IMMA|UTCMMA|UTCHMMA|DMMA") == 0                     # north_star: no tensor cores on this path


def test_factor_kernel_instruction_mix():
    s = sass("bdilu_factor.o", "factor_kernelILi32ELb1ELb0")
    assert count(s, r"\bDFMA\b") > 500
    assert count(s, r"\bCREDUX\b|\bREDUX\b") >= 90        # general loop: three reductions per elimination step
    assert count(s, r"\bSHFL\b") >= 60                     # optimistic 16 x 2 grid: multipliers handed over by shuffle
    assert count(s, r"HMMA|IMMA|UTCMMA|DMMA") == 0


def test_narrow_block_kernel_uses_cp_async():
    s = sass("bdilu_solve_pf.o", "solve_chain_pf_kernelILi8ELi2E")
    assert count(s, r"\bLDGSTS\b") >= 20 and count(s, r"\bLDGDEPBAR\b|\bDEPBAR\b") >= 2
    assert count(s, r"\bSHFL\b") >= 40
