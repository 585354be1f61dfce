#!/usr/bin/env python
"""Build recipe for the parity oracle (TEST INFRASTRUCTURE, not product code).

  oracle/_build/libbdilu_oracle.so   gcc build of oracle/bdilu_oracle.c (the C restatement, "port")
  oracle/_ref/libbdilu_ref.so        g++ build of oracle/ref_shim.cpp against the UNMODIFIED
                                     reference header where it lies under /root/reference,
                                     with -fopenmp -DBLOCK_DIAG_ILU_WITH_OPENMP (the macro the header
                                     really tests, block_diag_ilu.hpp:441) and LAPACK = scipy's bundled
                                     LP64 OpenBLAS (no system liblapack in this image), renamed at
                                     compile time (-Ddgetrf_=scipy_dgetrf_ ...).

The reference's own build system (setup.py / CMake) is NOT run: its setup.py never enables OpenMP
(SURVEY.md section 0.4).  Both outputs are git-ignored and travel to the GPU box with the gpurun
snapshot; /root/reference does not exist there, so the _ref build is skipped when it is absent and
the prebuilt file is used.
"""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE = os.environ.get("BDILU_REFERENCE_ROOT", "/root/reference")
LAPACK_SYMS = ["dgetrf", "sgetrf", "dgetrs", "sgetrs", "dgemv", "sgemv", "dgbmv", "sgbmv",
               "dgbtrf", "sgbtrf", "dgbtrs", "sgbtrs", "dgesvd", "sgesvd"]


def _newer(target, sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


def scipy_openblas():
    import scipy
    libdir = os.path.join(os.path.dirname(os.path.dirname(scipy.__file__)), "scipy.libs")
    libs = sorted(glob.glob(os.path.join(libdir, "libscipy_openblas-*.so")))
    if not libs:
        raise RuntimeError("scipy's bundled OpenBLAS not found under %s" % libdir)
    return libdir, os.path.basename(libs[0])


def build_port(verbose=True):
    out_dir = os.path.join(HERE, "_build")
    os.makedirs(out_dir, exist_ok=True)
    src = os.path.join(HERE, "bdilu_oracle.c")
    out = os.path.join(out_dir, "libbdilu_oracle.so")
    if _newer(out, [src]):
        return out
    # -ffp-contract=off: the reference is compiled -O2 by g++ on x86-64 where no FMA contraction
    # happens by default without -march flags; keep the restatement's rounding the same.
    cmd = ["gcc", "-O2", "-std=c11", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off",
           "-fvisibility=hidden", src, "-o", out, "-lm"]
    if verbose:
        print(" ".join(cmd))
    subprocess.check_call(cmd)
    return out


def build_ref(verbose=True):
    """Returns the path of libbdilu_ref.so, or None when neither sources nor a prebuilt file exist."""
    out_dir = os.path.join(HERE, "_ref")
    out = os.path.join(out_dir, "libbdilu_ref.so")
    hdr = os.path.join(REFERENCE, "block_diag_ilu", "include", "block_diag_ilu.hpp")
    if not os.path.exists(hdr):
        return out if os.path.exists(out) else None
    src = os.path.join(HERE, "ref_shim.cpp")
    if _newer(out, [src, hdr]):
        return out
    os.makedirs(out_dir, exist_ok=True)
    libdir, libname = scipy_openblas()
    cmd = ["g++", "-O2", "-DNDEBUG", "-std=c++14", "-fPIC", "-shared", "-fopenmp",
           "-DBLOCK_DIAG_ILU_WITH_OPENMP=1", "-fvisibility=hidden",
           "-I" + os.path.join(REFERENCE, "block_diag_ilu", "include"),
           "-I" + os.path.join(REFERENCE, "external", "anyode", "include")]
    cmd += ["-D%s_=scipy_%s_" % (s, s) for s in LAPACK_SYMS]
    cmd += [src, "-o", out, "-L" + libdir, "-l:" + libname, "-Wl,-rpath," + libdir]
    if verbose:
        print(" ".join(cmd))
    subprocess.check_call(cmd)
    return out


if __name__ == "__main__":
    print("port:", build_port())
    print("ref :", build_ref())
    sys.exit(0)
