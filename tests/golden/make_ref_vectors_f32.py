#!/usr/bin/env python
"""Generates tests/golden/ref_vectors_f32.npz from the UNMODIFIED reference header instantiated with Real_t = float.

Run in the build container (where /root/reference exists):

    python tests/golden/make_ref_vectors_f32.py

oracle/_ref/libbdilu_ref.so (oracle/ref_shim.cpp compiled against block_diag_ilu/include/block_diag_ilu.hpp where it
lies; LAPACK sgetrf from scipy's OpenBLAS) runs BlockDiagMatrix<float>::dot_vec, the ILU_inplace<float> constructor and
solve on seeded inputs (the float64 recipes of block_diag_ilu_b200.synth rounded to float32); inputs and outputs are
stored.  Families D and R only: their pivots are separated by orders of magnitude, so a single-precision rounding
difference between two correct eliminations cannot flip one (family P's near-ties can, at float's 6e-8).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from block_diag_ilu_b200 import synth  # noqa: E402
from oracle.loader import RefF32  # noqa: E402

# name: (nblocks, blockw, ndiag, ld, nsat, family, seed)
CASES = {
    "cfg1_D": (64, 4, 1, 16, 0, "D", 1_000_000),
    "cfg1_R_tight": (64, 4, 1, 4, 0, "R", 1_000_001),
    "w1_D": (9, 1, 2, 1, 0, "D", 11),
    "w5_ld8_R": (7, 5, 3, 8, 0, "R", 12),
    "w16_ld19_R": (6, 16, 1, 19, 0, "R", 13),
    "w8_sat2_R": (9, 8, 2, 8, 2, "R", 14),
    "w32_R": (6, 32, 2, 32, 0, "R", 3_000_001),
    "w32_D": (6, 32, 2, 32, 0, "D", 3_000_000),
    "w64_R": (3, 64, 2, 64, 0, "R", 4_000_001),
    "w48_ld56_D": (3, 48, 2, 56, 0, "D", 15),
}


def main():
    ref = RefF32()
    out = {}
    for name, (N, w, d, ld, nsat, fam, seed) in CASES.items():
        data, b = synth.make_system(N, w, d, fam, seed, ld=ld, nsat=nsat)
        data, b = data.astype(np.float32), b.astype(np.float32)
        out[name + "/shape"] = np.array([N, w, d, ld, nsat, seed], np.int64)
        out[name + "/family"] = np.array(fam)
        out[name + "/data"] = data.copy()
        out[name + "/b"] = b.copy()
        av = ref.dot_vec(data.copy(), N, w, d, nsat, ld, b)
        fac = data.copy()
        ipiv, rbc, cbr, info = ref.factor(fac, N, w, d, nsat, ld)
        assert info == 0
        x, flag = ref.solve(fac, N, w, d, nsat, ld, rbc, cbr, b)
        assert flag == 0
        out[name + "/factored"] = fac.copy()
        out[name + "/ipiv"] = ipiv
        out[name + "/rowbycol"] = rbc
        out[name + "/colbyrow"] = cbr
        out[name + "/x"] = x
        out[name + "/dot_vec"] = av
        print("%-14s N=%d w=%d d=%d ld=%d nsat=%d %s  non-identity pivots: %d/%d" % (
            name, N, w, d, ld, nsat, fam, int((ipiv != np.tile(np.arange(1, w + 1), N)).sum()), N * w))
    ref.release()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_vectors_f32.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
