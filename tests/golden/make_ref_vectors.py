#!/usr/bin/env python
"""Generates tests/golden/ref_vectors.npz from the UNMODIFIED reference header.

Run in the build container (where /root/reference exists):

    python tests/golden/make_ref_vectors.py

It loads oracle/_ref/libbdilu_ref.so (oracle/build_oracle.py compiles oracle/ref_shim.cpp against
block_diag_ilu/include/block_diag_ilu.hpp where it lies, LAPACK dgetrf from scipy's OpenBLAS), runs
ILU_inplace ctor + solve + dot_vec on seeded inputs, and stores inputs and outputs.  The GPU box has no
/root/reference: tests read only the committed .npz.
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from block_diag_ilu_b200 import synth  # noqa: E402
from oracle.loader import Ref  # noqa: E402

# name: (nblocks, blockw, ndiag, ld, nsat, family, seed, store_inputs)
CASES = {
    "cfg1_D": (64, 4, 1, 4, 0, "D", 1_000_000, True),
    "cfg1_P": (64, 4, 1, 4, 0, "P", 1_000_001, True),
    "w1_P": (9, 1, 2, 1, 0, "P", 11, True),
    "w5_ld8_P": (7, 5, 3, 8, 0, "P", 12, True),
    "w16_ld19_P": (6, 16, 1, 19, 0, "P", 13, True),
    "w8_sat2_D": (9, 8, 2, 8, 2, "D", 14, True),
    "w32_P": (6, 32, 2, 32, 0, "P", 3_000_001, False),
    "w32_D": (6, 32, 2, 32, 0, "D", 3_000_000, False),
    "w64_P": (3, 64, 2, 64, 0, "P", 4_000_001, False),
    "w48_ld56_D": (3, 48, 2, 56, 0, "D", 15, False),
}


def main():
    ref = Ref()
    out = {}
    for name, (N, w, d, ld, nsat, fam, seed, keep) in CASES.items():
        data, b = synth.make_system(N, w, d, fam, seed, ld=ld, nsat=nsat)
        out[name + "/shape"] = np.array([N, w, d, ld, nsat, seed], np.int64)
        out[name + "/family"] = np.array(fam)
        out[name + "/input_sha256"] = np.array(hashlib.sha256(data.tobytes() + b.tobytes()).hexdigest())
        if keep:
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
        print("%-12s N=%d w=%d d=%d ld=%d nsat=%d %s  non-identity pivots: %d/%d" % (
            name, N, w, d, ld, nsat, fam, int((ipiv != np.tile(np.arange(1, w + 1), N)).sum()), N * w))
    ref.release()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_vectors.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
