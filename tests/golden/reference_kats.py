"""Known-answer vectors held by the reference's own tests for the hot path, transcribed by hand.

Every entry cites where the numbers live in the chemreac/block_diag_ilu tree (paths relative to
/root/reference).  Data are given as the packed BlockDiagMatrix::m_data with TIGHT ld = blockw
(blocks column-major, then sub, then sup; block_diag_ilu.hpp:99-102, 203-212).
Expression-valued expectations (e.g. .2/(63+17/37.*13)) are kept as Python expressions so that they
round exactly like the C++ literals they come from.
"""
import numpy as np

# ---- permutation helpers: tests/test_block_diag.cpp:110-130 -----------------------------------
PIV_1BASED = [3, 4, 5, 4, 5]
ROWBYCOL = [2, 3, 4, 1, 0]
COLBYROW = [4, 3, 0, 1, 2]

# ---- m1: python_prototype/tests/test_fakelu.py:28-43, 121-125 (3x3 dense block, pivots) ------
M1 = dict(
    nblocks=1, blockw=3, ndiag=0,
    dense=np.array([[1, 3, 5], [2, 4, 7], [1, 1, 0]], dtype=np.float64),
    piv0=[1, 1, 2],  # 0-based pivots
    lu_merged=np.array([[2, 4, 7], [.5, 1, 1.5], [.5, -1, -2]], dtype=np.float64),
)

# ---- m2: tests/test_block_diag.cpp:7-39 (matrix), 132-164 (factor + solve), 394-415 (matvec);
#      python_prototype/tests/test_fakelu.py:46-63, 127-132 (merged LU, identity pivots) --------
M2 = dict(
    nblocks=3, blockw=2, ndiag=1,
    data=np.array([5, 5, 3, 8, 8, 4, 4, 4, 6, 2, 9, 7,   # blocks, column-major
                   1, 2, 3, 4,                            # sub
                   2, 3, 4, 5], dtype=np.float64),        # sup
    sub_after=[1 / 5., 2 / 5., 3 / 8., 4 / 2.],           # exact ==, :139-142
    sup_after=[2, 3, 4, 5],                               # exact ==, :145-148
    piv0=[[0, 1], [0, 1], [0, 1]],
    lu_merged=np.array([
        [5, 3, 2, 0, 0, 0],
        [1, 5, 0, 3, 0, 0],
        [1 / 5, 0, 8, 4, 4, 0],
        [0, 2 / 5, 1 / 2, 2, 0, 5],
        [0, 0, 3 / 8, 0, 6, 9],
        [0, 0, 0, 4 / 2, 1 / 3, 4]], dtype=np.float64),
    b=np.array([65, 202, 11, 65, 60, 121], dtype=np.float64),
    x=np.array([-31.47775, 53.42125, 31.0625, -43.36875, -19.25625, 19.5875]),
    x_tol=1e-14,                                          # :157-162
    matvec_x=np.array([3, 2, 6, 2, 5, 4], dtype=np.float64),
    matvec_b=np.array([5 * 3 + 3 * 2 + 2 * 6, 5 * 3 + 8 * 2 + 3 * 2, 8 * 6 + 4 * 2 + 4 * 5 + 1 * 3,
                       4 * 6 + 4 * 2 + 5 * 4 + 2 * 2, 6 * 5 + 9 * 4 + 3 * 6, 2 * 5 + 7 * 4 + 4 * 2],
                      dtype=np.float64),
    matvec_tol=1e-15,                                     # :409-414
    # tests/test_block_diag.cpp:539-560 (same matrix through the raw-pointer view)
    matvec2_x=np.array([-7, 13, 9, -4, -0.7, 42]),
    matvec2_b=np.array([22, 57, 46.2, 256, 400.8, 276.6]),
    matvec2_tol=1e-13,  # the reference asserts 1e-15 ABSOLUTE on values ~4e2 (i.e. bit-level); kept relative here
)

# ---- m3: python_prototype/tests/test_fakelu.py:66-88, 134-140, 181-215 -----------------------
_a = 63 + 17 / 37 * 13
_b = 72 + 11 / 42 * 24
_c = -57 + 14 / 72 * 13
M3 = dict(
    nblocks=2, blockw=2, ndiag=1,
    data=np.array([-17, 37, 63, 13, 11, -42, 72, 24, .1, .2, .2, .3], dtype=np.float64),
    piv0=[[1, 1], [1, 1]],
    rowbycol=[[1, 0], [1, 0]], colbyrow=[[1, 0], [1, 0]],
    sub_after=[0.1 / 37, .2 / _a],
    sup_after=[.2, .3],
    lu_merged=np.array([
        [37, 13, 0, .3],
        [-17 / 37, _a, .2, 0],
        [0, .2 / _a, -42, 24],
        [0.1 / 37, 0, -11 / 42, _b]], dtype=np.float64),
    b=np.array([-62, 207, 11, -14], dtype=np.float64),
)


def m3_xref():
    """Closed form worked in test_fakelu.py:196-207."""
    b = M3["b"]
    y = [b[1]]
    y = y + [b[0] + 17 / 37 * y[0]]
    y = y + [b[3] - .2 / _a * y[1]]
    y = y + [b[2] - 0.1 / 37 * y[0] + 11 / 42 * y[2]]
    x = [(y[3]) / _b]
    x = [(y[2] - 24 * x[-1]) / (-42)] + x
    x = [(y[1] - 0.2 * x[-2]) / _a] + x
    x = [(y[0] - 0.3 * x[-1] - 13 * x[-3]) / 37] + x
    return np.array(x)


# ---- m4: tests/test_block_diag.cpp:166-220; test_fakelu.py:91-118, 218-235 -------------------
M4 = dict(
    nblocks=3, blockw=2, ndiag=2,
    data=np.array([-17, 37, 63, 13, 11, -42, 72, 24, 72, 14, -13, -57,
                   .1, .2, -.1, .08, .03, -.1,
                   .2, .3, -.1, .2, .02, .03], dtype=np.float64),
    piv0=[[1, 1], [1, 1], [0, 1]],
    rowbycol=[[1, 0], [1, 0], [0, 1]], colbyrow=[[1, 0], [1, 0], [0, 1]],
    # sub(0,0,0), sub(0,0,1), sub(0,1,0), sub(0,1,1), sub(1,0,0), sub(1,0,1)   (:185-191, 1e-15)
    sub_after=[.1 / 37, .2 / (63 + 17 / 37. * 13), .1 / 42, .08 / (72 + 11 / 42. * 24), .03 / 37,
               -.1 / (63 + 17 / 37. * 13)],
    sub_tol=1e-15,
    sup_after=[0.2, 0.3, -.1, 0.2, 0.02, 0.03],           # exact ==, :194-199
    lu_merged=np.array([
        [37, 13, 0, .3, 0, .03],
        [-17 / 37, _a, .2, 0, .02, 0],
        [0, .2 / _a, -42, 24, 0, .2],
        [.1 / 37, 0, -11 / 42, _b, -.1, 0],
        [.03 / 37, 0, .1 / 42, 0, 72, -13],
        [0, -.1 / _a, 0, .08 / _b, 14 / 72, _c]], dtype=np.float64),
    b=np.array([-62, 207, 11, -14, 25, -167], dtype=np.float64),
    x=np.array([5.42317616680147374e+00, 4.78588898186963929e-01, 4.00565700557765081e-01,
                8.73749367816923223e-02, 9.14791409109598774e-01, 3.15378902934640371e+00]),
    x_tol=1e-15,                                          # :211-217
)

# ---- blockw = 1: tests/test_block_diag.cpp:714-826 ---------------------------------------------
W1 = dict(
    nblocks=6, blockw=1, ndiag=2,
    data=np.array([5, 8, 8, 4, 6, 7,
                   5, 0, 4, 0, 2, 1, 2, 3, 4,
                   3, 0, 4, 0, 9, 2, 3, 4, 5], dtype=np.float64),
    # sub(1, b, 0) for b = 0..3  (:765-769, 1e-15)
    sub1_after=[1 / 5., 2 / 8., 3 / 8., 4 / 4.],
    sup1_after=[2, 3, 4, 5],
    b=np.array([65, 202, 11, 65, 60, 121], dtype=np.float64),
    x=np.array([0.02968749999999858, 18.768973214285715, 4.272321428571431, -4.383928571428573,
                -4.660714285714287, 9.857142857142858]),
    x_tol=5e-15,
)

# ---- dot_vec2: tests/test_block_diag.cpp:494-537 (w=4, N=3, d=2, formula-filled) ---------------


def dotvec2_case():
    nblocks, blockw, ndiag = 3, 4, 2
    A = np.zeros((12, 12))
    for bi in range(nblocks):
        for ci in range(blockw):
            for ri in range(blockw):
                A[bi * blockw + ri, bi * blockw + ci] = 1.2 * bi + 4.1 * ci - 2.7 * ri
    for di in range(ndiag):
        for bi in range(nblocks - di - 1):
            for ci in range(blockw):
                A[(bi + di + 1) * blockw + ci, bi * blockw + ci] = 1.5 * bi + 0.6 * ci - 0.1 * di
                A[bi * blockw + ci, (bi + di + 1) * blockw + ci] = 3.7 * bi - 0.3 * ci + 0.2 * di
    x = np.array([1.5 * nblocks * blockw - 2.7 * i for i in range(12)])
    return dict(nblocks=nblocks, blockw=blockw, ndiag=ndiag, dense=A, x=x, b=A @ x, tol=5e-14)


# ---- satellites matvec: tests/test_block_diag.cpp:41-108 ----------------------------------------
SAT = dict(
    nblocks=3, blockw=2, ndiag=0, nsat=2,
    dense=np.array([[7, 2, 3, 0, 4, 0],
                    [6, 6, 0, 7, 0, 8],
                    [1, 0, 1, 3, 9, 0],
                    [0, 6, 5, 4, 0, 1],
                    [8, 0, 2, 0, 9, 7],
                    [0, 3, 0, 5, 2, 3]], dtype=np.float64),
    x=np.array([-7, 4, 2, 1, 5, 9], dtype=np.float64),
    b=np.array([-15, 61, 43, 47, 56, 54], dtype=np.float64),
    tol=1e-14,
)
