"""The Python mirror of the reference's Cython wrapper (block_diag_ilu_b200.Compressed*, ILU).

These restate the reference's own Python tests (block_diag_ilu/tests/test_block_diag_ilu.py) against the drop-in:
container behaviour runs on the CPU; everything that computes (dot_vec, scale_diag_add, ILU, solve) goes through
the C ABI to the GPU and is marked gpu.  Without a device the compute calls must fail loudly (no CPU fallback).
"""
import os

import numpy as np
import pytest

import block_diag_ilu_b200 as bdi
from block_diag_ilu_b200 import Compressed, Compressed_from_data, Compressed_from_dense, ILU, LU, _lib

gpu = pytest.mark.gpu


def A6():
    # test_block_diag_ilu.py:20-28
    return np.array([
        [9, 3, .1, 0, .01, 0],
        [-4, 18, 0, .5, 0, .02],
        [.2, 0, 7, -2, 1, 0],
        [0, -.1, 3, 5, 0, -.1],
        [.03, 0, .1, 0, -9, 2],
        [0, .04, 0, .05, -3, 7]])


def A6_data():
    # test_block_diag_ilu.py:31-35: blocks (column major), sub, sup; ld = 2
    return np.array([9, -4, 3, 18, 7, 3, -2, 5, -9, -3, 2, 7,
                     .2, -.1, .1, .05, .03, .04,
                     .1, .5, 1, -.1, .01, .02])


# ---- container (CPU) -------------------------------------------------------------------------------------

def test_get_include():
    inc = bdi.get_include()
    assert inc.endswith("include")
    assert "bdilu.h" in os.listdir(inc)
    assert "block_diag_ilu_b200.hpp" in os.listdir(inc)


def test_public_names():
    for name in ("Compressed", "Compressed_from_dense", "Compressed_from_data", "ILU", "LU", "diag_data_len",
                 "alloc_compressed", "get_compressed", "get_include", "__version__"):
        assert hasattr(bdi, name)


def test_to_dense_roundtrip():
    A = A6()
    assert np.array_equal(Compressed_from_dense(A, 3, 2, 2).to_dense(), A)


def test_from_data_indexing():
    A = A6()
    c = Compressed_from_data(A6_data(), 3, 2, 2, 0, 2)
    for idx, val in np.ndenumerate(A):
        assert c[idx] == val          # structural zeros read 0.0 (pyx:79-84)
    assert np.array_equal(c.data, A6_data())
    c2 = c.copy()
    c2.set_block(0, 0, 0, 42.0)
    assert c.get_block(0, 0, 0) == 9 and c2.get_block(0, 0, 0) == 42.0


def test_from_dense_satellites():
    # test_block_diag_ilu.py:74-88
    c = Compressed_from_dense(A6(), 3, 2, 0, 2)
    assert c.nsat == 2
    want_top = {(0, 0, 0): .01, (0, 0, 1): .02, (1, 0, 0): .1, (1, 0, 1): .5, (1, 1, 0): 1, (1, 1, 1): -.1}
    want_bot = {(0, 0, 0): .03, (0, 0, 1): .04, (1, 0, 0): .2, (1, 0, 1): -.1, (1, 1, 0): .1, (1, 1, 1): .05}
    for k, v in want_top.items():
        assert c.get_top(*k) == v
    for k, v in want_bot.items():
        assert c.get_bot(*k) == v
    assert np.array_equal(c.to_dense(), A6())


def test_default_ld_is_padded():
    c = Compressed(3, 2, 1)
    assert c.ld == 8 and c.data.size == (3 * 2 + 2 * 2) * 8   # block_diag_ilu.hpp:98, anyode_matrix.hpp:18-21


def test_constructor_errors():
    with pytest.raises(ValueError, match="A not square"):
        Compressed_from_dense(np.zeros((4, 6)), 2, 2, 1)
    with pytest.raises(ValueError, match="does not match"):
        Compressed_from_dense(np.zeros((6, 6)), 2, 2, 1)
    with pytest.raises(ValueError, match="too many diagonals"):
        Compressed_from_dense(np.zeros((6, 6)), 3, 2, 3)
    with pytest.raises(ValueError, match="Incompatible sizes"):
        Compressed_from_data(np.zeros(5), 3, 2, 2, 0, 2)


def test_lu_is_out_of_scope():
    with pytest.raises(NotImplementedError):
        LU(Compressed_from_dense(A6(), 3, 2, 2))


def test_compute_fails_loudly_without_device():
    if _lib.lib().bdilu_device_count() > 0:
        pytest.skip("a CUDA device is visible")
    c = Compressed_from_dense(A6(), 3, 2, 2)
    with pytest.raises(_lib.BdiluError):
        c.dot_vec(np.ones(6))
    with pytest.raises(_lib.BdiluError):
        ILU(c)
    with pytest.raises(_lib.BdiluError):
        c.scale_diag_add(c, 2.0, 1.0)


# ---- compute (GPU, through the C ABI) -----------------------------------------------------------------------

V6 = np.array([3, -2, 1, 5, 7, -6], dtype=np.float64)
B6 = np.array([-7, 3, 5, -4, 6, 1.5])


@gpu
@pytest.mark.parametrize("ndiag,nsat", [(2, 0), (0, 2)])
def test_dot_vec(ndiag, nsat):
    # test_block_diag_ilu.py:115-130
    c = Compressed_from_dense(A6(), 3, 2, ndiag, nsat)
    assert np.allclose(c.dot_vec(V6), A6().dot(V6), rtol=1e-14, atol=0)
    assert np.allclose(Compressed_from_data(A6_data(), 3, 2, 2, 0, 2).dot_vec(V6), A6().dot(V6), rtol=1e-14)


@gpu
@pytest.mark.parametrize("sat", [True, False])
def test_scale_diag_add(sat):
    # test_block_diag_ilu.py:91-102
    ndiag, nsat = (0, 2) if sat else (2, 0)
    A = A6()
    c1 = Compressed_from_dense(A, 3, 2, ndiag, nsat)
    c2 = Compressed_from_dense(A, 3, 2, ndiag, nsat)
    c2.scale_diag_add(c1, 5, 0)
    assert np.allclose(c2.dot_vec(V6), 5 * A.dot(V6))
    c2.scale_diag_add(c1, -0.25, 1)        # set_to_eye_plus_scaled_mtx(-gamma, J), block_diag_ilu.hpp:288-290
    assert np.allclose(c2.to_dense(), np.eye(6) - 0.25 * A, rtol=1e-15, atol=0)


@gpu
@pytest.mark.parametrize("ndiag,nsat", [(2, 0), (0, 2)])
def test_ilu_close_to_exact_on_dominant(ndiag, nsat):
    # test_block_diag_ilu.py:44-71, 105-112 (the exact solve comes from numpy here; LU is out of scope)
    c = Compressed_from_dense(A6(), 3, 2, ndiag, nsat)
    ix = ILU(c).solve(B6)
    assert np.allclose(np.linalg.solve(A6(), B6), ix, rtol=0.05, atol=1e-6)
    assert np.array_equal(c.to_dense(), A6())          # ILU copies its source (pyx:145)


@gpu
def test_ilu_solve_m1():
    # test_block_diag_ilu.py:215-221; pivots of test_fakelu.py:121-125 (0-based [1,1,2])
    A = np.array([[1, 3, 5], [2, 4, 7], [1, 1, 0]], dtype=np.float64)
    ilu = ILU(Compressed_from_dense(A, 1, 3, 0))
    assert ilu.ipiv.tolist() == [2, 2, 3]
    assert np.allclose(ilu.pbdm.to_dense(), [[2, 4, 7], [.5, 1, 1.5], [.5, -1, -2]], rtol=1e-15)
    assert np.allclose(ilu.solve(np.array([36., 51., 5.])), [2, 3, 5], rtol=1e-14)


def _m2():
    return np.array([[5, 3, 2, 0, 0, 0], [5, 8, 0, 3, 0, 0], [1, 0, 8, 4, 4, 0],
                     [0, 2, 4, 4, 0, 5], [0, 0, 3, 0, 6, 9], [0, 0, 0, 4, 2, 7]], dtype=np.float64)


@gpu
@pytest.mark.parametrize("sat", [True, False])
def test_ilu_solve_m2(sat):
    # test_block_diag_ilu.py:224-252; solution pinned by tests/test_block_diag.cpp:150-163
    A = _m2()
    ilu = ILU(Compressed_from_dense(A, 3, 2, 0, 2)) if sat else ILU(Compressed_from_dense(A, 3, 2, 1))
    b = np.array([65, 202, 11, 65, 60, 121], dtype=np.float64)
    x = ilu.solve(b)
    assert np.allclose(x, [-31.47775, 53.42125, 31.0625, -43.36875, -19.25625, 19.5875], rtol=1e-14)
    assert not np.allclose(x, [11, 12, -13, 17, 9, 5])       # the ILU error is intentional
    if not sat:
        # scaled sub-diagonals are exact quotients (tests/test_block_diag.cpp:138-143)
        assert [ilu.pbdm.get_sub(0, 0, 0), ilu.pbdm.get_sub(0, 0, 1), ilu.pbdm.get_sub(0, 1, 0),
                ilu.pbdm.get_sub(0, 1, 1)] == [1 / 5, 2 / 5, 3 / 8, 4 / 2]


def _m4():
    return np.array([[-17, 63, .2, 0, .02, 0], [37, 13, 0, .3, 0, .03], [.1, 0, 11, 72, -.1, 0],
                     [0, .2, -42, 24, 0, .2], [.03, 0, -.1, 0, 72, -13], [0, -.1, 0, .08, 14, -57]], dtype=np.float64)


@gpu
@pytest.mark.parametrize("nsat", [0, 1, 2])
def test_ilu_solve_m4(nsat):
    # test_block_diag_ilu.py:281-310; solution pinned by tests/test_block_diag.cpp:201-219
    A = _m4()
    ilu = ILU(Compressed_from_dense(A, 3, 2, 2 - nsat, nsat))
    assert ilu.ipiv.tolist() == [2, 2, 2, 2, 1, 2]
    assert ilu.rowbycol.tolist() == [1, 0, 1, 0, 0, 1] and ilu.colbyrow.tolist() == [1, 0, 1, 0, 0, 1]
    b = np.array([-62, 207, 11, -14, 25, -167], dtype=np.float64)
    x = ilu.solve(b)
    xref = [5.42317616680147374, 0.478588898186963929, 0.400565700557765081,
            0.0873749367816923223, 0.914791409109598774, 3.15378902934640371]
    assert np.allclose(x, xref, rtol=1e-14, atol=0)
    assert np.allclose(np.linalg.solve(A, b), x, rtol=0.03, atol=1e-6)


@gpu
def test_ilu_solve_m3_sat_and_multi_rhs():
    # test_block_diag_ilu.py:255-278 (m3) + the [nrhs, n] extension
    A = np.array([[-17, 63, .2, 0], [37, 13, 0, .3], [.1, 0, 11, 72], [0, .2, -42, 24]], dtype=np.float64)
    b = np.array([-62, 207, 11, -14], dtype=np.float64)
    xs = []
    for nd, ns in [(1, 0), (0, 1)]:
        ilu = ILU(Compressed_from_dense(A, 2, 2, nd, ns))
        x = ilu.solve(b.copy())
        assert np.allclose(np.linalg.solve(A, b), x, rtol=0.01, atol=1e-6)
        xs.append(x)
        X = ilu.solve(np.stack([b, 2 * b, -b]))
        assert X.shape == (3, 4) and np.array_equal(X[0], x) and np.allclose(X[1], 2 * x, rtol=1e-15)
    assert np.allclose(xs[0], xs[1], rtol=1e-14)


@gpu
def test_solve_flag_exceptions():
    # pyx:132-137: NaN in b -> ValueError; zero U diagonal -> ZeroDivisionError; singular block -> "ILU failed!"
    ilu = ILU(Compressed_from_dense(_m2(), 3, 2, 1))
    b = np.array([65, 202, np.nan, 65, 60, 121])
    with pytest.raises(ValueError, match="NaN in b"):
        ilu.solve(b)
    S = _m2()
    S[2:4, 2:4] = [[1, 2], [2, 4]]
    with pytest.raises(RuntimeError, match="ILU failed!"):
        ILU(Compressed_from_dense(S, 3, 2, 1))


@gpu
def test_from_tensor_device_resident():
    torch = pytest.importorskip("torch")
    from block_diag_ilu_b200 import synth
    N, w, d = 16, 8, 2
    data, b = synth.make_batch(3, N, w, d, "D", seed0=11)
    ilu = ILU.from_tensor(torch.from_numpy(data).cuda(), N, w, d)
    x = ilu.solve(torch.from_numpy(b).cuda())
    assert x.is_cuda and tuple(x.shape) == b.shape
    for s in range(3):
        host = ILU(Compressed_from_data(data[s], N, w, d, 0, w))
        assert np.array_equal(host.solve(b[s, 0]), x[s, 0].cpu().numpy())
    with pytest.raises(TypeError):
        ilu.solve(b[0, 0])
