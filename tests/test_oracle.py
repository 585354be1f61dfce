"""Pins the parity oracle (CPU only).

(1) the C restatement (oracle/bdilu_oracle.c, "port") against every known-answer vector the
    reference's own tests hold for the path (tests/golden/reference_kats.py);
(2) the port against outputs of the UNMODIFIED reference header run in the build container
    (tests/golden/ref_vectors.npz, made by tests/golden/make_ref_vectors.py);
(3) when oracle/_ref is loadable, the reference shim itself against (1) -- which checks the shim, the
    LAPACK renaming and the transcription in one go.
"""
import hashlib
import os

import numpy as np
import pytest

from block_diag_ilu_b200 import datastruct as ds
from block_diag_ilu_b200 import synth
from tests.golden import reference_kats as K

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_vectors.npz")


@pytest.fixture(params=["port", "reference"])
def impl(request, port):
    if request.param == "port":
        return port
    from oracle.loader import Ref
    try:
        return Ref()
    except (OSError, FileNotFoundError) as e:
        pytest.skip("reference shim unavailable: %s" % e)


def merged_from_packed(data, N, w, d, ld=None):
    return ds.unpack_dense(data, N, w, d, 0, ld or w)


def diag_blocks(A, N, w):
    """The prototype's LU_merged shows couplings at their ROW-PERMUTED places (test_fakelu.py:66-118); the
    header keeps them by column (block_diag_ilu.hpp:471-475), so only the diagonal blocks are comparable."""
    return np.stack([A[i * w:(i + 1) * w, i * w:(i + 1) * w] for i in range(N)])


def factor_solve(impl, case, ld=None):
    N, w, d = case["nblocks"], case["blockw"], case["ndiag"]
    ld = ld or w
    data = case["data"].copy()
    if ld != w:  # re-pack with a padded leading dimension
        data = ds.pack_dense(ds.unpack_dense(data, N, w, d, 0, w), N, w, d, 0, ld)
    ipiv, rbc, cbr, info = impl.factor(data, N, w, d, 0, ld)
    assert info == 0
    x = None
    if "b" in case:
        x, flag = impl.solve(data, N, w, d, 0, ld, rbc, cbr, case["b"])
        assert flag == 0
    return data, ipiv, rbc, cbr, x


def test_permutation_helpers(impl):
    rbc = impl.rowpiv2rowbycol(np.array(K.PIV_1BASED, np.int32))
    assert rbc.tolist() == K.ROWBYCOL
    assert impl.rowbycol2colbyrow(rbc).tolist() == K.COLBYROW


def test_m1_pivots_and_factors(impl):
    A = K.M1["dense"]
    data = ds.pack_dense(A, 1, 3, 0, 0, 3)
    ipiv, rbc, cbr, info = impl.factor(data, 1, 3, 0, 0, 3)
    assert info == 0
    assert (ipiv - 1).tolist() == K.M1["piv0"]
    assert np.allclose(ds.unpack_dense(data, 1, 3, 0, 0, 3), K.M1["lu_merged"], rtol=0, atol=1e-15)


@pytest.mark.parametrize("ld", [None, 5])
def test_m2(impl, ld):
    c = K.M2
    data, ipiv, rbc, cbr, x = factor_solve(impl, c, ld)
    N, w, d = c["nblocks"], c["blockw"], c["ndiag"]
    L = ld or w
    o_sub, o_sup, _, _ = ds.offsets(N, w, d, 0, L)
    sub = data[o_sub:o_sup].reshape(-1, L)[:, :w].ravel()
    sup = data[o_sup:o_sup + (o_sup - o_sub)].reshape(-1, L)[:, :w].ravel()
    assert sub.tolist() == c["sub_after"]          # exact, as the reference asserts with ==
    assert sup.tolist() == c["sup_after"]
    assert (ipiv.reshape(N, w) - 1).tolist() == c["piv0"]
    assert np.allclose(merged_from_packed(data, N, w, d, L), c["lu_merged"], rtol=0, atol=1e-15)
    assert np.abs(x - c["x"]).max() < c["x_tol"]


def test_m2_matvec(impl):
    c = K.M2
    N, w, d = c["nblocks"], c["blockw"], c["ndiag"]
    out = impl.dot_vec(c["data"].copy(), N, w, d, 0, w, c["matvec_x"])
    assert np.abs(out - c["matvec_b"]).max() < c["matvec_tol"]
    out = impl.dot_vec(c["data"].copy(), N, w, d, 0, w, c["matvec2_x"])
    assert np.abs(out - c["matvec2_b"]).max() <= c["matvec2_tol"] * np.abs(c["matvec2_b"]).max()


def test_m3(impl):
    c = K.M3
    data, ipiv, rbc, cbr, x = factor_solve(impl, c)
    N, w, d = c["nblocks"], c["blockw"], c["ndiag"]
    assert (ipiv.reshape(N, w) - 1).tolist() == c["piv0"]
    assert rbc.reshape(N, w).tolist() == c["rowbycol"]
    assert cbr.reshape(N, w).tolist() == c["colbyrow"]
    o_sub, o_sup, _, _ = ds.offsets(N, w, d, 0, w)
    assert np.allclose(data[o_sub:o_sup], c["sub_after"], rtol=0, atol=1e-15)
    assert data[o_sup:].tolist() == c["sup_after"]
    assert np.allclose(diag_blocks(merged_from_packed(data, N, w, d), N, w), diag_blocks(c["lu_merged"], N, w),
                       rtol=4e-16, atol=1e-15)  # one ulp: OpenBLAS scales by the reciprocal of the pivot
    assert np.allclose(x, K.m3_xref(), rtol=1e-14, atol=0)


@pytest.mark.parametrize("ld", [None, 8])
def test_m4(impl, ld):
    c = K.M4
    data, ipiv, rbc, cbr, x = factor_solve(impl, c, ld)
    N, w, d = c["nblocks"], c["blockw"], c["ndiag"]
    L = ld or w
    assert (ipiv.reshape(N, w) - 1).tolist() == c["piv0"]
    assert rbc.reshape(N, w).tolist() == c["rowbycol"]
    assert cbr.reshape(N, w).tolist() == c["colbyrow"]
    o_sub, o_sup, _, _ = ds.offsets(N, w, d, 0, L)
    sub = data[o_sub:o_sup].reshape(-1, L)[:, :w].ravel()
    sup = data[o_sup:o_sup + (o_sup - o_sub)].reshape(-1, L)[:, :w].ravel()
    assert np.abs(sub - np.array(c["sub_after"])).max() < c["sub_tol"]
    assert sup.tolist() == c["sup_after"]
    assert np.allclose(diag_blocks(merged_from_packed(data, N, w, d, L), N, w), diag_blocks(c["lu_merged"], N, w),
                       rtol=4e-16, atol=1e-15)  # one ulp: OpenBLAS scales by the reciprocal of the pivot
    assert np.abs(x - c["x"]).max() < c["x_tol"]


def test_blockw1(impl):
    c = K.W1
    data, ipiv, rbc, cbr, x = factor_solve(impl, c)
    N, d = c["nblocks"], c["ndiag"]
    o_sub, o_sup, _, _ = ds.offsets(N, 1, d, 0, 1)
    k = ds.diag_chunk(N, 1, 0)
    assert np.abs(data[o_sub + k:o_sub + k + 4] - np.array(c["sub1_after"])).max() < 1e-15
    assert data[o_sup + k:o_sup + k + 4].tolist() == c["sup1_after"]
    assert np.abs(x - c["x"]).max() < c["x_tol"]


def test_dotvec2_and_satellites(impl):
    c = K.dotvec2_case()
    data = ds.pack_dense(c["dense"], c["nblocks"], c["blockw"], c["ndiag"], 0, 4)
    out = impl.dot_vec(data, c["nblocks"], c["blockw"], c["ndiag"], 0, 4, c["x"])
    assert np.abs(out - c["b"]).max() < c["tol"] * max(1.0, np.abs(c["b"]).max())
    s = K.SAT
    data = ds.pack_dense(s["dense"], s["nblocks"], s["blockw"], s["ndiag"], s["nsat"], 8)
    out = impl.dot_vec(data, s["nblocks"], s["blockw"], s["ndiag"], s["nsat"], 8, s["x"])
    assert np.abs(out - s["b"]).max() < s["tol"]


def test_scale_diag_add(impl):
    # tests/test_block_diag.cpp:562-614: mv.scale_diag_add(src, 2, 1) == 2*src + I on the blocks
    c = K.M2
    N, w, d = c["nblocks"], c["blockw"], c["ndiag"]
    out = impl.scale_diag_add(c["data"].copy(), N, w, d, 0, w, 2.0, 1.0)
    want = 2 * ds.unpack_dense(c["data"], N, w, d, 0, w) + np.eye(N * w)
    assert np.array_equal(ds.unpack_dense(out, N, w, d, 0, w), want)


def test_solve_flags(port):
    # block_diag_ilu.hpp:500-501 (1-based index of the first NaN), :533-534 (N*w + index of a zero U diagonal)
    c = K.M2
    N, w, d = c["nblocks"], c["blockw"], c["ndiag"]
    data = c["data"].copy()
    ipiv, rbc, cbr, info = port.factor(data, N, w, d, 0, w)
    b = c["b"].copy()
    b[3] = np.nan
    _, flag = port.solve(data, N, w, d, 0, w, rbc, cbr, b)
    assert flag == 4
    data[(1 * w + 1) * w + 1] = 0.0  # U[1,1] of block 1
    _, flag = port.solve(data, N, w, d, 0, w, rbc, cbr, c["b"])
    assert flag == N * w + 1 * w + 1


# ---- port vs the reference's recorded outputs ------------------------------------------------------

def golden_cases():
    z = np.load(GOLDEN)
    return sorted({k.split("/")[0] for k in z.files})


@pytest.mark.parametrize("name", golden_cases())
def test_port_matches_reference_outputs(port, name):
    z = np.load(GOLDEN)
    N, w, d, ld, nsat, seed = (int(v) for v in z[name + "/shape"])
    fam = str(z[name + "/family"])
    if name + "/data" in z.files:
        data, b = z[name + "/data"].copy(), z[name + "/b"].copy()
    else:
        data, b = synth.make_system(N, w, d, fam, seed, ld=ld, nsat=nsat)
        if hashlib.sha256(data.tobytes() + b.tobytes()).hexdigest() != str(z[name + "/input_sha256"]):
            pytest.skip("numpy's generator stream differs from the one the fixture was made with")
    av = port.dot_vec(data.copy(), N, w, d, nsat, ld, b)
    assert np.abs(av - z[name + "/dot_vec"]).max() <= 1e-13 * np.abs(z[name + "/dot_vec"]).max()
    ipiv, rbc, cbr, info = port.factor(data, N, w, d, nsat, ld)
    assert info == 0
    assert np.array_equal(ipiv, z[name + "/ipiv"])          # LAPACK's pivots, bit-exact
    assert np.array_equal(rbc, z[name + "/rowbycol"])
    assert np.array_equal(cbr, z[name + "/colbyrow"])
    want = z[name + "/factored"]
    nb = N * w * ld
    g, r = data[:nb].reshape(N, -1), want[:nb].reshape(N, -1)
    tol = 1e-12 if fam == "D" else 1e-10  # non-dominant blocks: two fp64 eliminations differ by cond(block)*eps
    assert (np.abs(g - r).max(axis=1) / np.abs(r).max(axis=1)).max() < tol
    if want.size > nb:
        assert np.abs(data[nb:] - want[nb:]).max() <= tol * np.abs(want[nb:]).max()
    x, flag = port.solve(data, N, w, d, nsat, ld, rbc, cbr, b)
    assert flag == 0
    assert np.abs(x - z[name + "/x"]).max() <= tol * np.abs(z[name + "/x"]).max()


# ---- Real_t = float: the float build of the reference header and its recorded outputs -----------------------------------
GOLDEN_F32 = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_vectors_f32.npz")


def _golden_f32_names():
    z = np.load(GOLDEN_F32)
    return sorted({k.split("/")[0] for k in z.files})


def test_float_default_ld_and_ndata_match_the_reference():
    """bdilu_default_ld_f32 / bdilu_ndata against BlockDiagMatrix<float>'s own arithmetic (hpp:98-102)."""
    from block_diag_ilu_b200 import _lib
    from oracle.loader import RefF32
    try:
        r = RefF32()
    except (OSError, FileNotFoundError, AttributeError) as e:
        pytest.skip("float reference shim unavailable: %s" % e)
    L = _lib.lib()
    for w in (1, 2, 3, 5, 8, 15, 16, 17, 31, 32, 33, 48, 64):
        assert L.bdilu_default_ld_f32(w) == r.default_ld(w)
        for (N, d, nsat) in ((1, 0, 0), (3, 1, 0), (7, 2, 2), (9, 3, 1)):
            assert L.bdilu_ndata(N, w, d, nsat, L.bdilu_default_ld_f32(w)) == r.ndata(N, w, d, nsat, 0)


@pytest.mark.parametrize("name", _golden_f32_names())
def test_float_reference_reproduces_its_recorded_outputs(name):
    """The committed float fixtures are what the reference computes here (bit for bit): they pin the GPU float tests."""
    from oracle.loader import RefF32
    try:
        r = RefF32()
    except (OSError, FileNotFoundError, AttributeError) as e:
        pytest.skip("float reference shim unavailable: %s" % e)
    z = np.load(GOLDEN_F32)
    N, w, d, ld, nsat, seed = (int(v) for v in z[name + "/shape"])
    data, b = synth.make_system(N, w, d, str(z[name + "/family"]), seed, ld=ld, nsat=nsat)
    assert np.array_equal(data.astype(np.float32), z[name + "/data"]) and np.array_equal(b.astype(np.float32), z[name + "/b"])
    fac = z[name + "/data"].copy()
    assert np.array_equal(r.dot_vec(fac.copy(), N, w, d, nsat, ld, z[name + "/b"]), z[name + "/dot_vec"])
    ipiv, rbc, cbr, info = r.factor(fac, N, w, d, nsat, ld)
    assert info == 0 and np.array_equal(ipiv, z[name + "/ipiv"]) and np.array_equal(rbc, z[name + "/rowbycol"])
    assert np.array_equal(cbr, z[name + "/colbyrow"]) and np.array_equal(fac, z[name + "/factored"])
    x, flag = r.solve(fac, N, w, d, nsat, ld, rbc, cbr, z[name + "/b"])
    assert flag == 0 and np.array_equal(x, z[name + "/x"])
    r.release()
