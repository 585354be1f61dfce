"""Real_t = float: the *_f32 entry points of the C ABI against the reference header instantiated with
float (oracle/_ref, LAPACK sgetrf) and against its recorded outputs (tests/golden/ref_vectors_f32.npz).

Bars: ipiv / rowbycol / colbyrow bit-exact; factors within F32_TOL block-max-norm-relative; solutions within F32_TOL
in ||dx||_inf/||x||_inf; flags / info equal.  F32_TOL = 2e-5 (~300 float ulps; measured discrepancies are <= 4e-6: the
two eliminations round differently -- OpenBLAS' blocked sgetrf against the right-looking elimination here -- and a chain
of blocks carries that through the solve).  Inputs are families D and R: their pivots are separated by orders of
magnitude, so float rounding cannot flip one (family P's near-ties can; the double path covers P).
"""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from block_diag_ilu_b200 import _lib  # noqa: E402
from block_diag_ilu_b200 import device as dev  # noqa: E402
from block_diag_ilu_b200 import synth  # noqa: E402

import os  # noqa: E402

F32_TOL = 2e-5
GOLDEN_F32 = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_vectors_f32.npz")


@pytest.fixture(scope="module")
def ref32():
    from oracle.loader import RefF32
    try:
        r = RefF32()
    except (OSError, FileNotFoundError, AttributeError) as e:
        pytest.fail("oracle/_ref (float build of the reference header) did not load on the GPU box: %s" % e)
    yield r
    r.release()


def blocks_rel_err(got, want, N, w, ld):
    nb = N * w * ld
    g = got[:nb].reshape(N, w, ld)[:, :, :w].reshape(N, -1).astype(np.float64)
    r = want[:nb].reshape(N, w, ld)[:, :, :w].reshape(N, -1).astype(np.float64)
    e = (np.abs(g - r).max(axis=1) / np.abs(r).max(axis=1)).max()
    rest_g = got[nb:].reshape(-1, ld)[:, :w].astype(np.float64)
    rest_r = want[nb:].reshape(-1, ld)[:, :w].astype(np.float64)
    if rest_r.size:
        e = max(e, np.abs(rest_g - rest_r).max() / max(np.abs(rest_r).max(), 1e-30))
    return e


def make32(nsys, N, w, d, fam, seed0, ld, nrhs, nsat):
    data, b = synth.make_batch(nsys, N, w, d, fam, seed0=seed0, ld=ld, nrhs=nrhs, nsat=nsat)
    return data.astype(np.float32), b.astype(np.float32)


@pytest.mark.parametrize("fam", ["D", "R"])
@pytest.mark.parametrize("N,w,d,ld,nsat", [
    (64, 4, 1, 16, 0),      # BASELINE config 1 with the float default ld (64-byte rows)
    (64, 4, 1, 4, 0), (6, 1, 2, 1, 0), (5, 2, 1, 2, 0), (7, 3, 2, 3, 0), (9, 5, 3, 8, 0), (33, 8, 2, 8, 0),
    (40, 16, 1, 16, 0), (24, 32, 2, 32, 0), (12, 32, 2, 35, 0), (10, 24, 3, 24, 0), (8, 64, 3, 64, 0),
    (6, 48, 2, 56, 0), (1, 32, 0, 32, 0), (3, 7, 0, 8, 0),
    (9, 8, 2, 8, 2), (7, 32, 1, 32, 3), (5, 40, 1, 40, 1),   # satellites (periodic couplings)
    (1024, 16, 1, 16, 0),   # BASELINE config 2's shape
])
def test_f32_factor_solve_parity(ref32, N, w, d, ld, nsat, fam):
    nsys, nrhs = 3, 2
    data, b = make32(nsys, N, w, d, fam, 500, ld, nrhs, nsat)
    dd, bb = torch.from_numpy(data).cuda(), torch.from_numpy(b).cuda()
    ilu = dev.BatchedILU(dd, N, w, d, nsat=nsat, ld=ld, copy=True)
    x = ilu.solve(bb)
    av = dev.dot_vec(dd, bb, N, w, d, nsat=nsat, ld=ld)
    torch.cuda.synchronize()
    assert ilu.data.dtype == torch.float32 and x.dtype == torch.float32
    got, gx, gav = ilu.data.cpu().numpy(), x.cpu().numpy(), av.cpu().numpy()
    for s in range(nsys):
        for r in range(nrhs):
            want = ref32.dot_vec(data[s].copy(), N, w, d, nsat, ld, b[s, r])
            assert np.abs(gav[s, r] - want).max() <= F32_TOL * np.abs(want).max()
        fac = data[s].copy()
        ipiv, rbc, cbr, info = ref32.factor(fac, N, w, d, nsat, ld)
        assert info == 0 and int(ilu.info[s]) == 0
        assert np.array_equal(ilu.ipiv[s].cpu().numpy(), ipiv), "ipiv mismatch sys %d" % s
        assert np.array_equal(ilu.rowbycol[s].cpu().numpy(), rbc)
        assert np.array_equal(ilu.colbyrow[s].cpu().numpy(), cbr)
        assert blocks_rel_err(got[s], fac, N, w, ld) < F32_TOL
        for r in range(nrhs):
            xr, flag = ref32.solve(fac, N, w, d, nsat, ld, rbc, cbr, b[s, r])
            assert flag == int(ilu.flags[s, r]) == 0
            err = np.abs(gx[s, r].astype(np.float64) - xr).max() / np.abs(xr).max()
            assert err < F32_TOL, "x mismatch sys %d rhs %d: %g" % (s, r, err)


def test_f32_fused_entry_point_and_cfg3_shape(ref32):
    """bdilu_factor_solve_f32 at BASELINE config 3's shape (one system), against the float reference."""
    N, w, d = 8192, 32, 2
    data, b = make32(2, N, w, d, "D", 3_000_000, w, 1, 0)
    dd, bb = torch.from_numpy(data).cuda(), torch.from_numpy(b).cuda()
    x, ilu = dev.factor_solve(dd, bb, N, w, d)
    torch.cuda.synchronize()
    fac = data[1].copy()
    ipiv, rbc, cbr, info = ref32.factor(fac, N, w, d, 0, w)
    assert np.array_equal(ilu.ipiv[1].cpu().numpy(), ipiv)
    assert blocks_rel_err(ilu.data[1].cpu().numpy(), fac, N, w, w) < F32_TOL
    xr, flag = ref32.solve(fac, N, w, d, 0, w, rbc, cbr, b[1, 0])
    assert flag == 0 and int(ilu.flags[1, 0]) == 0
    assert np.abs(x[1, 0].cpu().numpy().astype(np.float64) - xr).max() / np.abs(xr).max() < F32_TOL


def golden_names():
    z = np.load(GOLDEN_F32)
    return sorted({k.split("/")[0] for k in z.files})


@pytest.mark.parametrize("name", golden_names())
def test_f32_reference_recorded_outputs_through_the_c_abi(name):
    """Inputs AND outputs of the float reference, recorded in the build container (tests/golden/make_ref_vectors_f32.py):
    no oracle in the loop."""
    z = np.load(GOLDEN_F32)
    N, w, d, ld, nsat, seed = (int(v) for v in z[name + "/shape"])
    data, b = z[name + "/data"].copy(), z[name + "/b"].copy()
    dd = torch.from_numpy(data[None]).cuda()
    bb = torch.from_numpy(b[None, None]).cuda()
    av = dev.dot_vec(dd, bb, N, w, d, nsat=nsat, ld=ld).cpu().numpy()[0, 0]
    assert np.abs(av - z[name + "/dot_vec"]).max() <= F32_TOL * np.abs(z[name + "/dot_vec"]).max()
    ilu = dev.BatchedILU(dd, N, w, d, nsat=nsat, ld=ld, copy=True)
    x = ilu.solve(bb).cpu().numpy()[0, 0]
    assert np.array_equal(ilu.ipiv[0].cpu().numpy(), z[name + "/ipiv"])
    assert np.array_equal(ilu.rowbycol[0].cpu().numpy(), z[name + "/rowbycol"])
    assert np.array_equal(ilu.colbyrow[0].cpu().numpy(), z[name + "/colbyrow"])
    assert blocks_rel_err(ilu.data[0].cpu().numpy(), z[name + "/factored"], N, w, ld) < F32_TOL
    assert int(ilu.flags[0, 0]) == 0
    xr = z[name + "/x"].astype(np.float64)
    assert np.abs(x - xr).max() <= F32_TOL * np.abs(xr).max()


def test_f32_flags_and_info_encodings(ref32):
    """NaN in b -> 1-based index of the first NaN; zero U diagonal -> n + row; singular block -> info word."""
    N, w, d = 5, 4, 1
    data, b = make32(1, N, w, d, "D", 77, w, 1, 0)
    bn = b.copy()
    bn[0, 0, 6] = np.nan
    bn[0, 0, 13] = np.nan
    ilu = dev.BatchedILU(torch.from_numpy(data).cuda(), N, w, d, copy=True)
    ilu.solve(torch.from_numpy(bn).cuda())
    fac = data[0].copy()
    _, rbc, cbr, _ = ref32.factor(fac, N, w, d, 0, w)
    _, flag = ref32.solve(fac, N, w, d, 0, w, rbc, cbr, bn[0, 0])
    assert int(ilu.flags[0, 0]) == flag == 7
    # a zero on U's diagonal (planted after the factorisation, as the reference's own test does)
    ilu.data[0, (2 * w + 1) * w + 1] = 0.0
    ilu.solve(torch.from_numpy(b).cuda())
    fac[(2 * w + 1) * w + 1] = 0.0
    _, flag = ref32.solve(fac, N, w, d, 0, w, rbc, cbr, b[0, 0])
    assert int(ilu.flags[0, 0]) == flag == N * w + 2 * w + 1
    # singular block 3 (column 2 zero): info = (block << 7) | (dgetrf info = 3); the reference throws
    sing = data.copy()
    blk = sing[0, :N * w * w].reshape(N, w, w)
    blk[3, 2, :] = 0.0
    il2 = dev.BatchedILU(torch.from_numpy(sing).cuda(), N, w, d, copy=True, check=False)
    assert int(il2.info[0]) == (3 << 7) | 3
    assert ref32.factor(sing[0].copy(), N, w, d, 0, w)[3] == -1
    with pytest.raises(RuntimeError, match="ILU failed!"):
        dev.BatchedILU(torch.from_numpy(sing).cuda(), N, w, d, copy=True)


def test_f32_host_entry_points(ref32):
    """bdilu_ilu_create_f32 / bdilu_ilu_solve_f32 / bdilu_host_dot_vec_f32 / bdilu_host_scale_diag_add_f32 on host memory."""
    L = _lib.lib()
    N, w, d, nsat = 9, 8, 2, 2
    ld = L.bdilu_default_ld_f32(w)
    assert ld == 16 == ref32.default_ld(w)
    data, b = make32(1, N, w, d, "R", 31, ld, 3, nsat)
    data, b = np.ascontiguousarray(data[0]), np.ascontiguousarray(b[0])
    out = np.zeros_like(b)
    _lib.check(L.bdilu_host_dot_vec_f32(_lib.ptr(data), N, w, d, nsat, ld, _lib.ptr(b), _lib.ptr(out), 3))
    for r in range(3):
        want = ref32.dot_vec(data.copy(), N, w, d, nsat, ld, b[r])
        assert np.abs(out[r] - want).max() <= F32_TOL * np.abs(want).max()
    sc = np.zeros_like(data)
    _lib.check(L.bdilu_host_scale_diag_add_f32(_lib.ptr(sc), _lib.ptr(data), N, w, d, nsat, ld, -0.25, 1.0))
    assert np.array_equal(sc, ref32.scale_diag_add(data.copy(), N, w, d, nsat, ld, -0.25, 1.0))
    fac = data.copy()
    n = N * w
    ipiv, rbc, cbr = (np.zeros(n, np.int32) for _ in range(3))
    info = np.zeros(1, np.int32)
    h = C.c_void_p()
    _lib.check(L.bdilu_ilu_create_f32(C.byref(h), _lib.ptr(fac), 0, 1, N, w, d, nsat, ld, 1, _lib.ptr(ipiv), _lib.ptr(rbc),
                                      _lib.ptr(cbr), _lib.ptr(info)))
    try:
        want = data.copy()
        ripiv, rrbc, rcbr, rinfo = ref32.factor(want, N, w, d, nsat, ld)
        assert info[0] == 0 and rinfo == 0
        assert np.array_equal(ipiv, ripiv) and np.array_equal(rbc, rrbc) and np.array_equal(cbr, rcbr)
        assert blocks_rel_err(fac, want, N, w, ld) < F32_TOL   # copy_back: the caller's view holds the factors
        x = np.zeros_like(b)
        flags = np.full(3, -1, np.int32)
        assert L.bdilu_ilu_solve_f32(h, _lib.ptr(b), _lib.ptr(x), 3, _lib.ptr(flags)) == 0
        assert not flags.any()
        for r in range(3):
            xr, _ = ref32.solve(want, N, w, d, nsat, ld, rrbc, rcbr, b[r])
            assert np.abs(x[r].astype(np.float64) - xr).max() / np.abs(xr).max() < F32_TOL
    finally:
        L.bdilu_ilu_destroy_f32(h)
