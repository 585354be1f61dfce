"""Regression tests for the round-1 review findings: per-device state, grid limits, reproducible norms, host-pointer
error paths."""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from block_diag_ilu_b200 import _lib  # noqa: E402
from block_diag_ilu_b200 import device as dev  # noqa: E402
from block_diag_ilu_b200 import synth  # noqa: E402


def test_more_than_65535_systems_in_matvec_and_scale(oracle):
    """grid.y / grid.z hold the system index: wide batches go out in slabs (they used to fail with invalid value)."""
    nsys, N, w, d = 70000, 2, 2, 1
    base_d, base_b = synth.make_batch(7, N, w, d, "D", seed0=5)
    data = np.tile(base_d, (nsys // 7 + 1, 1))[:nsys].copy()
    b = np.tile(base_b, (nsys // 7 + 1, 1, 1))[:nsys].copy()
    dd, bb = torch.from_numpy(data).cuda(), torch.from_numpy(b).cuda()
    out = dev.dot_vec(dd, bb, N, w, d).cpu().numpy()
    for s in (0, 6, 65534, 65535, 65536, nsys - 1):
        want = oracle.dot_vec(data[s].copy(), N, w, d, 0, w, b[s, 0])
        assert np.abs(out[s, 0] - want).max() <= 1e-13 * np.abs(want).max()
    n2 = dev.residual_norm2(dd, bb, torch.zeros_like(bb), N, w, d).cpu().numpy()
    assert np.allclose(n2[:, 0], (b[:, 0] ** 2).sum(axis=1), rtol=1e-13)
    sc = dev.scale_diag_add(dd, -0.5, 1.0, N, w, d).cpu().numpy()
    want = oracle.scale_diag_add(data[nsys - 1].copy(), N, w, d, 0, w, -0.5, 1.0)
    assert np.array_equal(sc[nsys - 1], want)
    x, ilu = dev.factor_solve(dd.clone(), bb, N, w, d)
    assert int(ilu.info.abs().max()) == 0 and int(ilu.flags.abs().max()) == 0
    assert torch.equal(x[0], x[7 * (nsys // 7 - 1)])


def test_residual_norms_are_bit_reproducible():
    N, w, d, nrhs = 700, 32, 2, 5
    data, b = synth.make_batch(3, N, w, d, "D", seed0=8, nrhs=nrhs)
    dd, bb = torch.from_numpy(data).cuda(), torch.from_numpy(b).cuda()
    xx = torch.from_numpy(np.random.default_rng(1).uniform(-1, 1, b.shape)).cuda()
    first = dev.residual_norm2(dd, bb, xx, N, w, d).clone()
    for _ in range(50):
        assert torch.equal(dev.residual_norm2(dd, bb, xx, N, w, d), first)
    # and they are the norms
    ax = dev.dot_vec(dd, xx, N, w, d)
    want = ((bb - ax) ** 2).sum(dim=2)
    assert torch.allclose(first, want, rtol=1e-12)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_one_process_two_devices(oracle):
    """Function attributes, handles and the host workspace are per device: the same process drives cuda:0 and cuda:1."""
    L = _lib.lib()
    for (N, w, d, nrhs) in [(100, 32, 2, 1), (40, 64, 3, 16), (300, 16, 1, 2)]:
        data, b = synth.make_batch(2, N, w, d, "D", seed0=N, nrhs=nrhs)
        xs = []
        for di in (0, 1, 0):
            with torch.cuda.device(di):
                ilu = dev.BatchedILU(torch.from_numpy(data).to("cuda:%d" % di), N, w, d, copy=True)
                xs.append(ilu.solve(torch.from_numpy(b).to("cuda:%d" % di)).cpu())
        assert torch.equal(xs[0], xs[1]) and torch.equal(xs[0], xs[2])
    # host-pointer entry points: a handle made on device 1 is used while device 0 is current
    N, w, d = 50, 8, 2
    data, b = synth.make_batch(3, N, w, d, "D", seed0=2)
    hx = np.zeros_like(b)
    info = np.zeros(3, np.int32)
    flags = np.zeros(3, np.int32)
    handle = C.c_void_p()
    with torch.cuda.device(1):
        assert L.bdilu_ilu_create(C.byref(handle), _lib.ptr(data.copy()), data.shape[1], 3, N, w, d, 0, w, 0, None, None,
                                  None, _lib.ptr(info)) == 0
    with torch.cuda.device(0):
        assert L.bdilu_ilu_solve(handle, _lib.ptr(b), _lib.ptr(hx), 1, _lib.ptr(flags)) == 0
        hx2 = np.zeros_like(b)
        assert L.bdilu_host_factor_solve(_lib.ptr(data), data.shape[1], 3, N, w, d, 0, w, _lib.ptr(b), _lib.ptr(hx2), 1,
                                         _lib.ptr(info), _lib.ptr(flags)) == 0
    L.bdilu_ilu_destroy(handle)
    assert np.array_equal(hx, hx2)
    ref = data[1].copy()
    _, rbc, cbr, _ = oracle.factor(ref, N, w, d, 0, w)
    xr, _ = oracle.solve(ref, N, w, d, 0, w, rbc, cbr, b[1, 0])
    assert np.abs(hx[1, 0] - xr).max() <= 1e-12 * np.abs(xr).max()


def test_host_solve_without_flags_and_after_errors():
    """bdilu_ilu_solve with flags == NULL (its own buffer, freed on every path) and argument errors leave the handle usable."""
    L = _lib.lib()
    N, w, d = 30, 4, 1
    data, b = synth.make_batch(2, N, w, d, "D", seed0=4, nrhs=3)
    handle = C.c_void_p()
    assert L.bdilu_ilu_create(C.byref(handle), _lib.ptr(data.copy()), data.shape[1], 2, N, w, d, 0, w, 0, None, None, None,
                              None) == 0
    hx = np.zeros_like(b)
    assert L.bdilu_ilu_solve(handle, _lib.ptr(b), _lib.ptr(hx), 3, None) == 0
    assert L.bdilu_ilu_solve(handle, None, _lib.ptr(hx), 3, None) == _lib.ERR_ARG
    bn = b.copy()
    bn[1, 2, 5] = np.nan
    assert L.bdilu_ilu_solve(handle, _lib.ptr(bn), _lib.ptr(hx), 3, None) == 6   # first nonzero flag: 1-based NaN index
    hx2 = np.zeros_like(b)
    assert L.bdilu_ilu_solve(handle, _lib.ptr(b), _lib.ptr(hx2), 3, None) == 0
    L.bdilu_ilu_destroy(handle)
    x, _ = dev.factor_solve(torch.from_numpy(data).cuda(), torch.from_numpy(b).cuda(), N, w, d)
    assert np.array_equal(x.cpu().numpy(), hx2)
