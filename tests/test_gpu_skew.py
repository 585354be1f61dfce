"""GPU parity of the experimental systolic blockw-32 solve kernel (csrc/bdilu_solve_skew.cu; opt-in through test hook 8,
not the default path: it is slower than quad32 on B200) and of the two-pass dispatch around it:
systems that never pivoted run on it, systems that pivoted / vectors with NaN / zero U diagonals are handed to the general
wavefront kernel, which also reports the reference's flags (block_diag_ilu.hpp:491-495).  Every case is compared with the
oracle through the C ABI; the debug counters prove which kernel solved what."""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from block_diag_ilu_b200 import _lib  # noqa: E402
from block_diag_ilu_b200 import device as dev  # noqa: E402
from block_diag_ilu_b200 import synth  # noqa: E402

from .test_gpu_parity import run_case  # noqa: E402


@pytest.fixture(autouse=True)
def force_skew():
    old = _lib.lib().bdilu_debug_force_chain_solve(8)
    yield
    _lib.lib().bdilu_debug_force_chain_solve(old)


def stats(reset=False):
    out = (C.c_int64 * 3)()
    assert _lib.lib().bdilu_debug_skew_stats(out, int(reset)) == 0
    return list(out)


@pytest.mark.parametrize("N", [8, 33, 63, 64, 65, 66, 71, 72, 73, 95, 96, 97, 100, 127, 128, 129, 255, 256, 257, 1000])
@pytest.mark.parametrize("d", [1, 2])
def test_chain_lengths(oracle, N, d):
    """Every alignment of the chain length against the 8-block groups and the 32-block rounds."""
    stats(reset=True)
    run_case(oracle, 2, N, 32, d, "D", nrhs=1, seed0=N)
    assert stats() == ([2, 0, 0] if N >= 64 else [0, 0, 0])   # chains shorter than 64 blocks stay on the chain kernel


@pytest.mark.parametrize("ld", [32, 34, 40])
@pytest.mark.parametrize("nrhs", [1, 3])
def test_padded_ld_and_several_vectors(oracle, ld, nrhs):
    stats(reset=True)
    run_case(oracle, 3, 77, 32, 2, "D", ld=ld, nrhs=nrhs, seed0=5)
    assert stats() == [3 * nrhs, 0, 0]


def test_mixed_batch_pivoting_systems_go_to_the_general_kernel(oracle):
    """Systems 1 and 3 of 5 are non-dominant (they pivot): the systolic kernel skips them, the second pass solves them."""
    N, w, d, nrhs = 150, 32, 2, 2
    dD, bD = synth.make_batch(5, N, w, d, "D", seed0=100, nrhs=nrhs)
    dP, bP = synth.make_batch(5, N, w, d, "P", seed0=200, nrhs=nrhs)
    data, b = dD.copy(), bD.copy()
    for s in (1, 3):
        data[s], b[s] = dP[s], bP[s]
    stats(reset=True)
    ilu = dev.BatchedILU(torch.from_numpy(data).cuda(), N, w, d, copy=True)
    x = ilu.solve(torch.from_numpy(b).cuda()).cpu().numpy()
    assert stats() == [3 * nrhs, 0, 2 * nrhs]
    for s in range(5):
        ref = data[s].copy()
        ipiv, rbc, cbr, info = oracle.factor(ref, N, w, d, 0, w)
        assert np.array_equal(ilu.ipiv[s].cpu().numpy(), ipiv)
        for r in range(nrhs):
            xr, flag = oracle.solve(ref, N, w, d, 0, w, rbc, cbr, b[s, r])
            assert flag == int(ilu.flags[s, r]) == 0
            tol = 1e-12 if s not in (1, 3) else 1e-9   # the pivoting systems are ill-conditioned chains (see test_gpu_parity)
            assert np.abs(x[s, r] - xr).max() <= tol * np.abs(xr).max()


def test_flags_nan_and_zero_diagonal(oracle):
    """NaN in b and an exactly zero U diagonal: handed back at the end, flags as the reference returns them."""
    N, w, d, nrhs = 70, 32, 2, 3
    data, b = synth.make_batch(2, N, w, d, "D", seed0=9, nrhs=nrhs)
    b[0, 1, 40 * w + 3] = np.nan
    b[0, 1, 60 * w + 30] = np.nan
    b[1, 2, 0] = np.nan
    stats(reset=True)
    ilu = dev.BatchedILU(torch.from_numpy(data).cuda(), N, w, d, copy=True)
    x = ilu.solve(torch.from_numpy(b).cuda()).cpu().numpy()
    want = np.zeros((2, nrhs), dtype=np.int32)
    want[0, 1], want[1, 2] = 40 * w + 3 + 1, 1
    assert np.array_equal(ilu.flags.cpu().numpy(), want)
    assert stats() == [4, 2, 0]
    # the clean vectors are untouched by their neighbours' NaN
    ref = data[0].copy()
    _, rbc, cbr, _ = oracle.factor(ref, N, w, d, 0, w)
    xr, _ = oracle.solve(ref, N, w, d, 0, w, rbc, cbr, b[0, 0])
    assert np.abs(x[0, 0] - xr).max() <= 1e-12 * np.abs(xr).max()
    # zero U diagonals in system 1 after the factorisation: flag = N*w + index of the LAST one met by the backward sweep
    fac = ilu.data.clone()
    for blk, k in [(5, 10), (33, 31), (33, 0)]:
        fac[1, (blk * w + k) * w + k] = 0.0
    ilu.data = fac
    b2 = np.nan_to_num(b)
    stats(reset=True)
    ilu.solve(torch.from_numpy(b2).cuda())
    fl = ilu.flags.cpu().numpy()
    assert list(fl[0]) == [0, 0, 0] and list(fl[1]) == [N * w + 33 * w + 31] * nrhs
    assert stats() == [3, 3, 0]


def test_full_shape_two_systems_and_repeatability(oracle):
    """cfg3's shape, and 50 repeats bit-identical (the kernel has no polling and no atomics on data)."""
    N, w, d = 8192, 32, 2
    ilu = run_case(oracle, 2, N, w, d, "D", seed0=31)
    _, b = synth.make_batch(2, N, w, d, "D", seed0=31)
    bb = torch.from_numpy(b).cuda()
    x0 = ilu.solve(bb).clone()
    for _ in range(50):
        x = ilu.solve(bb)
        assert torch.equal(x, x0)


def test_wide_batch_of_vectors(oracle):
    """More vectors than SMs: 40 systems x 8 vectors = 320 CTAs in three waves."""
    stats(reset=True)
    run_case(oracle, 40, 80, 32, 2, "D", nrhs=8, seed0=3)
    assert stats() == [320, 0, 0]
