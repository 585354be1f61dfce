"""The parity sets SURVEY.md section 8(d) prescribes, to the letter, on the AUTOMATICALLY dispatched kernels:

  cfg3  8 dominant + 2 pivoting systems at the full shape (w=32, N=8192, d=2)
  cfg4  the full shape with all 256 right-hand sides, 8 columns compared; one pivoting system
  cfg5  8 dominant + 2 pivoting systems inside batches of 512 / 1024 / 2048 / 4096 systems (the per-GPU slab sizes at
        8 / 4 / 2 / 1 GPUs, which the solve dispatcher treats differently)

plus the reference's recorded outputs (tests/golden/ref_vectors.npz) fed through the C ABI directly, and bit-identical
repeats of every wavefront solve kernel (their shared-memory protocols poll; a race would show as a changed bit).
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from block_diag_ilu_b200 import _lib  # noqa: E402
from block_diag_ilu_b200 import device as dev  # noqa: E402
from block_diag_ilu_b200 import synth  # noqa: E402

from .test_gpu_parity import PATHS, TOL, YARDSTICK_MAX, block_rel_err, run_case, two_cpu_yardstick  # noqa: E402
from .test_oracle import GOLDEN  # noqa: E402


def test_cfg3_contract_set(oracle):
    run_case(oracle, 8, 8192, 32, 2, "D", seed0=3_000_000)
    run_case(oracle, 2, 8192, 32, 2, "P", seed0=3_100_000, yardstick=True)
    run_case(oracle, 2, 8192, 32, 2, "R", seed0=3_200_000)   # pivots in every block, held to the full 1e-12


def test_cfg4_contract_set(oracle):
    """Full shape, 256 vectors on the automatic path (many-vector wavefront kernel), 8 of the 256 columns compared."""
    N, w, d, nrhs = 2048, 64, 3, 256
    data, b = synth.make_batch(1, N, w, d, "D", seed0=4_000_000, nrhs=nrhs)
    ilu = dev.BatchedILU(torch.from_numpy(data).cuda(), N, w, d, copy=True)
    x = ilu.solve(torch.from_numpy(b).cuda())
    torch.cuda.synchronize()
    ref = data[0].copy()
    ipiv, rbc, cbr, info = oracle.factor(ref, N, w, d, 0, w)
    assert info == 0 and np.array_equal(ilu.ipiv[0].cpu().numpy(), ipiv)
    assert block_rel_err(ilu.data[0].cpu().numpy(), ref, N, w, w) < TOL
    assert int(ilu.flags.abs().max()) == 0
    for r in (0, 1, 7, 8, 100, 127, 254, 255):
        xr, flag = oracle.solve(ref, N, w, d, 0, w, rbc, cbr, b[0, r])
        assert flag == 0
        assert np.abs(x[0, r].cpu().numpy() - xr).max() < TOL * np.abs(xr).max(), r
    run_case(oracle, 1, 2048, 64, 3, "P", seed0=4_100_000, yardstick=True)
    run_case(oracle, 1, 2048, 64, 3, "R", nrhs=3, seed0=4_200_000)


@pytest.mark.parametrize("nsys", [512, 1024, 2048, 4096])
def test_cfg5_contract_set_per_slab_size(oracle, nsys):
    """Ten distinct systems (8 D + 2 P) spread through a batch of the slab size a rank owns at 8 / 4 / 2 / 1 GPUs."""
    N, w, d = 512, 8, 2
    dD, bD = synth.make_batch(8, N, w, d, "D", seed0=5_000_000)
    dP, bP = synth.make_batch(2, N, w, d, "P", seed0=5_100_000)
    base_d, base_b = np.concatenate([dD, dP]), np.concatenate([bD, bP])
    reps = nsys // 10 + 1
    data = np.tile(base_d, (reps, 1))[:nsys].copy()
    b = np.tile(base_b, (reps, 1, 1))[:nsys].copy()
    ilu = dev.BatchedILU(torch.from_numpy(data).cuda(), N, w, d, copy=True)
    x = ilu.solve(torch.from_numpy(b).cuda()).cpu().numpy()
    got = ilu.data.cpu().numpy()
    assert int(ilu.info.abs().max()) == 0 and int(ilu.flags.abs().max()) == 0
    where = [0, 1, 2, 3, nsys // 2 // 10 * 10 + 4, nsys // 2 // 10 * 10 + 5, (nsys - 10) // 10 * 10 + 6,
             (nsys - 10) // 10 * 10 + 7, 8, (nsys - 10) // 10 * 10 + 9]   # first, middle, last copies of the ten
    for s in where:
        k = s % 10
        ref = base_d[k].copy()
        ipiv, rbc, cbr, info = oracle.factor(ref, N, w, d, 0, w)
        assert np.array_equal(ilu.ipiv[s].cpu().numpy(), ipiv) and np.array_equal(ilu.colbyrow[s].cpu().numpy(), cbr)
        ftol = xtol = TOL
        if k >= 8:
            yf, yx = two_cpu_yardstick(oracle, base_d[k], base_b[k, 0], N, w, d, 0, w)
            assert yf <= YARDSTICK_MAX and yx <= YARDSTICK_MAX
            ftol, xtol = max(TOL, 10 * yf), max(TOL, 10 * yx)
        assert block_rel_err(got[s], ref, N, w, w) < ftol
        xr, flag = oracle.solve(ref, N, w, d, 0, w, rbc, cbr, base_b[k, 0])
        assert flag == 0 and np.abs(x[s, 0] - xr).max() < xtol * np.abs(xr).max(), s
    # every copy of a system gives the same bits wherever it sits in the batch
    for k in range(10):
        assert np.array_equal(x[k], x[k + 10 * ((nsys - 10) // 10)])


def golden_names():
    z = np.load(GOLDEN)
    return sorted({k.split("/")[0] for k in z.files if k.endswith("/data")})


@pytest.mark.parametrize("name", golden_names())
def test_reference_recorded_outputs_through_the_c_abi(name):
    """tests/golden/ref_vectors.npz holds inputs AND outputs of the reference header (made by make_ref_vectors.py in the
    build container): the GPU is compared with those numbers directly, no oracle in the loop."""
    z = np.load(GOLDEN)
    N, w, d, ld, nsat, seed = (int(v) for v in z[name + "/shape"])
    fam = str(z[name + "/family"])
    data, b = z[name + "/data"].copy(), z[name + "/b"].copy()
    dd = torch.from_numpy(data[None]).cuda()
    bb = torch.from_numpy(b[None, None]).cuda()
    av = dev.dot_vec(dd, bb, N, w, d, nsat=nsat, ld=ld).cpu().numpy()[0, 0]
    assert np.abs(av - z[name + "/dot_vec"]).max() <= 1e-13 * np.abs(z[name + "/dot_vec"]).max()
    ilu = dev.BatchedILU(dd, N, w, d, nsat=nsat, ld=ld, copy=True)
    x = ilu.solve(bb).cpu().numpy()[0, 0]
    assert np.array_equal(ilu.ipiv[0].cpu().numpy(), z[name + "/ipiv"])
    assert np.array_equal(ilu.rowbycol[0].cpu().numpy(), z[name + "/rowbycol"])
    assert np.array_equal(ilu.colbyrow[0].cpu().numpy(), z[name + "/colbyrow"])
    tol = 1e-12 if fam == "D" else 1e-10   # same bars as the CPU restatement is held to (tests/test_oracle.py)
    want = z[name + "/factored"]
    got = ilu.data[0].cpu().numpy()
    nb = N * w * ld
    g = got[:nb].reshape(N, w, ld)[:, :, :w].reshape(N, -1)
    r = want[:nb].reshape(N, w, ld)[:, :, :w].reshape(N, -1)
    assert (np.abs(g - r).max(axis=1) / np.abs(r).max(axis=1)).max() < tol
    assert int(ilu.flags[0, 0]) == 0
    assert np.abs(x - z[name + "/x"]).max() <= tol * np.abs(z[name + "/x"]).max()


@pytest.mark.parametrize("path,shape", [
    ("auto", (8192, 32, 2, 1)),           # quad32, the headline kernel: in-band sentinel ring, counter-free recycling
    ("wave32_onerow", (4096, 32, 2, 1)),
    ("wave_generic", (2048, 16, 1, 1)),
    ("auto", (1024, 16, 1, 1)),           # cfg2's dispatch: the prefetched narrow-block kernel (cp.async ring, images read across lanes)
    ("mrhs_wave", (1024, 16, 1, 1)),      # the wavefront-group kernel at cfg2's shape
    ("pf", (512, 8, 2, 1)),               # cfg5's shape
    ("mrhs_wave", (2048, 64, 3, 16)),
    ("skew32", (4096, 32, 2, 1)),
    ("half32", (4096, 32, 2, 1)),         # two CTAs per chain: progress words in global memory, fences, throttle
])
def test_wavefront_kernels_repeat_bit_identically(path, shape):
    """200 solves of the same batch: any race in the polling protocols would surface as a differing bit (or a trap)."""
    N, w, d, nrhs = shape
    L = _lib.lib()
    old = L.bdilu_debug_force_chain_solve(PATHS[path])
    try:
        data, b = synth.make_batch(3, N, w, d, "D", seed0=91, nrhs=nrhs)
        ilu = dev.BatchedILU(torch.from_numpy(data).cuda(), N, w, d, copy=True)
        bb = torch.from_numpy(b).cuda()
        x0 = ilu.solve(bb).clone()
        f0 = ilu.flags.clone()
        for _ in range(200):
            x = ilu.solve(bb)
            assert torch.equal(x, x0) and torch.equal(ilu.flags, f0)
    finally:
        L.bdilu_debug_force_chain_solve(old)
