"""The integrator-facing adapter (anyode_iterative.hpp:19-88 restated for device-resident batches)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from block_diag_ilu_b200 import precond, synth  # noqa: E402


def test_adapter_surface_without_gpu():
    p = precond.BlockDiagPreconditioner(4, 3, 1, nsys=2, device="cpu")
    assert (p.m_njacvec_dot, p.m_nprec_setup, p.m_nprec_solve) == (0, 0, 0)
    with pytest.raises(RuntimeError):
        p.jtimes(torch.zeros(12, dtype=torch.float64))
    with pytest.raises(RuntimeError):
        p.solve(torch.zeros(12, dtype=torch.float64))
    with pytest.raises(NotImplementedError):
        p.prec_solve_left(0.0, None, None, torch.zeros(12, dtype=torch.float64), ewt=torch.ones(12))


@pytest.mark.gpu
@pytest.mark.parametrize("N,w,d", [(16, 4, 1), (64, 32, 2), (20, 8, 2)])
def test_prec_setup_solve_matches_oracle(oracle, N, w, d):
    from block_diag_ilu_b200 import device as dev
    nsys, gamma = 3, 0.05
    J, r = synth.make_batch(nsys, N, w, d, "P", seed0=5, nrhs=4)
    calls = []

    def jac_fn(t, y, fy, out):
        calls.append(t)
        out.copy_(torch.from_numpy(J))

    p = precond.BlockDiagPreconditioner(N, w, d, nsys=nsys, jac_fn=jac_fn)
    status, recomputed = p.prec_setup(0.0, None, None, False, gamma)
    assert status == precond.SUCCESS and recomputed and calls == [0.0]
    rd = torch.from_numpy(r).cuda()
    status, z = p.prec_solve_left(0.0, None, None, rd, gamma, 1e-3)
    assert status == precond.SUCCESS and z.shape == rd.shape
    zz = z.cpu().numpy()
    for s in range(nsys):
        M = oracle.scale_diag_add(J[s].copy(), N, w, d, 0, w, -gamma, 1.0)
        ipiv, rbc, cbr, info = oracle.factor(M, N, w, d, 0, w)
        assert info == 0 and np.array_equal(p.decomp.ipiv[s].cpu().numpy(), ipiv)
        for k in range(4):
            xr, flag = oracle.solve(M, N, w, d, 0, w, rbc, cbr, r[s, k])
            assert flag == 0
            assert np.abs(zz[s, k] - xr).max() / np.abs(xr).max() < 1e-12
    # a new gamma with jac_ok: J is reused, the factor storage too
    store = p.decomp.data.data_ptr()
    status, recomputed = p.prec_setup(0.1, None, None, True, 2 * gamma)
    assert status == precond.SUCCESS and not recomputed and calls == [0.0] and p.decomp.data.data_ptr() == store
    want = dev.BatchedILU(dev.scale_diag_add(p.jac, -2 * gamma, 1.0, N, w, d), N, w, d)
    assert torch.equal(p.decomp.data, want.data)
    # jtimes is the block-diagonal matvec of the cached Jacobian
    v = torch.from_numpy(r[:, 0]).cuda()
    status, jv = p.jtimes(v)
    assert torch.equal(jv, dev.dot_vec(p.jac, v, N, w, d))
    assert (p.m_njacvec_dot, p.m_nprec_setup, p.m_nprec_solve) == (1, 2, 1)


@pytest.mark.gpu
def test_preconditioned_richardson_converges():
    """The ILU of a diagonally dominant M is a good P: z <- z + P^-1 (r - M z) contracts fast."""
    from block_diag_ilu_b200 import device as dev
    N, w, d, nsys, gamma = 128, 32, 2, 4, 0.02
    J, r = synth.make_batch(nsys, N, w, d, "D", seed0=11)
    p = precond.BlockDiagPreconditioner(N, w, d, nsys=nsys)
    p.set_jacobian(torch.from_numpy(J).cuda())
    assert p.prec_setup(0.0, None, None, True, gamma) == (precond.SUCCESS, False)
    M = dev.scale_diag_add(p.jac, -gamma, 1.0, N, w, d)
    rd = torch.from_numpy(r).cuda()
    z = torch.zeros_like(rd)
    norms = []
    for _ in range(4):
        res = rd - dev.dot_vec(M, z, N, w, d)
        norms.append(float(res.norm()))
        z = z + p.prec_solve_left(0.0, None, None, res, check=False)[1]
    assert norms[1] < 1e-4 * norms[0] and norms[3] < 1e-12 * norms[0], norms


@pytest.mark.gpu
def test_singular_M_is_a_recoverable_error():
    N, w, d = 6, 4, 1
    J, _ = synth.make_batch(1, N, w, d, "D", seed0=3)
    J[0, : w * w] = 0.0            # block 0 of J is zero ...
    p = precond.BlockDiagPreconditioner(N, w, d)
    p.set_jacobian(torch.from_numpy(J).cuda())
    assert p.prec_setup(0.0, None, None, True, 0.5)[0] == precond.SUCCESS       # ... M's block 0 is the identity
    J[0, : w * w] = np.eye(w).ravel() * 2.0
    p.set_jacobian(torch.from_numpy(J).cuda())
    assert p.prec_setup(0.0, None, None, True, 0.5)[0] == precond.RECOVERABLE_ERROR  # I - 0.5*2*I = 0
