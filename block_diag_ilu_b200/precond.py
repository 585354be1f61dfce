"""Device-resident preconditioner adapter: the caller side of the hot path.

What an implicit integrator does with this library (external/anyode/include/anyode/anyode_iterative.hpp:19-88
with JacMat_t = BlockDiagMatrix, Decomp_t = ILU-like; anyode_decomposition.hpp:16-21 for factorize()/solve()):

    prec_setup(t, y, fy, jac_ok, gamma)  : (re)compute J unless jac_ok, M = I - gamma*J, factor M      (:38-63)
    prec_solve_left(t, y, fy, r, ...)    : solve P z = r with the ILU factors of M                     (:65-86)
    jtimes(vec, ...)                     : J @ vec                                                     (:19-36)

Same names, argument meaning, counters (m_njacvec_dot, m_nprec_setup, m_nprec_solve) and status codes as the
reference, with everything a batch wide and resident in HBM: J, M's factors, r and z are torch CUDA tensors, the
assembly of M is fused into the factor kernel's load (bdilu_factor_scaled), and the vectors of one Krylov block
([nsys, nrhs, n]) are solved in one launch.  No host round trip happens between GMRES iterations.
"""
import torch

from . import device as _dev

SUCCESS, RECOVERABLE_ERROR, UNRECOVERABLE_ERROR = 0, 1, -1  # AnyODE::Status (anyode.hpp)


class BlockDiagPreconditioner:
    """P ~= I - gamma*J for a batch of block-diagonal Jacobians with ndiag diagonal-only couplings.

    jac_fn(t, y, fy, out) -- optional callback that writes the packed Jacobians (BlockDiagMatrix::m_data per system,
    float64 [nsys, ndata]) into `out` on the device (the reference's dense_jac_cmaj slot); alternatively keep
    `self.jac` up to date yourself and call prec_setup(..., jac_ok=True).
    """

    def __init__(self, nblocks, blockw, ndiag, nsat=0, ld=None, nsys=1, device="cuda", jac_fn=None):
        self.shape = _dev.Shape(nblocks, blockw, ndiag, nsat, ld)
        self.nsys = int(nsys)
        self.device = torch.device(device)
        self.jac_fn = jac_fn
        self.jac = None        # m_jac_cache
        self._store = None     # m_M_cache: holds the factors of M after prec_setup
        self.decomp = None     # m_decomp_cache
        self.m_njacvec_dot = self.m_nprec_setup = self.m_nprec_solve = 0

    def _dims(self):
        s = self.shape
        return dict(nblocks=s.nblocks, blockw=s.blockw, ndiag=s.ndiag, nsat=s.nsat, ld=s.ld)

    def set_jacobian(self, jac):
        """Adopt `jac` (float64 CUDA [nsys, ndata], not copied) as the cached Jacobian."""
        jac = _dev._check_data(jac, self.shape)
        if jac.shape[0] != self.nsys:
            raise ValueError("expected %d systems, got %d" % (self.nsys, jac.shape[0]))
        self.jac = jac

    def jtimes(self, vec, t=None, y=None, fy=None):
        """(status, J @ vec) -- anyode_iterative.hpp:19-36 (the cached Jacobian is used; recompute it in jac_fn)."""
        if self.jac is None:
            raise RuntimeError("no Jacobian: call prec_setup or set_jacobian first")
        out = _dev.dot_vec(self.jac, vec, **self._dims())
        self.m_njacvec_dot += 1
        return SUCCESS, out

    def prec_setup(self, t, y, fy, jac_ok, gamma):
        """(status, jac_recomputed) -- anyode_iterative.hpp:38-63."""
        jac_recomputed = False
        if not jac_ok or self.jac is None:
            if self.jac_fn is None:
                if self.jac is None:
                    raise RuntimeError("no Jacobian and no jac_fn")
            else:
                if self.jac is None:
                    self.jac = torch.zeros((self.nsys, self.shape.ndata), dtype=torch.float64, device=self.device)
                self.jac_fn(t, y, fy, self.jac)
                jac_recomputed = True
        status = self.factorize(gamma)
        self.m_nprec_setup += 1
        return status, jac_recomputed

    def factorize(self, gamma):
        """M = I - gamma*J assembled and factored in one pass; the previous factor storage is reused."""
        self.decomp = _dev.BatchedILU.from_scaled(self.jac, -float(gamma), 1.0, out=self._store, check=False,
                                                  **self._dims())
        self._store = self.decomp.data
        return SUCCESS if not bool((self.decomp.info != 0).any().item()) else RECOVERABLE_ERROR

    def solve(self, r, out=None):
        """DecompositionBase::solve: z with the shape of r; per-vector flags in self.decomp.flags."""
        if self.decomp is None:
            raise RuntimeError("prec_setup / factorize first")
        return self.decomp.solve(r, out=out)

    def prec_solve_left(self, t, y, fy, r, gamma=None, delta=None, ewt=None, out=None, check=True):
        """(status, z) with P z = r -- anyode_iterative.hpp:65-86.  r: [n], [nsys, n] or [nsys, nrhs, n].

        check=False skips the device->host read of the flags (status is then SUCCESS unconditionally; read
        self.decomp.flags when convenient), which keeps a Krylov loop free of synchronisation points.
        """
        if ewt is not None:
            raise NotImplementedError("Not implemented: ewt in prec_solve_left")
        self.m_nprec_solve += 1
        z = self.solve(r, out=out)
        if check and bool((self.decomp.flags != 0).any().item()):
            return RECOVERABLE_ERROR, z
        return SUCCESS, z
