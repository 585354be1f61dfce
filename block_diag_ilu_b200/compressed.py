"""Host-side mirror of the reference's Cython wrapper (block_diag_ilu/_block_diag_ilu.pyx).

Same names, argument meaning and error behaviour as `Compressed` (= PyBlockDiagMatrix, pyx:15-93),
`Compressed_from_dense` (:95-117), `Compressed_from_data` (:120-129), `ILU` (= PyILU, :140-154) and the
solve-flag mapping (:132-137).  The container is a numpy array holding BlockDiagMatrix::m_data byte for byte
(block_diag_ilu.hpp:98-102, 203-221); every computation (dot_vec, scale_diag_add, factor, solve) is one call
into libbdilu_b200.so through the C ABI of include/bdilu.h and runs on the GPU -- there is no CPU fallback.

Beyond the reference: `ILU.solve` also takes [nrhs, n] arrays and CUDA float64 torch tensors ([n], [nrhs, n]),
which stay on the device; `ILU.from_tensor` builds a batched device-resident factorisation.
"""
import ctypes as C

import numpy as np

from . import _lib
from . import datastruct as ds


class Compressed:
    """BlockDiagMatrix<double>(nullptr, nblocks, blockw, ndiag, nsat, ld) (block_diag_ilu.hpp:110-123)."""

    def __init__(self, nblocks, blockw, ndiag, nsat=0, ld=0):
        self.nblocks, self.blockw, self.ndiag, self._nsat = int(nblocks), int(blockw), int(ndiag), int(nsat)
        self.ld = int(ld) if ld else ds.default_ld(self.blockw)
        if self.ld < self.blockw:
            raise ValueError("ld < blockw")
        o_sub, o_sup, o_bot, o_top = ds.offsets(self.nblocks, self.blockw, self.ndiag, self._nsat, self.ld)
        self._o_sub, self._o_sup, self._o_bot, self._o_top = o_sub, o_sup, o_bot, o_top
        self._data = np.zeros(ds.ndata(self.nblocks, self.blockw, self.ndiag, self._nsat, self.ld))

    # -- container -------------------------------------------------------------------------------
    @property
    def data(self):
        """A copy of m_data, as in the reference (pyx:27-31)."""
        return self._data.copy()

    @property
    def nsat(self):
        return self._nsat

    def copy(self):
        return Compressed_from_data(self._data, self.nblocks, self.blockw, self.ndiag, self._nsat, self.ld)

    def _dims(self):
        return (self.nblocks, self.blockw, self.ndiag, self._nsat, self.ld)

    def _blk(self, bi, ri, ci):
        return (bi * self.blockw + ci) * self.ld + ri

    def _diag(self, di, bi, ci):
        return ds.diag_chunk(self.nblocks, di, bi) * self.ld + ci

    def _sat(self, si, bi, ci):
        return ds.sat_chunk(si, bi) * self.ld + ci

    def get_block(self, bi, ri, ci):
        return float(self._data[self._blk(bi, ri, ci)])

    def get_sub(self, di, bi, ci):
        return float(self._data[self._o_sub + self._diag(di, bi, ci)])

    def get_sup(self, di, bi, ci):
        return float(self._data[self._o_sup + self._diag(di, bi, ci)])

    def get_bot(self, si, bi, ci):
        return float(self._data[self._o_bot + self._sat(si, bi, ci)])

    def get_top(self, si, bi, ci):
        return float(self._data[self._o_top + self._sat(si, bi, ci)])

    def set_block(self, bi, ri, ci, value):
        self._data[self._blk(bi, ri, ci)] = value

    def set_sub(self, di, bi, ci, value):
        self._data[self._o_sub + self._diag(di, bi, ci)] = value

    def set_sup(self, di, bi, ci, value):
        self._data[self._o_sup + self._diag(di, bi, ci)] = value

    def set_bot(self, si, bi, ci, value):
        self._data[self._o_bot + self._sat(si, bi, ci)] = value

    def set_top(self, si, bi, ci, value):
        self._data[self._o_top + self._sat(si, bi, ci)] = value

    # -- indexing (valid_index / operator(), block_diag_ilu.hpp:222-275) -----------------------------
    def _locate(self, ri, ci):
        """Offset into m_data of dense element (ri, ci), or None where it is structurally zero."""
        w, N, d, ns = self.blockw, self.nblocks, self.ndiag, self._nsat
        bri, bci = ri // w, ci // w
        lri, lci = ri - bri * w, ci - bci * w
        if bri == bci:
            return self._blk(bri, lri, lci)
        if lri != lci:
            return None
        if bri > bci:
            if bri - bci > d:
                if N - bri + bci <= ns:
                    return self._o_bot + self._sat(N - 1 - bri + bci, bci, lci)
                return None
            return self._o_sub + self._diag(bri - bci - 1, bci, lci)
        if bci - bri > d:
            if N - bci + bri <= ns:
                return self._o_top + self._sat(N - bci + bri - 1, bri, lri)
            return None
        return self._o_sup + self._diag(bci - bri - 1, bri, lri)

    def __getitem__(self, key):
        ri, ci = key
        k = self._locate(int(ri), int(ci))
        return 0.0 if k is None else float(self._data[k])

    def to_dense(self):
        return ds.unpack_dense(self._data, self.nblocks, self.blockw, self.ndiag, self._nsat, self.ld)

    # -- compute: one C-ABI call each, on the GPU ------------------------------------------------------
    def dot_vec(self, vec):
        """BlockDiagMatrix::dot_vec (block_diag_ilu.hpp:295-326) -> bdilu_host_dot_vec."""
        vec = np.ascontiguousarray(vec, dtype=np.float64)
        n = self.nblocks * self.blockw
        if vec.ndim != 1 or vec.size != n:
            raise ValueError("vec must be 1-D with nblocks*blockw elements")
        out = np.empty(n)
        _lib.check(_lib.lib().bdilu_host_dot_vec(_lib.ptr(self._data), *self._dims(), _lib.ptr(vec), _lib.ptr(out), 1))
        return out

    def scale_diag_add(self, other, scale, diag_add):
        """self = scale*other + diag_add*I (block_diag_ilu.hpp:276-287) -> bdilu_host_scale_diag_add."""
        if not isinstance(other, Compressed):
            raise TypeError("other must be a Compressed")
        if other._data.size != self._data.size:
            raise ValueError("Incompatible sizes")
        _lib.check(_lib.lib().bdilu_host_scale_diag_add(_lib.ptr(self._data), _lib.ptr(other._data), *self._dims(),
                                                        float(scale), float(diag_add)))


def Compressed_from_dense(A, nblocks, blockw, ndiag, nsat=0):
    """Dense -> packed with tight ld = blockw; same checks and messages as pyx:95-117, packed in bulk."""
    A = np.asarray(A, dtype=np.float64)
    if A.ndim != 2 or A.shape[0] != A.shape[1]:
        raise ValueError("A not square")
    if A.shape[0] != nblocks * blockw:
        raise ValueError("A shape does not match nblocks & blockw")
    if ndiag > nblocks - 1:
        raise ValueError("too many diagonals")
    cmprs = Compressed(nblocks, blockw, ndiag, nsat, ld=blockw)
    cmprs._data[:] = ds.pack_dense(A, nblocks, blockw, ndiag, nsat, blockw)
    return cmprs


def Compressed_from_data(data, nblocks, blockw, ndiag, nsat, ld):
    """Packed m_data -> Compressed (copy); ValueError('Incompatible sizes') as pyx:120-129."""
    cmprs = Compressed(nblocks, blockw, ndiag, nsat, ld)
    data = np.asarray(data, dtype=np.float64)
    if data.ndim != 1 or data.size != cmprs._data.size:
        raise ValueError('Incompatible sizes')
    cmprs._data[:] = data
    return cmprs


def _check_solve_flag(flag, N):
    """pyx:132-137."""
    if flag != 0:
        if flag < N:
            raise ValueError("NaN in b")
        raise ZeroDivisionError("Rank deficient matrix")


def _is_torch(x):
    return type(x).__module__.split(".")[0] == "torch"


class ILU:
    """PyILU (pyx:140-154): copies the matrix, factors the copy (on the GPU), keeps the factors device-resident.

    Unlike the reference binding (whose .pxd has no `except +`, so a singular block aborts the process,
    block_diag_ilu.pxd:41-44) a failed factorisation raises RuntimeError("ILU failed!") -- the message of the
    C++ throw at block_diag_ilu.hpp:483-484.
    """

    def __init__(self, cmprs):
        if not isinstance(cmprs, Compressed):
            raise TypeError("expected a Compressed")
        self._dev = None
        self._h = C.c_void_p()
        self.nblocks, self.blockw = cmprs.nblocks, cmprs.blockw
        self._n = cmprs.nblocks * cmprs.blockw
        pbdm = cmprs.copy()                      # pyx:145
        n = self._n
        self.ipiv = np.empty(n, dtype=np.int32)      # m_ipiv / m_rowbycol / m_colbyrow (hpp:428-429)
        self.rowbycol = np.empty(n, dtype=np.int32)
        self.colbyrow = np.empty(n, dtype=np.int32)
        info = np.zeros(1, dtype=np.int32)
        L = _lib.lib()
        _lib.check(L.bdilu_ilu_create(C.byref(self._h), _lib.ptr(pbdm._data), pbdm._data.size, 1, *pbdm._dims(), 1,
                                      _lib.ptr(self.ipiv), _lib.ptr(self.rowbycol), _lib.ptr(self.colbyrow),
                                      _lib.ptr(info)))
        self.pbdm = pbdm                         # now holds L\U and the scaled sub-diagonals (ILU_inplace semantics)
        self.info = int(info[0])
        if self.info != 0:
            self._destroy()
            raise RuntimeError("ILU failed!")

    @classmethod
    def from_tensor(cls, data, nblocks, blockw, ndiag, nsat=0, ld=None, copy=True):
        """Batched, device-resident: data is a CUDA float64 tensor [nsys, ndata] (or [ndata]).  No host copies."""
        from . import device
        self = cls.__new__(cls)
        self._h = C.c_void_p()
        self.nblocks, self.blockw = int(nblocks), int(blockw)
        self._n = self.nblocks * self.blockw
        self._dev = device.BatchedILU(data, nblocks, blockw, ndiag, nsat, ld, copy=copy, check=True)
        self.pbdm = None
        self.ipiv, self.rowbycol, self.colbyrow = self._dev.ipiv, self._dev.rowbycol, self._dev.colbyrow
        self.info = 0
        return self

    def solve(self, b):
        """x = ILU_inplace::solve(b) (block_diag_ilu.hpp:486-538); raises as pyx:132-137.

        numpy [n] (the reference's signature) or [nrhs, n] -> numpy; a CUDA float64 torch tensor stays on the
        device (only for an ILU made by from_tensor, whose factors already live there)."""
        if _is_torch(b):
            if self._dev is None:
                raise TypeError("torch tensors need a device-resident ILU: use ILU.from_tensor")
            x = self._dev.solve(b)
            flags = self._dev.flags.flatten()
            bad = flags[flags != 0]
            if bad.numel():
                _check_solve_flag(int(bad[0].item()), self._n)
            return x
        if self._dev is not None:
            raise TypeError("this ILU is device-resident: pass a CUDA float64 torch tensor")
        b = np.ascontiguousarray(b, dtype=np.float64)
        if b.ndim not in (1, 2) or b.shape[-1] != self._n:
            raise ValueError("b must have nblocks*blockw elements per vector")
        nrhs = 1 if b.ndim == 1 else b.shape[0]
        x = np.zeros_like(b)
        if nrhs == 0:
            return x
        flags = np.zeros(nrhs, dtype=np.int32)
        rc = _lib.lib().bdilu_ilu_solve(self._h, _lib.ptr(b), _lib.ptr(x), nrhs, _lib.ptr(flags))
        _lib.check(rc)
        _check_solve_flag(int(rc), self._n)
        return x

    def _destroy(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            _lib.lib().bdilu_ilu_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self._destroy()
        except Exception:
            pass


class LU:
    """PyLU (pyx:157-172) is the exact banded LU (dgbtrf/dgbtrs) -- off the hot path, not built."""

    def __init__(self, cmprs):
        raise NotImplementedError("LU (exact banded solver) is outside the accelerated path; use scipy.linalg.solve_banded")
