"""Device-resident batched API: torch CUDA tensors in, torch CUDA tensors out, no host round trip.

PyTorch is plumbing here (HBM allocations, streams); every operation is one call into the C ABI
(include/bdilu.h) on raw device pointers.  Shapes:

    data  float64 [nsys, ndata]     (float32 tensors take the *_f32 entry points: factor, solve, dot_vec, scale_diag_add)
                                    one packed BlockDiagMatrix::m_data per row (block_diag_ilu.hpp:99-102)
    b, x  float64 [nsys, nrhs, n]   n = nblocks*blockw  ([nsys, n] and [n] are accepted and restored)
    ipiv, rowbycol, colbyrow  int32 [nsys, n];  info int32 [nsys];  flags int32 [nsys, nrhs]
"""
import torch

from . import _lib
from .datastruct import ndata as _ndata


def _stream():
    return torch.cuda.current_stream().cuda_stream


class Shape:
    """nblocks, blockw, ndiag, nsat, ld of one system (ld=None -> tight ld = blockw)."""

    def __init__(self, nblocks, blockw, ndiag, nsat=0, ld=None):
        self.nblocks, self.blockw, self.ndiag, self.nsat = int(nblocks), int(blockw), int(ndiag), int(nsat)
        self.ld = int(ld) if ld else int(blockw)
        self.n = self.nblocks * self.blockw
        self.ndata = _ndata(self.nblocks, self.blockw, self.ndiag, self.nsat, self.ld)

    def dims(self):
        return (self.nblocks, self.blockw, self.ndiag, self.nsat, self.ld)


def _check_data(data, shape):
    if not (isinstance(data, torch.Tensor) and data.is_cuda and data.dtype in (torch.float64, torch.float32)):
        raise TypeError("data must be a CUDA float64 (or float32: the *_f32 entry points) tensor")
    if data.dim() == 1:
        data = data.unsqueeze(0)
    if data.dim() != 2 or data.shape[1] < shape.ndata or data.stride(1) != 1:
        raise ValueError("data must be [nsys, >=ndata] with unit inner stride (ndata=%d)" % shape.ndata)
    return data


def _vec3(v, nsys, n, dtype=torch.float64):
    """-> ([nsys, nrhs, n] contiguous view/copy, restore-shape)"""
    if not (isinstance(v, torch.Tensor) and v.is_cuda and v.dtype == dtype):
        raise TypeError("vectors must be CUDA %s tensors" % str(dtype).replace("torch.", ""))
    shp = v.shape
    if v.dim() == 1:
        v = v.view(1, 1, -1)
    elif v.dim() == 2:
        v = v.unsqueeze(1)
    if v.dim() != 3 or v.shape[0] != nsys or v.shape[2] != n:
        raise ValueError("vector shape %s does not match nsys=%d, n=%d" % (tuple(shp), nsys, n))
    return v.contiguous(), shp


class BatchedILU:
    """ILU_inplace for a batch of systems living in HBM (block_diag_ilu.hpp:426-539).

    The constructor factors (as the reference's does); `data` is overwritten with L\\U and the scaled
    sub-diagonals unless copy=True (the `ILU` owning-wrapper semantics, :541-551).
    Raises RuntimeError("ILU failed!") if any block is singular and check=True (:483-484).
    """

    def __init__(self, data, nblocks, blockw, ndiag, nsat=0, ld=None, copy=False, check=True):
        _lib.require_device()
        self.shape = Shape(nblocks, blockw, ndiag, nsat, ld)
        data = _check_data(data, self.shape)
        self.data = data.clone() if copy else data
        self.nsys = self.data.shape[0]
        dev = self.data.device
        n = self.shape.n
        self.ipiv = torch.empty((self.nsys, n), dtype=torch.int32, device=dev)
        self.rowbycol = torch.empty_like(self.ipiv)
        self.colbyrow = torch.empty_like(self.ipiv)
        self.info = torch.empty((self.nsys,), dtype=torch.int32, device=dev)
        self.flags = None
        factor = _lib.lib().bdilu_factor_f32 if self.data.dtype == torch.float32 else _lib.lib().bdilu_factor
        with torch.cuda.device(dev):
            _lib.check(factor(
                _lib.ptr(self.data), self.data.stride(0), self.nsys, *self.shape.dims(),
                _lib.ptr(self.ipiv), _lib.ptr(self.rowbycol), _lib.ptr(self.colbyrow), _lib.ptr(self.info),
                _stream()))
        if check and bool((self.info != 0).any().item()):
            raise RuntimeError("ILU failed!")

    @classmethod
    def from_scaled(cls, src, scale, diag_add, nblocks, blockw, ndiag, nsat=0, ld=None, out=None, check=True):
        """ILU of scale*src + diag_add*I, assembled inside the factor kernel's load (bdilu_factor_scaled).

        The pair `m.set_to_eye_plus_scaled_mtx(-gamma, J); ILU_inplace(&m)` of an implicit integrator
        (block_diag_ilu.hpp:276-290, 430-485) in one pass over the matrix: `src` (J) is only read, the factors
        go to `out` (a new tensor by default; pass the previous factor storage to re-factor with a new gamma).
        """
        _lib.require_device()
        self = cls.__new__(cls)
        self.shape = Shape(nblocks, blockw, ndiag, nsat, ld)
        src = _check_data(src, self.shape)
        if src.dtype != torch.float64:
            raise TypeError("from_scaled: the fused assembly is float64 only")
        self.data = torch.empty_like(src) if out is None else _check_data(out, self.shape)
        if self.data.shape[0] != src.shape[0]:
            raise ValueError("out must hold as many systems as src")
        self.nsys = src.shape[0]
        dev, n = src.device, self.shape.n
        self.ipiv = torch.empty((self.nsys, n), dtype=torch.int32, device=dev)
        self.rowbycol = torch.empty_like(self.ipiv)
        self.colbyrow = torch.empty_like(self.ipiv)
        self.info = torch.empty((self.nsys,), dtype=torch.int32, device=dev)
        self.flags = None
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().bdilu_factor_scaled(
                _lib.ptr(self.data), self.data.stride(0), _lib.ptr(src), src.stride(0), self.nsys,
                *self.shape.dims(), float(scale), float(diag_add),
                _lib.ptr(self.ipiv), _lib.ptr(self.rowbycol), _lib.ptr(self.colbyrow), _lib.ptr(self.info),
                _stream()))
        if check and bool((self.info != 0).any().item()):
            raise RuntimeError("ILU failed!")
        return self

    def solve(self, b, out=None):
        """x with the shape of b; per-vector reference return codes land in self.flags [nsys, nrhs]."""
        dt = self.data.dtype
        b3, shp = _vec3(b, self.nsys, self.shape.n, dt)
        nrhs = b3.shape[1]
        x3 = torch.empty_like(b3) if out is None else _vec3(out, self.nsys, self.shape.n, dt)[0]
        self.flags = torch.empty((self.nsys, nrhs), dtype=torch.int32, device=b3.device)
        head = (_lib.ptr(self.data), self.data.stride(0), self.nsys, *self.shape.dims(),
                _lib.ptr(self.rowbycol), _lib.ptr(self.colbyrow), _lib.ptr(b3), _lib.ptr(x3),
                nrhs, self.shape.n, self.shape.n * nrhs)
        with torch.cuda.device(b3.device):
            if dt == torch.float32:
                _lib.check(_lib.lib().bdilu_solve_f32(*head, _lib.ptr(self.flags), _stream()))
            else:
                _lib.check(_lib.lib().bdilu_solve(*head, None, _lib.ptr(self.flags), _stream()))
        return x3.view(shp)


def factor_solve(data, b, nblocks, blockw, ndiag, nsat=0, ld=None):
    """Fused ILU(A) + solve(b): factors `data` in place, returns (x, ilu) with ilu.flags / ilu.info set."""
    _lib.require_device()
    ilu = BatchedILU.__new__(BatchedILU)
    ilu.shape = Shape(nblocks, blockw, ndiag, nsat, ld)
    ilu.data = _check_data(data, ilu.shape)
    ilu.nsys = ilu.data.shape[0]
    dev, n = ilu.data.device, ilu.shape.n
    b3, shp = _vec3(b, ilu.nsys, n, ilu.data.dtype)
    nrhs = b3.shape[1]
    ilu.ipiv = torch.empty((ilu.nsys, n), dtype=torch.int32, device=dev)
    ilu.rowbycol = torch.empty_like(ilu.ipiv)
    ilu.colbyrow = torch.empty_like(ilu.ipiv)
    ilu.info = torch.empty((ilu.nsys,), dtype=torch.int32, device=dev)
    ilu.flags = torch.empty((ilu.nsys, nrhs), dtype=torch.int32, device=dev)
    x3 = torch.empty_like(b3)
    fs = _lib.lib().bdilu_factor_solve_f32 if ilu.data.dtype == torch.float32 else _lib.lib().bdilu_factor_solve
    with torch.cuda.device(dev):
        _lib.check(fs(
            _lib.ptr(ilu.data), ilu.data.stride(0), ilu.nsys, *ilu.shape.dims(),
            _lib.ptr(ilu.ipiv), _lib.ptr(ilu.rowbycol), _lib.ptr(ilu.colbyrow), _lib.ptr(ilu.info),
            _lib.ptr(b3), _lib.ptr(x3), nrhs, n, n * nrhs, _lib.ptr(ilu.flags), _stream()))
    return x3.view(shp), ilu


class GraphedFactorSolve:
    """`ILU<double> ilu(A); ilu.solve(b, x)` for small batches, captured once in a CUDA graph.

    A batch that fits one wave (BASELINE configs 1, 2, and one-GPU slabs of config 5) is bound by launch latency: the
    copy of A into the work buffer (the owning-ILU deep copy, block_diag_ilu.hpp:546-547), the factor launch, its
    info memset and the solve launch are recorded once and replayed with a single cudaGraphLaunch.  `a` and `b` are
    the STATIC input tensors (update them in place between replays); results land in `x`, `info`, `flags`.
    """

    def __init__(self, a, b, nblocks, blockw, ndiag, nsat=0, ld=None):
        _lib.require_device()
        self.shape = Shape(nblocks, blockw, ndiag, nsat, ld)
        self.a = _check_data(a, self.shape)
        if self.a.dtype != torch.float64:
            raise TypeError("GraphedFactorSolve is float64 only")
        self.nsys = self.a.shape[0]
        dev, n = self.a.device, self.shape.n
        self.b, self._shp = _vec3(b, self.nsys, n)
        nrhs = self.b.shape[1]
        self.work = torch.empty_like(self.a)
        self.ipiv = torch.empty((self.nsys, n), dtype=torch.int32, device=dev)
        self.rowbycol = torch.empty_like(self.ipiv)
        self.colbyrow = torch.empty_like(self.ipiv)
        self.info = torch.empty((self.nsys,), dtype=torch.int32, device=dev)
        self.flags = torch.empty((self.nsys, nrhs), dtype=torch.int32, device=dev)
        self._x3 = torch.empty_like(self.b)
        self.x = self._x3.view(self._shp)

        def body():
            self.work.copy_(self.a)
            _lib.check(_lib.lib().bdilu_factor_solve(
                _lib.ptr(self.work), self.work.stride(0), self.nsys, *self.shape.dims(),
                _lib.ptr(self.ipiv), _lib.ptr(self.rowbycol), _lib.ptr(self.colbyrow), _lib.ptr(self.info),
                _lib.ptr(self.b), _lib.ptr(self._x3), nrhs, n, n * nrhs, _lib.ptr(self.flags), _stream()))

        with torch.cuda.device(dev):
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                body()  # warm-up outside the capture (function attributes, lazy module load)
            torch.cuda.current_stream().wait_stream(side)
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                body()

    def replay(self):
        self.graph.replay()
        return self.x


def dot_vec(data, vec, nblocks, blockw, ndiag, nsat=0, ld=None, out=None):
    """A @ vec for the UNFACTORED packed batch (BlockDiagMatrix::dot_vec, block_diag_ilu.hpp:295-326)."""
    _lib.require_device()
    shape = Shape(nblocks, blockw, ndiag, nsat, ld)
    data = _check_data(data, shape)
    v3, shp = _vec3(vec, data.shape[0], shape.n, data.dtype)
    out = torch.empty_like(v3) if out is None else _vec3(out, data.shape[0], shape.n, data.dtype)[0]
    mv = _lib.lib().bdilu_dot_vec_f32 if data.dtype == torch.float32 else _lib.lib().bdilu_dot_vec
    with torch.cuda.device(data.device):
        _lib.check(mv(
            _lib.ptr(data), data.stride(0), data.shape[0], *shape.dims(), _lib.ptr(v3), _lib.ptr(out),
            v3.shape[1], shape.n, shape.n * v3.shape[1], _stream()))
    return out.view(shp)


def residual_norm2(data, b, x, nblocks, blockw, ndiag, nsat=0, ld=None):
    """||b - A x||_2^2 per (system, vector) -> float64 [nsys, nrhs]; A is the UNFACTORED batch."""
    _lib.require_device()
    shape = Shape(nblocks, blockw, ndiag, nsat, ld)
    data = _check_data(data, shape)
    if data.dtype != torch.float64:
        raise TypeError("residual_norm2 is float64 only")
    b3, _ = _vec3(b, data.shape[0], shape.n)
    x3, _ = _vec3(x, data.shape[0], shape.n)
    out = torch.empty((data.shape[0], b3.shape[1]), dtype=torch.float64, device=data.device)
    with torch.cuda.device(data.device):
        _lib.check(_lib.lib().bdilu_residual_norm2(
            _lib.ptr(data), data.stride(0), data.shape[0], *shape.dims(), _lib.ptr(b3), _lib.ptr(x3),
            b3.shape[1], shape.n, shape.n * b3.shape[1], _lib.ptr(out), _stream()))
    return out


def scale_diag_add(src, scale, diag_add, nblocks, blockw, ndiag, nsat=0, ld=None, out=None):
    """out = scale*src + diag_add*I on the packed batch (block_diag_ilu.hpp:276-287)."""
    _lib.require_device()
    shape = Shape(nblocks, blockw, ndiag, nsat, ld)
    src = _check_data(src, shape)
    dst = torch.empty_like(src) if out is None else _check_data(out, shape)
    sda = _lib.lib().bdilu_scale_diag_add_f32 if src.dtype == torch.float32 else _lib.lib().bdilu_scale_diag_add
    with torch.cuda.device(src.device):
        _lib.check(sda(
            _lib.ptr(dst), _lib.ptr(src), src.stride(0), src.shape[0], *shape.dims(),
            float(scale), float(diag_add), _stream()))
    return dst
