"""ctypes front-end to the parity oracle (TEST INFRASTRUCTURE, not product code).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  Two back ends with one interface:

  Port : oracle/_build/libbdilu_oracle.so  -- the C restatement (bdilu_oracle.c)
  Ref  : oracle/_ref/libbdilu_ref.so       -- the untouched reference header + LAPACK dgetrf

Both factor IN PLACE on a float64 numpy buffer laid out as BlockDiagMatrix::m_data
(block_diag_ilu.hpp:99-102) and return (ipiv, rowbycol, colbyrow, info).
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_ip = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")


def _opt(arr):
    return arr.ctypes.data_as(C.c_void_p) if arr is not None else None


class Port:
    kind = "port"

    def __init__(self, path=None):
        path = path or os.path.join(HERE, "_build", "libbdilu_oracle.so")
        if not os.path.exists(path):
            from . import build_oracle
            path = build_oracle.build_port(verbose=False)
        L = self.lib = C.CDLL(path)
        L.oracle_ndata.restype = C.c_int64
        L.oracle_ndata.argtypes = [C.c_int] * 5
        L.oracle_default_ld.argtypes = [C.c_int]
        L.oracle_getrf.argtypes = [C.c_int, _dp, C.c_int, _ip]
        L.oracle_factor.argtypes = [_dp] + [C.c_int] * 5 + [_ip, _ip, _ip]
        L.oracle_solve.argtypes = [_dp] + [C.c_int] * 5 + [_ip, _ip, _dp, _dp, _dp]
        L.oracle_dot_vec.argtypes = [_dp] + [C.c_int] * 5 + [_dp, _dp]
        L.oracle_scale_diag_add.argtypes = [_dp, _dp] + [C.c_int] * 5 + [C.c_double, C.c_double]
        L.oracle_rowpiv2rowbycol.argtypes = [C.c_int, _ip, _ip]
        L.oracle_rowbycol2colbyrow.argtypes = [C.c_int, _ip, _ip]
        L.oracle_factor_solve_batch.restype = C.c_double
        L.oracle_factor_solve_batch.argtypes = ([_dp, C.c_int64, C.c_int] + [C.c_int] * 5 +
                                                [_ip, _ip, _ip, _dp, _dp, C.c_int, _ip, C.c_int])

    def ndata(self, nblocks, blockw, ndiag, nsat=0, ld=0):
        ld = ld or self.default_ld(blockw)
        return int(self.lib.oracle_ndata(nblocks, blockw, ndiag, nsat, ld))

    def default_ld(self, blockw):
        return int(self.lib.oracle_default_ld(blockw))

    def getrf(self, a, n, lda):
        ipiv = np.zeros(n, np.int32)
        info = self.lib.oracle_getrf(n, a, lda, ipiv)
        return ipiv, info

    def factor(self, data, nblocks, blockw, ndiag, nsat, ld):
        n = nblocks * blockw
        ipiv, rbc, cbr = (np.zeros(n, np.int32) for _ in range(3))
        info = self.lib.oracle_factor(data, nblocks, blockw, ndiag, nsat, ld, ipiv, rbc, cbr)
        return ipiv, rbc, cbr, int(info)

    def solve(self, data, nblocks, blockw, ndiag, nsat, ld, rowbycol, colbyrow, b):
        b = np.ascontiguousarray(b, np.float64)
        x = np.zeros_like(b)
        y = np.zeros_like(b)
        flag = self.lib.oracle_solve(data, nblocks, blockw, ndiag, nsat, ld, rowbycol, colbyrow, b, x, y)
        return x, int(flag)

    def dot_vec(self, data, nblocks, blockw, ndiag, nsat, ld, vec):
        vec = np.ascontiguousarray(vec, np.float64)
        out = np.full_like(vec, np.nan)
        self.lib.oracle_dot_vec(data, nblocks, blockw, ndiag, nsat, ld, vec, out)
        return out

    def scale_diag_add(self, src, nblocks, blockw, ndiag, nsat, ld, scale, diag_add):
        dst = np.empty_like(src)
        self.lib.oracle_scale_diag_add(dst, src, nblocks, blockw, ndiag, nsat, ld, scale, diag_add)
        return dst

    def rowpiv2rowbycol(self, piv):
        piv = np.ascontiguousarray(piv, np.int32)
        out = np.zeros_like(piv)
        self.lib.oracle_rowpiv2rowbycol(piv.size, piv, out)
        return out

    def rowbycol2colbyrow(self, rbc):
        rbc = np.ascontiguousarray(rbc, np.int32)
        out = np.zeros_like(rbc)
        self.lib.oracle_rowbycol2colbyrow(rbc.size, rbc, out)
        return out

    def factor_solve_batch(self, data, nsys, nblocks, blockw, ndiag, nsat, ld, b, nrhs=1, nthreads=0, mode=1):
        """data: [nsys, ndata] (factored in place); b: [nsys, nrhs, n].  Returns dict."""
        n = nblocks * blockw
        data2 = data.reshape(nsys, -1)
        ipiv, rbc, cbr = (np.zeros((nsys, n), np.int32) for _ in range(3))
        b = np.ascontiguousarray(b, np.float64).reshape(nsys, nrhs, n)
        x = np.zeros_like(b)
        infos = np.zeros(nsys, np.int32)
        secs = self.lib.oracle_factor_solve_batch(data2.reshape(-1), data2.shape[1], nsys, nblocks, blockw, ndiag,
                                                  nsat, ld, ipiv.reshape(-1), rbc.reshape(-1), cbr.reshape(-1),
                                                  b.reshape(-1), x.reshape(-1), nrhs, infos, nthreads)
        return dict(seconds=float(secs), ipiv=ipiv, rowbycol=rbc, colbyrow=cbr, x=x, infos=infos)


class Ref:
    """The reference itself.  `factor` keeps the reference's ILU_inplace object alive for `solve`."""
    kind = "reference"

    def __init__(self, path=None):
        path = path or os.path.join(HERE, "_ref", "libbdilu_ref.so")
        if not os.path.exists(path):
            from . import build_oracle
            path = build_oracle.build_ref(verbose=False)
            if path is None:
                raise FileNotFoundError("oracle/_ref/libbdilu_ref.so missing and /root/reference absent")
        L = self.lib = C.CDLL(path)
        L.ref_ndata.restype = C.c_int64
        L.ref_ndata.argtypes = [C.c_int] * 5
        L.ref_default_ld.argtypes = [C.c_int]
        L.ref_ilu_create.restype = C.c_void_p
        L.ref_ilu_create.argtypes = [_dp] + [C.c_int] * 5 + [_ip, _ip, _ip, C.POINTER(C.c_int)]
        L.ref_ilu_solve.argtypes = [C.c_void_p, _dp, _dp]
        L.ref_ilu_destroy.argtypes = [C.c_void_p]
        L.ref_dot_vec.argtypes = [_dp] + [C.c_int] * 5 + [_dp, _dp]
        L.ref_scale_diag_add.argtypes = [_dp, _dp] + [C.c_int] * 5 + [C.c_double, C.c_double]
        L.ref_rowpiv2rowbycol.argtypes = [C.c_int, _ip, _ip]
        L.ref_rowbycol2colbyrow.argtypes = [C.c_int, _ip, _ip]
        L.ref_set_factor_threads.argtypes = [C.c_int]
        L.ref_factor_solve_batch.restype = C.c_double
        L.ref_factor_solve_batch.argtypes = ([_dp, C.c_int64, C.c_int] + [C.c_int] * 5 +
                                             [C.c_void_p, C.c_void_p, C.c_void_p, _dp, _dp, C.c_int,
                                              _ip, C.c_int, C.c_int])
        self._handles = {}

    def max_threads(self):
        return int(self.lib.ref_max_threads())

    def ndata(self, nblocks, blockw, ndiag, nsat=0, ld=0):
        return int(self.lib.ref_ndata(nblocks, blockw, ndiag, nsat, ld))

    def default_ld(self, blockw):
        return int(self.lib.ref_default_ld(blockw))

    def factor(self, data, nblocks, blockw, ndiag, nsat, ld):
        """info is 0, or -1 when the reference threw "ILU failed!" (block_diag_ilu.hpp:483-484)."""
        n = nblocks * blockw
        ipiv, rbc, cbr = (np.zeros(n, np.int32) for _ in range(3))
        threw = C.c_int(0)
        h = self.lib.ref_ilu_create(data, nblocks, blockw, ndiag, nsat, ld, ipiv, rbc, cbr, C.byref(threw))
        key = data.ctypes.data
        old = self._handles.pop(key, None)
        if old:
            self.lib.ref_ilu_destroy(old[0])
        if h:
            self._handles[key] = (h, data)  # keep the buffer alive: ILU_inplace holds a raw pointer
        return ipiv, rbc, cbr, (-1 if threw.value else 0)

    def solve(self, data, nblocks, blockw, ndiag, nsat, ld, rowbycol, colbyrow, b):
        h = self._handles[data.ctypes.data][0]
        b = np.ascontiguousarray(b, np.float64)
        x = np.zeros_like(b)
        flag = self.lib.ref_ilu_solve(h, b, x)
        return x, int(flag)

    def dot_vec(self, data, nblocks, blockw, ndiag, nsat, ld, vec):
        vec = np.ascontiguousarray(vec, np.float64)
        out = np.full_like(vec, np.nan)
        self.lib.ref_dot_vec(data, nblocks, blockw, ndiag, nsat, ld, vec, out)
        return out

    def scale_diag_add(self, src, nblocks, blockw, ndiag, nsat, ld, scale, diag_add):
        dst = np.empty_like(src)
        self.lib.ref_scale_diag_add(dst, src, nblocks, blockw, ndiag, nsat, ld, scale, diag_add)
        return dst

    def rowpiv2rowbycol(self, piv):
        piv = np.ascontiguousarray(piv, np.int32)
        out = np.zeros_like(piv)
        self.lib.ref_rowpiv2rowbycol(piv.size, piv, out)
        return out

    def rowbycol2colbyrow(self, rbc):
        rbc = np.ascontiguousarray(rbc, np.int32)
        out = np.zeros_like(rbc)
        self.lib.ref_rowbycol2colbyrow(rbc.size, rbc, out)
        return out

    def release(self):
        for h, _ in self._handles.values():
            self.lib.ref_ilu_destroy(h)
        self._handles.clear()

    def factor_solve_batch(self, data, nsys, nblocks, blockw, ndiag, nsat, ld, b, nrhs=1, nthreads=0, mode=1):
        n = nblocks * blockw
        data2 = data.reshape(nsys, -1)
        ipiv, rbc, cbr = (np.zeros((nsys, n), np.int32) for _ in range(3))
        b = np.ascontiguousarray(b, np.float64).reshape(nsys, nrhs, n)
        x = np.zeros_like(b)
        infos = np.zeros(nsys, np.int32)
        secs = self.lib.ref_factor_solve_batch(data2.reshape(-1), data2.shape[1], nsys, nblocks, blockw, ndiag,
                                               nsat, ld, _opt(ipiv), _opt(rbc), _opt(cbr),
                                               b.reshape(-1), x.reshape(-1), nrhs, infos, mode, nthreads)
        return dict(seconds=float(secs), ipiv=ipiv, rowbycol=rbc, colbyrow=cbr, x=x, infos=infos)


class RefF32:
    """The reference header instantiated with Real_t = float (sgetrf).  Same interface as Ref on float32 buffers."""
    kind = "reference-f32"

    def __init__(self, path=None):
        path = path or os.path.join(HERE, "_ref", "libbdilu_ref.so")
        if not os.path.exists(path):
            from . import build_oracle
            path = build_oracle.build_ref(verbose=False)
            if path is None:
                raise FileNotFoundError("oracle/_ref/libbdilu_ref.so missing and /root/reference absent")
        fp = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
        L = self.lib = C.CDLL(path)
        L.ref_ndata_f32.restype = C.c_int64
        L.ref_ndata_f32.argtypes = [C.c_int] * 5
        L.ref_default_ld_f32.argtypes = [C.c_int]
        L.ref_ilu_create_f32.restype = C.c_void_p
        L.ref_ilu_create_f32.argtypes = [fp] + [C.c_int] * 5 + [_ip, _ip, _ip, C.POINTER(C.c_int)]
        L.ref_ilu_solve_f32.argtypes = [C.c_void_p, fp, fp]
        L.ref_ilu_destroy_f32.argtypes = [C.c_void_p]
        L.ref_dot_vec_f32.argtypes = [fp] + [C.c_int] * 5 + [fp, fp]
        L.ref_scale_diag_add_f32.argtypes = [fp, fp] + [C.c_int] * 5 + [C.c_float, C.c_float]
        self._handles = {}

    def ndata(self, nblocks, blockw, ndiag, nsat=0, ld=0):
        return int(self.lib.ref_ndata_f32(nblocks, blockw, ndiag, nsat, ld))

    def default_ld(self, blockw):
        return int(self.lib.ref_default_ld_f32(blockw))

    def factor(self, data, nblocks, blockw, ndiag, nsat, ld):
        assert data.dtype == np.float32
        n = nblocks * blockw
        ipiv, rbc, cbr = (np.zeros(n, np.int32) for _ in range(3))
        threw = C.c_int(0)
        h = self.lib.ref_ilu_create_f32(data, nblocks, blockw, ndiag, nsat, ld, ipiv, rbc, cbr, C.byref(threw))
        key = data.ctypes.data
        old = self._handles.pop(key, None)
        if old:
            self.lib.ref_ilu_destroy_f32(old[0])
        if h:
            self._handles[key] = (h, data)
        return ipiv, rbc, cbr, (-1 if threw.value else 0)

    def solve(self, data, nblocks, blockw, ndiag, nsat, ld, rowbycol, colbyrow, b):
        h = self._handles[data.ctypes.data][0]
        b = np.ascontiguousarray(b, np.float32)
        x = np.zeros_like(b)
        flag = self.lib.ref_ilu_solve_f32(h, b, x)
        return x, int(flag)

    def dot_vec(self, data, nblocks, blockw, ndiag, nsat, ld, vec):
        vec = np.ascontiguousarray(vec, np.float32)
        out = np.full_like(vec, np.nan)
        self.lib.ref_dot_vec_f32(data, nblocks, blockw, ndiag, nsat, ld, vec, out)
        return out

    def scale_diag_add(self, src, nblocks, blockw, ndiag, nsat, ld, scale, diag_add):
        dst = np.empty_like(src)
        self.lib.ref_scale_diag_add_f32(dst, src, nblocks, blockw, ndiag, nsat, ld, scale, diag_add)
        return dst

    def release(self):
        for h, _ in self._handles.values():
            self.lib.ref_ilu_destroy_f32(h)
        self._handles.clear()


def best():
    """The reference when its compiled shim is available (here, or prebuilt on the GPU box), else the port."""
    try:
        return Ref()
    except (OSError, FileNotFoundError):
        return Port()
