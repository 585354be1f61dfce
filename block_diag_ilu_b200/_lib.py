"""ctypes binding of libbdilu_b200.so -- the ONLY compute back end of this package.

There is no CPU fallback: if the library is missing it is built in-tree with nvcc; if that fails, or
there is no CUDA device when a compute entry point is called, the call raises.
"""
import ctypes as C
import os

from . import _build

ERR_ARG, ERR_UNSUPPORTED, ERR_CUDA = -1, -2, -3

_i, _i64, _d, _p = C.c_int, C.c_int64, C.c_double, C.c_void_p
_DIMS = [_i] * 5  # nblocks, blockw, ndiag, nsat, ld

_SIGNATURES = {
    "bdilu_version": (_i, []),
    "bdilu_last_error": (C.c_char_p, []),
    "bdilu_device_count": (_i, []),
    "bdilu_default_ld": (_i, [_i]),
    "bdilu_ndata": (_i64, _DIMS),
    "bdilu_factor": (_i, [_p, _i64, _i] + _DIMS + [_p, _p, _p, _p, _p]),
    "bdilu_solve": (_i, [_p, _i64, _i] + _DIMS + [_p, _p, _p, _p, _i, _i64, _i64, _p, _p, _p]),
    "bdilu_factor_solve": (_i, [_p, _i64, _i] + _DIMS + [_p, _p, _p, _p, _p, _p, _i, _i64, _i64, _p, _p]),
    "bdilu_dot_vec": (_i, [_p, _i64, _i] + _DIMS + [_p, _p, _i, _i64, _i64, _p]),
    "bdilu_residual_norm2": (_i, [_p, _i64, _i] + _DIMS + [_p, _p, _i, _i64, _i64, _p, _p]),
    "bdilu_residual_norm2_static": (_i, [_p, _i64, _i] + _DIMS + [_p, _p, _i, _i64, _i64, _p, _p]),
    "bdilu_scale_diag_add": (_i, [_p, _p, _i64, _i] + _DIMS + [_d, _d, _p]),
    "bdilu_factor_scaled": (_i, [_p, _i64, _p, _i64, _i] + _DIMS + [_d, _d, _p, _p, _p, _p, _p]),
    "bdilu_workspace_bytes": (C.c_size_t, [_i] + _DIMS + [_i]),
    "bdilu_ilu_create": (_i, [C.POINTER(_p), _p, _i64, _i] + _DIMS + [_i, _p, _p, _p, _p]),
    "bdilu_ilu_solve": (_i, [_p, _p, _p, _i, _p]),
    "bdilu_ilu_destroy": (None, [_p]),
    "bdilu_host_dot_vec": (_i, [_p] + _DIMS + [_p, _p, _i]),
    "bdilu_host_scale_diag_add": (_i, [_p, _p] + _DIMS + [_d, _d]),
    "bdilu_host_factor_solve": (_i, [_p, _i64, _i] + _DIMS + [_p, _p, _i, _p, _p]),
    "bdilu_default_ld_f32": (_i, [_i]),
    "bdilu_factor_f32": (_i, [_p, _i64, _i] + _DIMS + [_p, _p, _p, _p, _p]),
    "bdilu_solve_f32": (_i, [_p, _i64, _i] + _DIMS + [_p, _p, _p, _p, _i, _i64, _i64, _p, _p]),
    "bdilu_factor_solve_f32": (_i, [_p, _i64, _i] + _DIMS + [_p, _p, _p, _p, _p, _p, _i, _i64, _i64, _p, _p]),
    "bdilu_dot_vec_f32": (_i, [_p, _i64, _i] + _DIMS + [_p, _p, _i, _i64, _i64, _p]),
    "bdilu_scale_diag_add_f32": (_i, [_p, _p, _i64, _i] + _DIMS + [C.c_float, C.c_float, _p]),
    "bdilu_ilu_create_f32": (_i, [C.POINTER(_p), _p, _i64, _i] + _DIMS + [_i, _p, _p, _p, _p]),
    "bdilu_ilu_solve_f32": (_i, [_p, _p, _p, _i, _p]),
    "bdilu_ilu_destroy_f32": (None, [_p]),
    "bdilu_host_dot_vec_f32": (_i, [_p] + _DIMS + [_p, _p, _i]),
    "bdilu_host_scale_diag_add_f32": (_i, [_p, _p] + _DIMS + [C.c_float, C.c_float]),
    "bdilu_launch_count": (_i64, []),
    "bdilu_launch_count_reset": (None, []),
}

# include/bdilu_internal.h: test and measurement hooks, not part of the drop-in boundary
_INTERNAL = {
    "bdilu_debug_force_chain_solve": (_i, [_i]),
    "bdilu_debug_skew_stats": (_i, [_p, _i]),
}

EXPORTED = tuple(_SIGNATURES)
EXPORTED_INTERNAL = tuple(_INTERNAL)
_lib = None


class BdiluError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libbdilu_b200 error %d: %s" % (code, msg))
        self.code = code


def lib():
    """Load (building first if needed) the shared library.  Raises if it cannot be had."""
    global _lib
    if _lib is None:
        path = os.environ.get("BDILU_LIB_PATH") or _build.build()  # override: instrumented debug builds
        L = C.CDLL(path)
        for name, (res, args) in list(_SIGNATURES.items()) + list(_INTERNAL.items()):
            fn = getattr(L, name)  # AttributeError here == ABI drift: fail loudly
            fn.restype = res
            fn.argtypes = args
        L._path = path
        _lib = L
    return _lib


def check(rc):
    if rc < 0:
        raise BdiluError(rc, lib().bdilu_last_error().decode())
    return rc


def require_device():
    if lib().bdilu_device_count() < 1:
        raise BdiluError(ERR_CUDA, "no CUDA device visible: block_diag_ilu_b200 has no CPU fallback")


def ptr(a):
    """Raw address of a numpy array or torch tensor (or None)."""
    if a is None:
        return None
    if hasattr(a, "data_ptr"):
        return C.c_void_p(a.data_ptr())
    return C.c_void_p(a.ctypes.data)
