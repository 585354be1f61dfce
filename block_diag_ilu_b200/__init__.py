"""block_diag_ilu_b200 -- B200-native drop-in for the hot path of chemreac/block_diag_ilu.

Public names follow block_diag_ilu/__init__.py:7-19 of the reference; `device` adds the batched,
device-resident API (torch CUDA tensors).  All compute goes through libbdilu_b200.so (include/bdilu.h).
"""
import os

from .compressed import Compressed, Compressed_from_dense, Compressed_from_data, ILU, LU
from .datastruct import diag_data_len, alloc_compressed, get_compressed

__version__ = "0.1.0"


def get_include():
    """Directory holding bdilu.h and the C++ header shim block_diag_ilu_b200.hpp."""
    return os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include")
