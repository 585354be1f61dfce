"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/bdilu.h
declares, validates arguments before touching CUDA, and fails loudly (no CPU fallback) without a device.
No compute call is expected to succeed here."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from block_diag_ilu_b200 import _lib, datastruct as ds

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "bdilu.h")


INTERNAL_HEADER = os.path.join(ROOT, "include", "bdilu_internal.h")


def declared_symbols(header=HEADER):
    src = open(header).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"BDILU_API\s+[^;(]*?\b(bdilu_\w+)\s*\(", src)))


def test_header_declares_what_the_binding_binds():
    assert set(declared_symbols()) == set(_lib.EXPORTED)
    # test hooks live in their own header, outside the drop-in boundary
    assert set(declared_symbols(INTERNAL_HEADER)) == set(_lib.EXPORTED_INTERNAL)
    assert not any("debug" in s for s in declared_symbols())


def test_library_exports_every_declared_symbol():
    L = _lib.lib()
    for name in declared_symbols():
        assert hasattr(L, name), name
    out = subprocess.check_output(["nm", "-D", "--defined-only", L._path]).decode()
    exported = {ln.split()[-1] for ln in out.splitlines() if " T " in ln}
    assert set(declared_symbols()) <= exported
    # nothing but the C ABI leaks out (kernels and helpers are hidden)
    assert exported == set(declared_symbols()) | set(declared_symbols(INTERNAL_HEADER)), \
        sorted(exported ^ (set(declared_symbols()) | set(declared_symbols(INTERNAL_HEADER))))[:5]


def test_header_compiles_as_plain_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "bdilu.h"\nint main(void){return bdilu_version() ? 0 : 1;}\n')
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", str(src),
                           "-o", str(tmp_path / "t.o")])


def test_layout_helpers_match_reference_arithmetic(port):
    L = _lib.lib()
    assert L.bdilu_version() >= 100
    for w in (1, 2, 3, 7, 8, 9, 16, 31, 32, 33, 64):
        assert L.bdilu_default_ld(w) == ds.default_ld(w) == port.default_ld(w)
    for (N, w, d, nsat, ld) in [(3, 2, 1, 0, 6), (64, 4, 1, 0, 4), (8192, 32, 2, 0, 32), (2048, 64, 3, 0, 64),
                                (512, 8, 2, 0, 8), (9, 5, 3, 2, 8), (3, 2, 0, 2, 0)]:
        assert L.bdilu_ndata(N, w, d, nsat, ld) == ds.ndata(N, w, d, nsat, ld) == port.ndata(N, w, d, nsat, ld)
    # SURVEY.md section 8 table
    assert L.bdilu_ndata(8192, 32, 2, 0, 32) == 9_436_992
    assert L.bdilu_ndata(512, 8, 2, 0, 8) == 49_104


def test_argument_errors_come_before_any_cuda_call():
    L = _lib.lib()
    buf = np.zeros(64)
    idx = np.zeros(64, np.int32)
    P = _lib.ptr
    args = (P(idx), P(idx), P(idx), P(idx), None)
    assert L.bdilu_factor(P(buf), 0, 1, 2, 4, 1, 0, 3, *args) == _lib.ERR_ARG          # ld < blockw
    assert b"ld" in L.bdilu_last_error()
    assert L.bdilu_factor(P(buf), 0, 1, 2, 4, 2, 0, 4, *args) == _lib.ERR_ARG          # ndiag > nblocks-1
    assert L.bdilu_factor(P(buf), 0, 1, 2, 65, 1, 0, 65, *args) == _lib.ERR_UNSUPPORTED  # blockw > 64
    assert L.bdilu_factor(P(buf), 8, 2, 2, 2, 1, 0, 2, *args) == _lib.ERR_ARG          # sys_stride < ndata
    assert L.bdilu_factor(None, 0, 1, 2, 2, 1, 0, 2, *args) == _lib.ERR_ARG            # null data
    assert L.bdilu_solve(P(buf), 0, 1, 2, 2, 1, 0, 2, P(idx), P(idx), P(buf), P(buf), 1, 4, 4, None, None,
                         None) == _lib.ERR_ARG                                         # b aliases x
    assert L.bdilu_factor(P(buf), 0, 0, 2, 2, 1, 0, 2, *args) == 0                     # empty batch is a no-op


def test_no_cpu_fallback_without_a_device():
    L = _lib.lib()
    if L.bdilu_device_count() > 0:
        pytest.skip("a CUDA device is visible")
    with pytest.raises(_lib.BdiluError):
        _lib.require_device()
    data = np.ones(12 + 8)
    out = C.c_void_p()
    rc = L.bdilu_ilu_create(C.byref(out), _lib.ptr(data), 0, 1, 3, 2, 1, 0, 2, 1, None, None, None, None)
    assert rc == _lib.ERR_CUDA and not out.value
    x = np.zeros(6)
    rc = L.bdilu_host_dot_vec(_lib.ptr(data), 3, 2, 1, 0, 2, _lib.ptr(np.ones(6)), _lib.ptr(x), 1)
    assert rc == _lib.ERR_CUDA and not x.any()


def test_product_never_touches_the_oracle():
    """The oracle is test infrastructure: nothing under block_diag_ilu_b200/ or include/ may reference it."""
    bad = []
    for base in ("block_diag_ilu_b200", "include"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, base)):
            if "build" in dirpath.split(os.sep) or "__pycache__" in dirpath:
                continue
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                    txt = open(os.path.join(dirpath, f), errors="ignore").read()
                    if re.search(r"\boracle\b", txt) and "oracle" not in f:
                        bad.append(os.path.join(dirpath, f))
    assert not bad, bad
