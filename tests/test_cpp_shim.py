"""The C++ header shim (include/block_diag_ilu_b200.hpp): compiles against the C ABI with plain g++, and
tests/cpp/test_shim.cpp -- the reference's tests/test_block_diag.cpp cases restated -- passes on a GPU.
Without a device the same binary must report that every compute call threw (exit code 3): no CPU fallback."""
import os
import subprocess

import pytest

from block_diag_ilu_b200 import _build, _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _compile(tmpdir):
    libdir = os.path.dirname(_build.build())
    exe = os.path.join(str(tmpdir), "test_shim")
    cmd = ["g++", "-std=c++14", "-O1", "-Wall", "-I" + os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", "test_shim.cpp"), "-L" + libdir, "-lbdilu_b200",
           "-Wl,-rpath," + libdir, "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "warning" not in r.stderr
    return exe


def test_shim_compiles_and_fails_loudly_without_device(tmp_path):
    exe = _compile(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    if _lib.lib().bdilu_device_count() > 0:
        assert r.returncode == 0, r.stdout
    else:
        assert r.returncode == 3, r.stdout
        assert "no CUDA device" in r.stdout


@pytest.mark.gpu
def test_shim_on_gpu(tmp_path):
    exe = _compile(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "shim ok" in r.stdout
