"""bdilu_residual_norm2_static (include/bdilu.h): the residual norms with their scratch in a file-scope device array,
for callers that may not allocate.  No counterpart in the reference; checked against the allocating entry point
bdilu_residual_norm2 (itself checked against the oracle in test_gpu_parity.py).  Collected last on purpose."""

import numpy as np
import pytest

from block_diag_ilu_b200 import _lib, synth

pytestmark = pytest.mark.gpu

STATIC_PARTIALS = 131072  # BDILU_STATIC_NORM2_PARTIALS


def _args(L, dd, bb, xx, out, nsys, N, w, d, nrhs):
    n = N * w
    return (_lib.ptr(dd), dd.stride(0), nsys, N, w, d, 0, 0, _lib.ptr(bb), _lib.ptr(xx), nrhs, n, n * nrhs,
            _lib.ptr(out), None)


@pytest.mark.parametrize("shape", [(3, 40, 32, 2, 1), (2, 17, 8, 1, 5), (2, 9, 64, 3, 9)])
def test_static_scratch_matches_allocating_entry_point(shape):
    import torch
    nsys, N, w, d, nrhs = shape
    L = _lib.lib()
    dev = torch.device("cuda:0")
    for call in range(2):  # two calls in a row share the array: the second sees the first one's leftovers
        data, _ = synth.make_batch(nsys, N, w, d, "D", seed0=900 + call)
        rng = np.random.default_rng(77 + call)
        b = rng.standard_normal((nsys, nrhs, N * w))
        x = rng.standard_normal((nsys, nrhs, N * w))
        dd, bb, xx = (torch.from_numpy(a).to(dev) for a in (data, b, x))
        ref = torch.full((nsys, nrhs), -1.0, dtype=torch.float64, device=dev)
        got = torch.full((nsys, nrhs), -2.0, dtype=torch.float64, device=dev)
        with torch.cuda.device(dev):
            _lib.check(L.bdilu_residual_norm2(*_args(L, dd, bb, xx, ref, nsys, N, w, d, nrhs)))
            torch.cuda.synchronize()
            _lib.check(L.bdilu_residual_norm2_static(*_args(L, dd, bb, xx, got, nsys, N, w, d, nrhs)))
            torch.cuda.synchronize()
        r, g = ref.cpu().numpy(), got.cpu().numpy()
        assert np.all(r > 0)
        # same kernels, same summation order: bit-identical (the suite's tolerance for norms is 1e-13 relative)
        assert np.array_equal(r, g), (call, np.abs(r - g).max())
        np.testing.assert_allclose(g, r, rtol=1e-13, atol=0)


def test_static_scratch_rejects_sizes_beyond_the_array():
    import torch
    L = _lib.lib()
    dev = torch.device("cuda:0")
    # nsys*nrhs*ceil(n/128) = 1 * 513 * 256 = 131328 > 131072; the check runs before anything touches the buffers
    N, w, d, nrhs = 1024, 32, 1, 513
    ndata = L.bdilu_ndata(N, w, d, 0, 0)
    dd = torch.zeros((1, ndata), dtype=torch.float64, device=dev)
    vv = torch.zeros((1, 1, N * w), dtype=torch.float64, device=dev)  # never dereferenced
    out = torch.zeros((1, nrhs), dtype=torch.float64, device=dev)
    assert 1 * nrhs * ((N * w + 127) // 128) > STATIC_PARTIALS
    rc = L.bdilu_residual_norm2_static(_lib.ptr(dd), dd.stride(0), 1, N, w, d, 0, 0, _lib.ptr(vv), _lib.ptr(vv), nrhs,
                                       N * w, N * w * nrhs, _lib.ptr(out), None)
    assert rc == _lib.ERR_UNSUPPORTED
    msg = L.bdilu_last_error().decode()
    assert "static scratch" in msg and str(STATIC_PARTIALS) in msg
    # right at the limit is accepted: 512 vectors * 256 CTAs = 131072
    nrhs = 512
    b = torch.ones((1, nrhs, N * w), dtype=torch.float64, device=dev)
    x = torch.zeros((1, nrhs, N * w), dtype=torch.float64, device=dev)
    out = torch.zeros((1, nrhs), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(L.bdilu_residual_norm2_static(_lib.ptr(dd), dd.stride(0), 1, N, w, d, 0, 0, _lib.ptr(b), _lib.ptr(x),
                                                 nrhs, N * w, N * w * nrhs, _lib.ptr(out), None))
        torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), np.full((1, nrhs), float(N * w)))
