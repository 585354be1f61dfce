"""Seeded synthetic inputs in the packed BlockDiagMatrix layout.

The reference ships no data files; its demo scripts draw uniform(-1,1) blocks with a boosted
diagonal and uniform(-1,1)*off_diag_factor couplings (scripts/ilu_vs_blocks.py:24-56,
python_prototype/demo.py:35-60).  Two families:

  "D"  dominant : blk ~ U(-1,1), blk[c,c] += sign(blk[c,c])*w ; sub,sup ~ 1e-3*U(-1,1)  (timed + parity)
  "P"  pivoting : as D without the diagonal boost (about 87 % non-identity pivots at w=32; parity only)

Layout follows block_diag_ilu.hpp:99-102 (blocks | sub | sup | bot | top), ld >= blockw.
"""
import numpy as np

from .datastruct import ndata as _ndata


def make_system(nblocks, blockw, ndiag, family="D", seed=0, ld=None, nsat=0, coupling=1e-3):
    """One packed system as float64 numpy [ndata] plus b [nblocks*blockw]."""
    ld = ld or blockw
    rng = np.random.default_rng(seed)
    w = blockw
    blk = rng.uniform(-1.0, 1.0, size=(nblocks, w, w))  # [b, c, r]: column-major blocks
    if family == "D":
        idx = np.arange(w)
        d = blk[:, idx, idx]
        blk[:, idx, idx] = d + np.where(d >= 0, 1.0, -1.0) * w
    elif family != "P":
        raise ValueError("family must be 'D' or 'P'")
    n_all = _ndata(nblocks, blockw, ndiag, nsat, ld)
    data = np.zeros(n_all)
    nblk = nblocks * w * ld
    data[:nblk].reshape(nblocks, w, ld)[:, :, :w] = blk
    rest = data[nblk:].reshape(-1, ld)
    rest[:, :w] = coupling * rng.uniform(-1.0, 1.0, size=(rest.shape[0], w))
    b = rng.uniform(-1.0, 1.0, size=nblocks * w)
    return data, b


def make_batch(nsys, nblocks, blockw, ndiag, family="D", seed0=0, ld=None, nrhs=1, nsat=0):
    """[nsys, ndata] data and [nsys, nrhs, n] right-hand sides; system s uses seed0 + s."""
    ld = ld or blockw
    nd = _ndata(nblocks, blockw, ndiag, nsat, ld)
    data = np.empty((nsys, nd))
    b = np.empty((nsys, nrhs, nblocks * blockw))
    for s in range(nsys):
        data[s], b0 = make_system(nblocks, blockw, ndiag, family, seed0 + s, ld, nsat)
        b[s, 0] = b0
        if nrhs > 1:
            b[s, 1:] = np.random.default_rng(10_000_019 * (seed0 + s) + 7).uniform(
                -1, 1, size=(nrhs - 1, nblocks * blockw))
    return data, b


def make_batch_device(nsys, nblocks, blockw, ndiag, device, seed=0, nrhs=1, family="D"):
    """Same recipe generated directly in HBM with torch (ld = blockw, nsat = 0) for the multi-GB
    bench batches.  Returns (data [nsys, ndata], b [nsys, nrhs, n]) float64 tensors."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    w = blockw
    nd = _ndata(nblocks, blockw, ndiag, 0, w)
    data = torch.empty((nsys, nd), dtype=torch.float64, device=device)
    nblk = nblocks * w * w
    data[:, :nblk].uniform_(-1.0, 1.0, generator=g)
    if family == "D":
        blocks = data[:, :nblk].view(nsys, nblocks, w, w)
        diag = torch.diagonal(blocks, dim1=2, dim2=3)
        diag.add_(torch.where(diag >= 0, 1.0, -1.0) * w)
    data[:, nblk:].uniform_(-1e-3, 1e-3, generator=g)
    b = torch.empty((nsys, nrhs, nblocks * w), dtype=torch.float64, device=device)
    b.uniform_(-1.0, 1.0, generator=g)
    return data, b
