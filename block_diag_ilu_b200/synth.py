"""Seeded synthetic inputs in the packed BlockDiagMatrix layout.

The reference ships no data files; its demo scripts draw uniform(-1,1) blocks with a boosted
diagonal and uniform(-1,1)*off_diag_factor couplings (scripts/ilu_vs_blocks.py:24-56,
python_prototype/demo.py:35-60).  Two families:

  "D"  dominant : blk ~ U(-1,1), blk[c,c] += sign(blk[c,c])*w ; sub,sup ~ 1e-3*U(-1,1)  (timed + parity)
  "P"  pivoting : as D without the diagonal boost (about 87 % non-identity pivots at w=32; parity only).  Blocks in
                  which two pivot candidates come within PIVOT_GAP relative at any elimination step are redrawn
                  ("non-degenerate": a rounding difference between two correct eliminations must not flip a pivot)
  "R"  permuted : a family-D block with its rows shuffled: every block pivots, and partial pivoting undoes the
                  shuffle, so the chain is as well conditioned as D's (pivot parity at the full 1e-12 bar)
  "H"  halves   : as R with the shuffle confined to the upper and the lower half of each block

Layout follows block_diag_ilu.hpp:99-102 (blocks | sub | sup | bot | top), ld >= blockw.
"""
import numpy as np

from .datastruct import ndata as _ndata


PIVOT_GAP = 1e-10


def min_pivot_gap(blk):
    """Per block, the smallest relative gap between the two largest pivot candidates over all elimination steps of
    LU with partial pivoting.  blk: [nblocks, w, w] column-major blocks ([b, c, r])."""
    A = np.ascontiguousarray(np.transpose(blk, (0, 2, 1)))  # [b, r, c]
    nb, w, _ = A.shape
    gap = np.full(nb, np.inf)
    idx = np.arange(nb)
    for k in range(w - 1):
        col = np.abs(A[:, k:, k])
        part = np.partition(col, -2, axis=1)
        top, second = part[:, -1], part[:, -2]
        with np.errstate(divide="ignore", invalid="ignore"):
            gap = np.minimum(gap, np.where(top > 0, (top - second) / top, 0.0))
        p = col.argmax(axis=1) + k
        rk, rp = A[idx, k, :].copy(), A[idx, p, :].copy()
        A[idx, k, :], A[idx, p, :] = rp, rk
        piv = A[:, k, k]
        with np.errstate(divide="ignore", invalid="ignore"):
            l = A[:, k + 1:, k] / piv[:, None]
        A[:, k + 1:, k + 1:] -= l[:, :, None] * A[:, None, k, k + 1:]
    return gap


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
    elif family == "R":
        idx = np.arange(w)
        d = blk[:, idx, idx]
        blk[:, idx, idx] = d + np.where(d >= 0, 1.0, -1.0) * w
        prng = np.random.default_rng(7_000_003 * (seed + 1))
        for bi in range(nblocks):
            blk[bi] = blk[bi][:, prng.permutation(w)]  # rows of the column-major block [c, r]
    elif family == "H":
        # as R, but the rows are shuffled inside the upper and the lower half of each block separately: every block
        # pivots, yet no row interchange crosses the middle (the two-SM blockw-32 solve kernel's condition)
        idx = np.arange(w)
        d = blk[:, idx, idx]
        blk[:, idx, idx] = d + np.where(d >= 0, 1.0, -1.0) * w
        prng = np.random.default_rng(7_000_003 * (seed + 1) + 1)
        h = w // 2
        for bi in range(nblocks):
            perm = np.concatenate([prng.permutation(h), h + prng.permutation(w - h)])
            blk[bi] = blk[bi][:, perm]
    elif family == "P":
        if w > 1:
            # redraw the (very rare) blocks with a near-tie between pivot candidates; the main stream is untouched
            bad = np.nonzero(min_pivot_gap(blk) < PIVOT_GAP)[0]
            redraw = np.random.default_rng(9_000_011 * (seed + 1))
            while bad.size:
                blk[bad] = redraw.uniform(-1.0, 1.0, size=(bad.size, w, w))
                bad = bad[min_pivot_gap(blk[bad]) < PIVOT_GAP]
    else:
        raise ValueError("family must be 'D', 'P', 'R' or 'H'")
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
