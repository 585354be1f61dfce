"""Packed-layout arithmetic and numpy packing helpers.

Mirrors the reference's public helpers `diag_data_len`, `alloc_compressed`, `get_compressed`
(block_diag_ilu/datastruct.py:7-65, re-exported from block_diag_ilu/__init__.py:14) with the same
names, argument meaning and errors, but packed with whole-array numpy indexing instead of
element-by-element Python loops.  The layout itself is BlockDiagMatrix's flat `m_data`
(block_diag_ilu/include/block_diag_ilu.hpp:98-102, 203-221):

    [ blocks : nblocks*blockw*ld ][ sub : D*ld ][ sup : D*ld ][ bot : S*ld ][ top : S*ld ]
    D = ndiag*nblocks - (ndiag^2+ndiag)/2,  S = (nsat^2+nsat)/2
    block(b,r,c) = data[(b*blockw + c)*ld + r]                (column-major per block)
    sub(di,b,c)  = data[blk + (di*nblocks - (di^2+di)/2 + b)*ld + c]   = A[(b+di+1)*w + c, b*w + c]
    sup(di,b,c)  = same + D*ld                                         = A[b*w + c, (b+di+1)*w + c]
"""
import numpy as np

ALIGNMENT_BYTES = 64  # anyode_matrix.hpp:21


def default_ld(blockw):
    """blockw rounded up to a 64-byte multiple of doubles (block_diag_ilu.hpp:98, anyode_matrix.hpp:18-20)."""
    return ((blockw * 8 + ALIGNMENT_BYTES - 1) & ~(ALIGNMENT_BYTES - 1)) // 8


def diag_rows(nblocks, ndiag):
    """Number of ld-long chunks in the sub (or sup) region."""
    return ndiag * nblocks - (ndiag * ndiag + ndiag) // 2


def sat_rows(nsat):
    return (nsat * nsat + nsat) // 2


def ndata(nblocks, blockw, ndiag, nsat=0, ld=0):
    """BlockDiagMatrix::m_ndata (block_diag_ilu.hpp:99-102)."""
    ld = ld or default_ld(blockw)
    return (nblocks * blockw + 2 * (diag_rows(nblocks, ndiag) + sat_rows(nsat))) * ld


def offsets(nblocks, blockw, ndiag, nsat=0, ld=0):
    """(sub, sup, bot, top) start offsets inside m_data."""
    ld = ld or default_ld(blockw)
    blk = nblocks * blockw * ld
    dh = diag_rows(nblocks, ndiag) * ld
    sh = sat_rows(nsat) * ld
    return blk, blk + dh, blk + 2 * dh, blk + 2 * dh + sh


def diag_chunk(nblocks, di, bi):
    """Chunk index of (di, bi) inside the sub/sup region (block_diag_ilu.hpp:206)."""
    return di * nblocks - (di * di + di) // 2 + bi


def sat_chunk(si, bi):
    """Chunk index of (si, bi) inside the bot/top region (block_diag_ilu.hpp:214)."""
    return (si * si + si) // 2 + bi


# ---- the reference's public helpers (tight ld = blockw, no satellites in get_compressed) ----------

def diag_data_len(N, n, ndiag):
    return n * (N * ndiag - ((ndiag + 1) * (ndiag + 1) - (ndiag + 1)) // 2)


def alloc_compressed(nblocks, blockw, ndiag, nsat):
    n_block_elem = blockw * blockw * nblocks
    n_diag_elem = 2 * diag_data_len(nblocks, blockw, ndiag)
    n_sat_elem = (nsat * nsat + nsat) * blockw
    return np.zeros(n_block_elem + n_diag_elem + n_sat_elem)


def pack_dense(A, nblocks, blockw, ndiag, nsat=0, ld=None, cmaj=True):
    """Dense (nblocks*blockw)^2 -> packed m_data with leading dimension ld (vectorised)."""
    ld = ld or blockw
    w, N = blockw, nblocks
    A = np.asarray(A, dtype=np.float64)
    out = np.zeros(ndata(N, w, ndiag, nsat, ld))
    o_sub, o_sup, o_bot, o_top = offsets(N, w, ndiag, nsat, ld)
    blocks = A.reshape(N, w, N, w)[np.arange(N), :, np.arange(N), :]  # [b, r, c]
    dst = out[:o_sub].reshape(N, w, ld)
    dst[:, :, :w] = blocks.transpose(0, 2, 1) if cmaj else blocks
    ci = np.arange(w)
    for di in range(ndiag):
        for bi in range(N - di - 1):
            k = diag_chunk(N, di, bi) * ld
            out[o_sub + k:o_sub + k + w] = A[(bi + di + 1) * w + ci, bi * w + ci]
            out[o_sup + k:o_sup + k + w] = A[bi * w + ci, (bi + di + 1) * w + ci]
    for si in range(nsat):
        for bi in range(si + 1):
            k = sat_chunk(si, bi) * ld
            out[o_top + k:o_top + k + w] = A[bi * w + ci, (N - si - 1 + bi) * w + ci]
            out[o_bot + k:o_bot + k + w] = A[(N - si - 1 + bi) * w + ci, bi * w + ci]
    return out


def get_compressed(A, N, n, ndiag, cmaj=True):
    """Dense matrix -> packed blocks (column major by default), then sub-, then super-diagonals.

    Same contract as the reference's get_compressed (datastruct.py:21-65): tight ld = n, no
    satellites, ValueError when A.shape != (n*N, n*N).
    """
    A = np.asarray(A)
    if A.shape != (n * N, n * N):
        raise ValueError("Shape of A != (n*N, n*N)")
    return pack_dense(A, N, n, ndiag, 0, n, cmaj)


def unpack_dense(data, nblocks, blockw, ndiag, nsat=0, ld=None):
    """Packed m_data -> dense matrix (structural zeros elsewhere)."""
    ld = ld or blockw
    w, N = blockw, nblocks
    data = np.asarray(data, dtype=np.float64)
    A = np.zeros((N * w, N * w))
    o_sub, o_sup, o_bot, o_top = offsets(N, w, ndiag, nsat, ld)
    blocks = data[:o_sub].reshape(N, w, ld)[:, :, :w]  # [b, c, r]
    for bi in range(N):
        A[bi * w:(bi + 1) * w, bi * w:(bi + 1) * w] = blocks[bi].T
    ci = np.arange(w)
    for di in range(ndiag):
        for bi in range(N - di - 1):
            k = diag_chunk(N, di, bi) * ld
            A[(bi + di + 1) * w + ci, bi * w + ci] = data[o_sub + k:o_sub + k + w]
            A[bi * w + ci, (bi + di + 1) * w + ci] = data[o_sup + k:o_sup + k + w]
    for si in range(nsat):
        for bi in range(si + 1):
            k = sat_chunk(si, bi) * ld
            A[bi * w + ci, (N - si - 1 + bi) * w + ci] = data[o_top + k:o_top + k + w]
            A[(N - si - 1 + bi) * w + ci, bi * w + ci] = data[o_bot + k:o_bot + k + w]
    return A
