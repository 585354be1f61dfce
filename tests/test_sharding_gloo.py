"""N > 1 path on the CPU: slab arithmetic and the only collectives of the path (norm gather, max timing),
world_size 2 over gloo.  The data path itself never communicates."""
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.multiprocessing as mp  # noqa: E402

from block_diag_ilu_b200 import sharding  # noqa: E402


def test_slab_bounds_cover_and_are_contiguous():
    for nsys in (0, 1, 7, 64, 4096, 4097):
        for world in (1, 2, 3, 4, 8):
            edges = [sharding.slab_bounds(nsys, world, r) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == nsys
            for (a, b), (c, d) in zip(edges[:-1], edges[1:]):
                assert b == c and b - a >= d - c >= b - a - 1
    assert sharding.slab_sizes(4096, 8) == [512] * 8          # BASELINE config 5
    assert sharding.slab_sizes(64, 4) == [16] * 4
    with pytest.raises(ValueError):
        sharding.slab_bounds(4, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, nsys, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = sharding.slab_bounds(nsys, world, rank)
        # each rank "solves" its own slab: the per-system value depends only on the global system index
        local = torch.arange(lo, hi, dtype=torch.float64) ** 2 + 0.5
        allv = sharding.gather_norms(local, nsys)
        t = sharding.max_over_ranks(1.0 + rank)
        q.put((rank, allv.numpy().tolist(), t))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("nsys", [8, 7])
def test_gather_norms_world2_gloo(nsys):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, nsys, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = (np.arange(nsys, dtype=np.float64) ** 2 + 0.5).tolist()
    for rank, allv, t in res:
        assert allv == want
        assert t == 2.0


def test_single_process_passthrough():
    v = torch.arange(5, dtype=torch.float64)
    assert sharding.gather_norms(v) is v
    assert sharding.max_over_ranks(3.5) == 3.5
