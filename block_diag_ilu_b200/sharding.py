"""Batch sharding over ranks: one contiguous slab of systems per GPU, no data-path collective.

Systems of a batch share nothing, so rank r of a `torchrun` job factors and solves systems
[slab_bounds(nsys, world, r)) entirely on its own device.  The only communication is after the hot path:
`gather_norms` all-gathers the per-system residual norms (a few KB) and `max_over_ranks` reduces timings.
Both run on whatever backend the process group has (NCCL over NVLink on the GPU box, gloo in the CPU tests).
"""
import torch
import torch.distributed as dist


def slab_bounds(nsys, world, rank):
    """[lo, hi) of the systems rank `rank` owns: contiguous, sizes differ by at most one, earlier ranks larger."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad world/rank")
    base, extra = divmod(int(nsys), world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def slab_sizes(nsys, world):
    return [slab_bounds(nsys, world, r)[1] - slab_bounds(nsys, world, r)[0] for r in range(world)]


def gather_norms(local, nsys=None, group=None):
    """All-gather per-system values (1-D tensor, this rank's slab) -> the whole batch in system order.

    Slabs may have unequal sizes (pass the global nsys); tensors are padded to the largest slab for the
    collective and trimmed afterwards."""
    if not (dist.is_available() and dist.is_initialized()):
        return local
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    sizes = [local.numel()] * world if nsys is None else slab_sizes(nsys, world)
    if sizes[rank] != local.numel():
        raise ValueError("local slab has %d systems, expected %d" % (local.numel(), sizes[rank]))
    m = max(sizes)
    pad = torch.zeros(m, dtype=local.dtype, device=local.device)
    pad[:local.numel()] = local.reshape(-1)
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad, group=group)
    return torch.cat([o[:s] for o, s in zip(out, sizes)])


def max_over_ranks(value, device=None, group=None):
    """MAX all-reduce of a python float (timings, worst residual, info codes)."""
    if not (dist.is_available() and dist.is_initialized()):
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())
