"""Multi-GPU plumbing: one process per GPU over torch.distributed.

The hot path has no data-path collective.  Query points are split into contiguous shards
(evaluation), value columns likewise (dds); the parameters -- coefficients, and, if a rank did
not build them itself, exponents and generating points -- are broadcast once, and results stay
sharded unless the caller asks for one all_gather.  The same code runs on `nccl` (GPU) and on
`gloo` (CPU tests of this host logic).
"""
import numpy as np


def shard_bounds(n, rank, world):
    """[start, stop) of rank's contiguous share of n items: n // world each, the remainder to
    the last rank."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank / world size")
    per = n // world
    start = rank * per
    stop = n if rank == world - 1 else start + per
    return start, stop


def shard_sizes(n, world):
    return [b - a for a, b in (shard_bounds(n, r, world) for r in range(world))]


def broadcast_parameters(tensors, src=0):
    """Broadcast a list of tensors from `src` in place (one call per tensor, before the timed
    region); a no-op without an initialised process group."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return tensors
    for t in tensors:
        dist.broadcast(t, src=src)
    return tensors


def gather_values(local, n_total):
    """All-gather the (rows_r, k) result shards of a contiguous split of n_total rows into the
    full (n_total, k) tensor on every rank.  Shards are padded to the largest one because
    all_gather_into_tensor wants equal sizes."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return local
    world = dist.get_world_size()
    sizes = shard_sizes(n_total, world)
    if local.shape[0] != sizes[dist.get_rank()]:
        raise ValueError("local shard does not match the contiguous split")
    biggest = max(sizes)
    k = local.shape[1]
    padded = local
    if local.shape[0] != biggest:
        padded = torch.zeros((biggest, k), dtype=local.dtype, device=local.device)
        padded[: local.shape[0]] = local
    out = torch.empty((world * biggest, k), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, padded.contiguous())
    if all(s == biggest for s in sizes):
        return out
    parts = [out[r * biggest: r * biggest + sizes[r]] for r in range(world)]
    return torch.cat(parts, dim=0)


def eval_sharded(evaluate, xx_full, coeffs, gather=True):
    """Evaluate at this rank's shard of `xx_full` with `evaluate(xx_shard, coeffs) -> (rows, k)`
    (e.g. ``plan.eval``); returns the gathered (P, k) values, or the local shard."""
    import torch.distributed as dist
    rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
    world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    a, b = shard_bounds(xx_full.shape[0], rank, world)
    local = evaluate(xx_full[a:b], coeffs)
    return gather_values(local, xx_full.shape[0]) if gather else local


def dds_sharded(transform, values, gather=True):
    """Divided differences with the k value columns split over the ranks (columns are
    independent: no exchange): `transform(values[:, a:b]) -> (N, b-a)`."""
    import torch
    import torch.distributed as dist
    rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
    world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    a, b = shard_bounds(values.shape[1], rank, world)
    local = transform(values[:, a:b].contiguous())
    if not gather or world == 1:
        return local
    full_t = gather_values(local.t().contiguous(), values.shape[1])
    return full_t.t().contiguous()
