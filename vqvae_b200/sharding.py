"""Contiguous sharding helpers shared by the pipeline (query rows, 512-edge batches) and the K-medoids sources."""
import torch


def shard_range(total, rank, world, multiple=1):
    """Contiguous [lo,hi) of `total` items for `rank`, boundaries at multiples of `multiple`."""
    units = -(-total // multiple)
    per, extra = divmod(units, world)
    lo_u = rank * per + min(rank, extra)
    hi_u = lo_u + per + (1 if rank < extra else 0)
    return min(lo_u * multiple, total), min(hi_u * multiple, total)


def all_gather_rows(x, total_rows, world, rank, group, dist, multiple=1):
    """Concatenate per-rank row shards (shard_range layout) into the full tensor on every rank."""
    if world == 1:
        return x
    sizes = [shard_range(total_rows, r, world, multiple) for r in range(world)]
    mx = max(hi - lo for lo, hi in sizes)
    pad = torch.zeros((mx,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    pad[: x.shape[0]] = x
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad, group=group)
    return torch.cat([b[: hi - lo] for b, (lo, hi) in zip(bufs, sizes)], 0)


def allreduce_min_assignment(dmin32, label, dist, group=None):
    """Combine per-rank partial assignments (dmin32 f32[n] >= 0 or +inf, label int32[n] or -1), in place.

    One allreduce-MIN over packed (float32 bits << 32 | label) int64 keys: non-negative floats order like their
    bit patterns, so the minimum key carries the smallest distance and, on float32 ties, the lowest label -- the
    reference's `argmin over k of float32(D[k,v])` rule (kmeans_optimized.py:100).  label -1 packs as 0xFFFFFFFF and
    loses every tie; a vertex no rank reaches keeps (+inf, -1)."""
    key = (dmin32.view(torch.int32).to(torch.int64) << 32) | (label.to(torch.int64) & 0xFFFFFFFF)
    dist.all_reduce(key, op=dist.ReduceOp.MIN, group=group)
    dmin32.copy_((key >> 32).to(torch.int32).view(torch.float32))
    label.copy_((key & 0xFFFFFFFF).to(torch.int32))
    return dmin32, label
