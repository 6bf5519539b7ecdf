"""View sharding over ``torch.distributed`` ranks (one process per GPU, NCCL over NVLink; ``gloo`` in CPU tests).

The path shards naturally by projection view: the phantom and the tables are replicated, rank r owns a
contiguous view range, traces and decomposes it with no communication, and the only collective is the all-gather that
assembles the view-major sinogram.  The reference itself is single-process.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def world():
    """(rank, world_size) of the default process group, (0, 1) when none is initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(n, rank, world_size):
    """Contiguous, balanced partition of range(n): the first n % world_size ranks get one extra item."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError(f'bad rank {rank} / world {world_size}')
    base, extra = divmod(n, world_size)
    begin = rank * base + min(rank, extra)
    return begin, base + (1 if rank < extra else 0)


def all_agree(flag, group=None):
    """True only if ``flag`` is true on EVERY rank (one MIN all-reduce of a byte; plain ``bool(flag)`` without a process
    group).  For set-up steps that may fail on one rank only -- mapping peer memory, say -- after which all ranks must take
    the same fallback or the next collective hangs."""
    rank, ws = world()
    if ws == 1:
        return bool(flag)
    dev = torch.device('cuda', torch.cuda.current_device()) if dist.get_backend(group) == 'nccl' else torch.device('cpu')
    t = torch.tensor([1 if flag else 0], dtype=torch.int32, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
    return bool(int(t.item()))


def shard_sizes(n, world_size):
    return [shard_range(n, r, world_size)[1] for r in range(world_size)]


def all_gather_views(local, n_total, dim=0, group=None):
    """Assemble the full tensor from per-rank shards that partition ``dim`` as :func:`shard_range` does.

    Equal shards use one ``all_gather_into_tensor`` (a single NCCL all-gather when ``dim`` is the outermost axis of a
    contiguous tensor); ragged shards are padded to the largest one and trimmed after the gather.
    """
    rank, ws = world()
    if ws == 1:
        return local
    sizes = shard_sizes(n_total, ws)
    if local.shape[dim] != sizes[rank]:
        raise ValueError(f'rank {rank} holds {local.shape[dim]} items along dim {dim}, expected {sizes[rank]}')
    x = local.movedim(dim, 0).contiguous()
    biggest = max(sizes)
    if biggest == 0:
        return local
    if x.shape[0] < biggest:
        pad = torch.zeros((biggest - x.shape[0],) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        x = torch.cat([x, pad], dim=0)
    out = torch.empty((ws * biggest,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    dist.all_gather_into_tensor(out, x, group=group)
    if min(sizes) != biggest:
        out = torch.cat([out[r * biggest:r * biggest + sizes[r]] for r in range(ws)], dim=0)
    return out.movedim(0, dim)
