"""RHS-column sharding across ranks (primary mode): every column of X is an
independent matvec, so rank r applies the replicated plan to its own contiguous slab of columns and
no collective sits on the data path.  `gather_columns` is the optional reassembly a caller that
wants the full block on every rank pays for (one all_gather of the Y slabs).

Host logic only: `apply_fn` is Plan.apply_device / Plan.apply_host in the product; the CPU tests
(gloo, world_size 2) drive it with the table interpreter.
"""
from __future__ import annotations

from typing import Callable, Tuple


def column_range(nrhs: int, world: int, rank: int) -> Tuple[int, int]:
    """Columns [c0, c1) of an n x nrhs block owned by `rank`: contiguous, sizes differ by <= 1,
    lower ranks take the remainder (rank r gets nrhs // world + (r < nrhs % world) columns)."""
    if world <= 0 or not 0 <= rank < world or nrhs < 0:
        raise ValueError("bad (nrhs, world, rank)")
    base, rem = divmod(nrhs, world)
    c0 = rank * base + min(rank, rem)
    return c0, c0 + base + (1 if rank < rem else 0)


def apply_column_shard(apply_fn: Callable, x, world: int, rank: int):
    """Apply to this rank's slab of a (replicated or locally generated) n x nrhs block `x`.
    Returns (c0, c1, y_local) with y_local = A x[:, c0:c1]; an empty slab skips the apply."""
    c0, c1 = column_range(x.shape[1], world, rank)
    if c1 == c0:
        return c0, c1, None
    return c0, c1, apply_fn(x[:, c0:c1])


def gather_columns(y_local, rows_out: int, nrhs: int, group=None):
    """Reassemble the full rows_out x nrhs result on every rank from the per-rank slabs (torch
    tensors; works with gloo on CPU tensors and nccl on CUDA tensors).  Slabs are padded to the
    widest one so a plain all_gather suffices."""
    import torch
    import torch.distributed as dist

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    width = max(column_range(nrhs, world, r)[1] - column_range(nrhs, world, r)[0] for r in range(world))
    c0, c1 = column_range(nrhs, world, rank)
    ref = y_local if y_local is not None else None
    if ref is None:
        raise ValueError("gather_columns needs a tensor (use an empty (rows_out, 0) slab for an idle rank)")
    pad = torch.zeros((rows_out, width), dtype=ref.dtype, device=ref.device)
    pad[:, : c1 - c0] = y_local
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    out = torch.empty((rows_out, nrhs), dtype=ref.dtype, device=ref.device)
    for r in range(world):
        a, b = column_range(nrhs, world, r)
        out[:, a:b] = parts[r][:, : b - a]
    return out
