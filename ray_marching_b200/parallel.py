"""Multi-GPU plumbing: one process per GPU, ``torch.distributed`` (NCCL over NVLink on the box,
gloo in the CPU tests).  Rays are independent (the stencil samples the SDF, not neighbouring
pixels; ``rendering/ray_marching.py:116``), so the data path never exchanges pixels.  Only two tiny
collectives exist:

* ``allreduce_shader_range``: one MAX over the 8-float reduction scratch (minima stored negated) between ``rm_render`` and
  ``rm_shade_finalize`` when a normalising shader (distance / proximity / laplacian; global
  ``.min()/.max()`` over the whole camera batch, ``shader.py:35-36,52-53,84``) runs on a sharded batch;
* ``allreduce_gradients``: SUM over one flat buffer holding every scene-parameter and camera
  gradient (inverse rendering, config 4) -- latency bound (<= ~21 KB), hence a single call.
"""
from __future__ import annotations

from typing import Iterable, List, Sequence, Tuple

import torch
import torch.distributed as dist

# the reduction scratch (include/rm_b200.h) stores minima negated: every entry combines with MAX

def world() -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(total: int, rank: int, world_size: int, align: int = 4) -> Tuple[int, int]:
    """Contiguous ``[begin, end)`` slice of ``total`` rays for ``rank``; boundaries are multiples of
    ``align`` (the kernels' 4-rays-per-thread, 128-bit store granularity) except the last."""
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside world of {world_size}")
    units = (total + align - 1) // align
    base, extra = divmod(units, world_size)
    begin_u = rank * base + min(rank, extra)
    end_u = begin_u + base + (1 if rank < extra else 0)
    return min(begin_u * align, total), min(end_u * align, total)


def shard_views(n_views: int, rank: int, world_size: int) -> List[int]:
    """Round-robin view assignment: rank g renders views g, g+G, ...."""
    return list(range(rank, n_views, world_size))


def allreduce_shader_range(reduce: torch.Tensor, group=None) -> torch.Tensor:
    """In-place MAX all-reduce of the 8-float shader reduction scratch (no-op for a single process):
    one latency-bound collective, no staging kernels."""
    if world()[1] > 1:
        dist.all_reduce(reduce, op=dist.ReduceOp.MAX, group=group)
    return reduce


def allreduce_gradients(params: Sequence[torch.Tensor], average: bool = True, group=None) -> int:
    """One SUM all-reduce over the concatenation of every ``.grad`` (missing grads count as zero so
    that all ranks agree on the layout); results are written back in place.  Returns the number of
    floats reduced."""
    rank, ws = world()
    grads = []
    for p in params:
        if p.grad is None:
            p.grad = torch.zeros_like(p)
        grads.append(p.grad)
    if not grads:
        return 0
    flat = torch.cat([g.reshape(-1) for g in grads])
    if ws > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        if average:
            flat /= ws
    off = 0
    for g in grads:
        g.copy_(flat[off:off + g.numel()].view_as(g))
        off += g.numel()
    return flat.numel()
