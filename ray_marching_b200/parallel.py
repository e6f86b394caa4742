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


def shard_strips(total: int, rank: int, world_size: int, strip_rays: int) -> dict:
    """Interleaved shard of ``total`` rays (row-interleaved tiles): strips of
    ``strip_rays`` consecutive rays dealt round-robin, rank g taking strips g, g+G, ...  Returns the
    keyword arguments of ``RenderLoop.render_planes`` / ``inverse_loss``:
    ``ray_begin, n_rays, strip_rays, strip_stride`` (``n_rays`` = this rank's local ray count; the last
    strip of the batch may be short).  Uneven per-ray cost (culling, early exit, open scenes) then
    averages out over the image instead of loading the rank that owns the busy region."""
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside world of {world_size}")
    if strip_rays <= 0 or strip_rays % 4:
        raise ValueError("strip_rays must be a positive multiple of 4 (128-bit plane stores)")
    n_strips = (total + strip_rays - 1) // strip_rays
    mine = len(range(rank, n_strips, world_size))
    n = mine * strip_rays
    if mine and (rank + (mine - 1) * world_size) == n_strips - 1:          # I own the (possibly short) last strip
        n -= n_strips * strip_rays - total
    return dict(ray_begin=rank * strip_rays, n_rays=n, strip_rays=strip_rays,
                strip_stride=world_size * strip_rays)


def strip_ray_index(shard: dict, device=None) -> torch.Tensor:
    """Global ray index of every local ray of an interleaved shard (int64 ``(n_rays,)``): scatter a
    local plane into the full image with ``full[strip_ray_index(shard)] = plane``."""
    l = torch.arange(shard["n_rays"], dtype=torch.int64, device=device)
    s = shard["strip_rays"]
    return shard["ray_begin"] + (l // s) * shard["strip_stride"] + l % s


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


class PipelinedFrames:
    """Frames of a view-sharded job with the collective taken off the critical path.

    A normalising shader (distance / proximity / laplacian) needs the global range over the whole
    camera batch, i.e. one 8-float MAX all-reduce between ``rm_render`` and ``rm_shade_finalize``.  On
    config 2 that is ~25 us of latency (NCCL launch, peer skew, the finalize pass) behind a 180 us
    render.  Here frame i's {all-reduce, finalize} run on a side stream while the main stream renders
    frame i+1 into the other buffer: steady-state time per frame = the render alone.  Results are
    identical to the unpipelined sequence (same kernels, same inputs).

    ``submit(q, t)`` renders on the current stream and returns the frame's ``RenderResult``; its
    normalised planes are complete once :meth:`wait` (for that result, or for everything in flight)
    has been called on the consuming stream.
    """

    def __init__(self, loop, modes: Sequence[int], *, marching_steps: int = 32, bounces: int = 0,
                 depth: int = 2, group=None):
        self.loop, self.modes, self.steps, self.bounces, self.group = loop, tuple(modes), marching_steps, bounces, group
        self.mask = 0
        for m in self.modes:
            self.mask |= 1 << (m % 8)
        self.normalised = bool(self.mask & ((1 << 1) | (1 << 2) | (1 << 5)))
        self.bufs = [None] * max(depth, 2)
        self.done = [None] * len(self.bufs)          # event: buffer finalised
        self.side = None
        self.i = 0

    def submit(self, orientations: torch.Tensor, translations: torch.Tensor):
        from . import ops
        b = self.i % len(self.bufs)
        self.i += 1
        main = torch.cuda.current_stream(orientations.device)
        if self.side is None:
            self.side = torch.cuda.Stream(orientations.device)
        if self.done[b] is not None:
            main.wait_event(self.done[b])            # the buffer's previous frame has been finalised
        res = self.loop.render_planes(orientations, translations, modes=self.modes,
                                      marching_steps=self.steps, bounces=self.bounces, finalize=False,
                                      out=self.bufs[b])
        self.bufs[b] = res
        if self.normalised:
            rendered = torch.cuda.Event()
            rendered.record(main)
            with torch.cuda.stream(self.side):
                self.side.wait_event(rendered)
                allreduce_shader_range(res.reduce, self.group)
                ops.shade_finalize(res, self.mask)
                self.done[b] = torch.cuda.Event()
                self.done[b].record(self.side)
        res._pipeline_slot = b
        return res

    def wait(self, res=None) -> None:
        """Makes the current stream wait until ``res`` (default: every frame in flight) is finalised."""
        slots = range(len(self.bufs)) if res is None else [res._pipeline_slot]
        for b in slots:
            if self.done[b] is not None:
                torch.cuda.current_stream().wait_event(self.done[b])
