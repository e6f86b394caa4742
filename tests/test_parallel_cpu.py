"""CPU (gloo, world_size 2) coverage of the multi-GPU host logic: ray sharding, the MIN/MAX
all-reduce of the shader range scratch, and the flat gradient all-reduce."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ray_marching_b200 import parallel


def test_shard_range_partitions_exactly():
    for total in (0, 1, 3, 4, 5, 1023, 307200, 2073600, 3600):
        for ws in (1, 2, 3, 4, 8):
            cuts = [parallel.shard_range(total, r, ws) for r in range(ws)]
            assert cuts[0][0] == 0 and cuts[-1][1] == total
            for (a, b), (c, d) in zip(cuts[:-1], cuts[1:]):
                assert b == c and a <= b
            for a, b in cuts[:-1]:
                assert (a % 4 == 0 or a == total) and (b % 4 == 0 or b == total)
            sizes = [b - a for a, b in cuts]
            assert max(sizes) - min(sizes) <= 4 + 3
    with pytest.raises(ValueError):
        parallel.shard_range(10, 2, 2)
    assert parallel.shard_views(64, 3, 8) == list(range(3, 64, 8))
    assert sum(len(parallel.shard_views(10, r, 4)) for r in range(4)) == 10


def test_shard_strips_partitions_exactly():
    """Interleaved strips: every ray belongs to exactly one rank; ranks differ by at most one strip."""
    for total, strip in ((3840 * 2160, 8 * 3840), (1000, 64), (1000, 1000), (1000, 8), (7, 4), (0, 4), (64, 64)):
        for ws in (1, 2, 3, 8):
            shards = [parallel.shard_strips(total, r, ws, strip) for r in range(ws)]
            idx = torch.cat([parallel.strip_ray_index(s) for s in shards])
            assert idx.numel() == total and torch.equal(idx.sort().values, torch.arange(total))
            sizes = [s["n_rays"] for s in shards]
            assert max(sizes) - min(sizes) <= strip
            for r, s in enumerate(shards):
                assert s["ray_begin"] == r * strip and s["strip_stride"] == ws * strip
    with pytest.raises(ValueError):
        parallel.shard_strips(100, 0, 2, 6)
    with pytest.raises(ValueError):
        parallel.shard_strips(100, 2, 2, 8)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, ws, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        # shader range scratch: per-rank partial (-min, max, -min, max, max|lap|, 3 NaN flags);
        # minima are stored negated so that one MAX all-reduce combines every entry
        red = torch.tensor([-(1.0 + rank), 5.0 - rank, 2.0 * rank, 0.5 * rank, 3.0 + rank, 0.0,
                            float(rank == 1), 0.0])
        parallel.allreduce_shader_range(red)
        # gradients: one parameter has no grad on rank 1
        a = torch.nn.Parameter(torch.ones(3))
        b = torch.nn.Parameter(torch.ones(2, 2))
        a.grad = torch.full((3,), float(rank + 1))
        if rank == 0:
            b.grad = torch.full((2, 2), 4.0)
        n = parallel.allreduce_gradients([a, b], average=True)
        out.put((rank, red.tolist(), a.grad.tolist(), b.grad.tolist(), n))
    finally:
        dist.destroy_process_group()


def test_collectives_world_size_2_gloo():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(out.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, red, ga, gb, n in res:
        assert red == [-1.0, 5.0, 2.0, 0.5, 4.0, 0.0, 1.0, 0.0]      # min 1, max 5, min -2, max .5
        assert ga == [1.5, 1.5, 1.5] and gb == [[2.0, 2.0], [2.0, 2.0]] and n == 7


def test_single_process_is_a_noop():
    red = torch.arange(8.0)
    assert parallel.allreduce_shader_range(red.clone()).equal(red)
    p = torch.nn.Parameter(torch.ones(2))
    p.grad = torch.tensor([3.0, 4.0])
    assert parallel.allreduce_gradients([p]) == 2 and p.grad.tolist() == [3.0, 4.0]
