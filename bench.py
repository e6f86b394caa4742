#!/usr/bin/env python
"""Benchmark of the sphere-tracing hot path (contract: task statement / DESIGN.md §7).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c1|c2|c3|c4|c4b|c5] [--impl reference]

Headline = BASELINE.json config 3 ("the 4K multi-primitive scene" of north_star): one 3840x2160 frame
of the synthetic-64 smooth union in its shell, 3 march segments (2 reflection bounces), lambertian.
A *step* is one pass of the hot path over that frame.  With N > 1 the SAME frame is sharded over the
ranks in row-interleaved strips (``parallel.shard_strips``; no data-path collective, strong scaling).
``value`` = whole-job Mrays/s with everything resident in HBM, ``e2e`` the same through the public API
from pinned host buffers (pose in, image out), ``roofline`` = the render kernel timed alone with CUDA
events against the measured FFMA peak.  The other BASELINE configs ride along as ``other_configs``
(short in-process runs, each with its own value / ms_per_step / roofline.frac / kernel_ms), so that c1, the
tape-specialised c2, forward+backward (c4, c4b) and the 512-primitive view batch (c5) are measured by the
same command.  ``--impl reference`` times the CPU oracle port (oracle/sdf_oracle.py -- the reference's
own PyTorch algorithm restated) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

if "reference" in sys.argv or int(os.environ.get("WORLD_SIZE", "1")) == 1:
    # the CPU arm uses every host core (torchrun exports OMP_NUM_THREADS=1 to its workers)
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)

import torch  # noqa: E402

PX = 3.45e-6            # main.py:25
STEPS = 32              # main.py:26
EPS = 5e-2              # main.py:42
DEFAULT_WORKLOAD = "c3"
STRIP_ROWS = 8          # interleaved sharding of one image: strips of 8 rows
C5_VIEWS = 8            # views per step of the c5 batch (first 8 of the 64-view ring)

# shader flop counts of SURVEY.md §8(d)
SHADER_FLOPS = {0: 8, 1: 18, 2: 7, 3: 7, 4: 6, 5: 7, 6: 20, 7: 20}
STENCIL_MODES = {0, 4, 5, 6, 7}


def workload(name: str):
    """BASELINE.json configs as concrete inputs.  ``sharding`` says how the
    config maps onto N ranks: 'strips' = one image in row-interleaved strips (strong), 'view-batch' = a
    fixed batch of views dealt round-robin (strong), 'views' = one view per rank (weak)."""
    from ray_marching_b200 import scene as S
    syn64 = lambda: S.make_synthetic_scene(64, 0, True)
    table = {
        "c1": dict(scene=S.make_test_scene2, desc="make_test_scene2 640x480 distance+proximity",
                   width=640, height=480, modes=(1, 2), bounces=0, cam_t=(0.0, 0.0, 1.0), backward=False,
                   sharding="views", heavy=False),
        "c2": dict(scene=S.make_test_scene2,
                   desc="make_test_scene2 1920x1080 normal+laplacian+lambertian+vignette from one march",
                   width=1920, height=1080, modes=(4, 5, 0, 3), bounces=0, cam_t=(0.0, 0.0, 1.0),
                   backward=False, sharding="views", heavy=False),
        "c3": dict(scene=syn64,
                   desc="synthetic-64 smooth union in shell, 3840x2160, 2 reflection bounces, lambertian",
                   width=3840, height=2160, modes=(0,), bounces=2, cam_t=(0.0, 0.0, -4.0), backward=False,
                   sharding="strips", heavy=True),
        "c4": dict(scene=S.make_test_scene2,
                   desc="inverse rendering fwd+bwd, make_test_scene2 1024x1024, lambertian+depth MSE, Adam",
                   width=1024, height=1024, modes=(0,), bounces=0, cam_t=(0.0, 0.1, 1.0), backward=True,
                   sharding="views", heavy=False),
        "c4b": dict(scene=syn64,
                    desc="inverse rendering fwd+bwd, synthetic-64 in shell (650 params) 1024x1024, "
                         "lambertian+depth MSE, Adam",
                    width=1024, height=1024, modes=(0,), bounces=0, cam_t=(0.0, 0.0, -4.0), backward=True,
                    sharding="views", heavy=True),
        "c5": dict(scene=lambda: S.make_synthetic_scene(512, 0, True),
                   desc=f"synthetic-512 smooth union in shell, batch of {C5_VIEWS} ring views at 3840x2160 "
                        f"(views 0..{C5_VIEWS - 1} of the 64-view ring), lambertian",
                   width=3840, height=2160, modes=(0,), bounces=0, cam_t=(0.0, 0.0, -4.0), backward=False,
                   sharding="view-batch", heavy=True, views=C5_VIEWS, ring=64),
    }
    if name not in table:
        raise SystemExit(f"unknown workload {name}")
    return dict(table[name], name=name)


def flops_per_ray(tape, wl, steps=STEPS) -> float:
    """Algorithmic forward flops per ray, SURVEY.md §8(d) counting rule."""
    F = tape.flops_per_eval
    stencil = bool(set(wl["modes"]) & STENCIL_MODES) or wl["bounces"] > 0
    seg = steps * (F + 6) + F + (4 * F + 45 if stencil else 0)
    total = 63 + seg * (wl["bounces"] + 1) + sum(SHADER_FLOPS[m] for m in wl["modes"])
    if wl["bounces"]:
        total += wl["bounces"] * 20            # hit test + reflection + lift
    return float(total)


def mufu_per_ray(tape, wl, steps=STEPS) -> float:
    """Special-function-unit operations per ray the formulas need (one per sqrt / exp / log / division;
    SURVEY.md 6.3: the XU pipe runs at 1/8 of the FP32 lane rate and is the next ceiling)."""
    from ray_marching_b200 import tape as T
    per_leaf = {T.OP_SPHERE: 1, T.OP_BOX: 1, T.OP_PLANE: 0, T.OP_LINE: 1, T.OP_DISK: 2, T.OP_TORUS: 2}
    ev = sum(per_leaf[k] for k in tape.leaf_kinds)
    smooth = [m for m in tape.root.modules() if type(m).__name__ == "SDFSmoothUnion"]
    ev += sum(len(m.sdfs) + 2 for m in smooth)           # one exp per child, one log + one division per union
    stencil = bool(set(wl["modes"]) & STENCIL_MODES) or wl["bounces"] > 0
    evals = steps + 1 + (4 if stencil else 0)
    return float(ev * evals * (wl["bounces"] + 1) + 8)


def ring_pose(j: int, n: int, base_t, device):
    """View j of n on a ring around the y axis looking at the origin."""
    from ray_marching_b200 import quaternion as Q
    th = 2 * math.pi * j / max(n, 1)
    q = torch.tensor([[math.cos(th / 2), 0.0, math.sin(th / 2), 0.0]])
    t = Q.rotation(torch.tensor([list(base_t)]), q)
    return q.to(device), t.to(device)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the GPU is under load."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.FIELDS}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def measure_fp32_peak(device) -> dict:
    """FFMA-only microbenchmark on all SMs -> TFLOP/s (the FP32 roofline denominator)."""
    from ray_marching_b200 import _lib
    lib = _lib.lib()
    sink = torch.zeros(4, device=device)
    blocks, iters = 148 * 8, 4096
    st = torch.cuda.current_stream(device).cuda_stream
    for _ in range(3):
        _lib.check(lib.rm_probe_ffma(blocks, iters, sink.data_ptr(), st), "rm_probe_ffma")
    best = float("inf")
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(lib.rm_probe_ffma(blocks, iters, sink.data_ptr(), st), "rm_probe_ffma")
        e1.record()
        e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    flops = blocks * 256 * iters * 128 * 2
    return {"tflops": flops / (best * 1e-3) / 1e12, "ms": best}


PEAK_SOURCE = ("measured: FFMA microbenchmark (rm_probe_ffma) on this GPU; nominal 74.5 TFLOP/s at 1965 MHz "
               "(MEASURED_PEAKS.json holds HBM and bf16 tensor figures only)")


def ncu_traffic(kernel: str, workload: str):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel (bytes per launch) from the
    newest COMMITTED `ncu --set full` summary of this workload (profiles/rNN/*_<kernel>_<wl>_summary.txt);
    None when no capture of this workload is committed.  (bench.py never runs under a profiler, so this
    figure comes from a file, not from this run -- labelled as such in the JSON.)"""
    import glob
    import re
    tag = {"render_kernel": "render", "render_bwd_kernel": "render_bwd",
           "render_loss_bwd_kernel": "render_loss_bwd", "render_loss_onfly_kernel": "render_loss_onfly", "render_loss_onfly4_kernel": "render_loss_onfly4",
           "loss_replay_kernel + render_loss_reverse_kernel": "render_loss_reverse"}[kernel]
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*", f"*_{tag}_{workload}_summary.txt")))
    if not files:
        return None, None
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    total, seen = 0.0, 0
    for line in open(files[-1]):
        m = re.match(r"dram__bytes_(read|write)\.sum\s+(\w+)\s+([0-9.,]+)", line)
        if m:
            total += float(m.group(3).replace(",", "")) * scale.get(m.group(2), 1.0)
            seen += 1
    return (total, "committed ncu capture: " + os.path.relpath(files[-1], ROOT)) if seen == 2 else (None, None)


# ---------------------------------------------------------------------------- GPU arm: plumbing
class Ctx:
    """Rank / device / collectives shared by every workload of one bench process."""

    def __init__(self, args):
        import torch.distributed as dist
        self.dist = dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=self.dev)
        self.args = args
        self.flush_buf = torch.empty(192 * 1024 * 1024 // 4, device=self.dev)      # 192 MB > 126 MB L2
        self.peak = measure_fp32_peak(self.dev)
        self.numa = bind_numa(self.local) if args.numa else {"bound": False, "why": "--numa 0"}

    def flush(self):
        self.flush_buf.zero_()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        torch.cuda.synchronize()

    def gather(self, x: float):
        """Every rank's value of ``x`` (list, rank order)."""
        if self.world == 1:
            return [float(x)]
        t = torch.tensor([float(x)], device=self.dev, dtype=torch.float64)
        out = [torch.zeros_like(t) for _ in range(self.world)]
        self.dist.all_gather(out, t)
        return [float(o.item()) for o in out]

    @staticmethod
    def stats(xs):
        s = sorted(xs)
        return {"min": s[0], "median": s[len(s) // 2], "max": s[-1]}


def bind_numa(local_rank: int) -> dict:
    """Best effort: pin this process (and with it the first-touch placement of its pinned host buffers)
    to the CPUs of the GPU's NUMA node, so that N ranks do not all stage their images through one node."""
    try:
        prop = torch.cuda.get_device_properties(local_rank)
        bus = f"{prop.pci_domain_id:04x}:{prop.pci_bus_id:02x}:{prop.pci_device_id:02x}.0"
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node < 0:
            return {"bound": False, "why": "numa_node = -1 (not exposed in this container)"}
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return {"bound": False, "why": f"no allowed CPU on node {node}"}
        os.sched_setaffinity(0, cpus)
        return {"bound": True, "node": node, "cpus": len(cpus)}
    except Exception as exc:                                     # noqa: BLE001 -- diagnostics only
        return {"bound": False, "why": str(exc)[:80]}


def time_total(ctx, steps, body, join=None, flush=False):
    """EXACTLY `steps` calls of body(i) inside ONE CUDA-event pair on the current stream, bracketed by a
    barrier + synchronize on both sides; `join()` makes the current stream wait for side streams before
    the closing event.  Returns (device ms, wall ms) of this rank."""
    ctx.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.perf_counter()
    e0.record()
    for i in range(steps):
        if flush:
            ctx.flush()                                  # inside the timed region (stated in config.l2)
        body(i)
    if join is not None:
        join()
    e1.record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - w0) * 1e3
    ctx.barrier()
    return e0.elapsed_time(e1), wall


def time_each(ctx, steps, body):
    """Per-step CUDA-event pairs on ONE stream (every kernel and collective of the step is launched on
    it), L2 flushed between steps outside the pairs.  Returns (sum of device ms, wall ms incl. flushes)."""
    ctx.barrier()
    evs = []
    w0 = time.perf_counter()
    for i in range(steps):
        ctx.flush()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        body(i)
        e1.record()
        evs.append((e0, e1))
    torch.cuda.synchronize()
    wall = (time.perf_counter() - w0) * 1e3
    ctx.barrier()
    return sum(a.elapsed_time(b) for a, b in evs), wall


def kernel_alone(ctx, launch, reps):
    """Average CUDA-event duration of one kernel launch (L2 flushed before each, untimed)."""
    ks = []
    for _ in range(reps):
        ctx.flush()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        launch()
        e1.record()
        e1.synchronize()
        ks.append(e0.elapsed_time(e1))
    return sum(ks) / len(ks)


# ---------------------------------------------------------------------------- GPU arm: forward workloads
def run_forward(ctx, wl, steps, warmup, headline):
    """One forward workload on all ranks -> its record (rank 0 fills in the whole-job figures)."""
    from ray_marching_b200 import RenderLoop, ops, parallel
    from ray_marching_b200.tape import tape_of
    args, dev, world, rank = ctx.args, ctx.dev, ctx.world, ctx.rank
    W, H = wl["width"], wl["height"]
    scene = wl["scene"]()
    modes, bounces = wl["modes"], wl["bounces"]
    mode_mask = sum(1 << m for m in modes)
    needs_reduce = any(m in (1, 2, 5) for m in modes)

    # ---- what this rank renders --------------------------------------------------------------
    shard, n_cam = {}, 1
    if wl["sharding"] == "strips":                      # ONE image, row-interleaved strips (strong)
        q, t = ring_pose(0, 1, wl["cam_t"], dev)
        total_rays = W * H
        if world > 1:
            shard = parallel.shard_strips(total_rays, rank, world, STRIP_ROWS * W)
        my_rays = shard["n_rays"] if shard else total_rays
        scaling, views, how = "strong", 1, (f"one frame in row-interleaved strips of {STRIP_ROWS} rows "
                                            f"(parallel.shard_strips), no collective" if world > 1 else "one GPU")
    elif wl["sharding"] == "view-batch":                # fixed batch of views, round-robin (strong)
        mine = parallel.shard_views(wl["views"], rank, world)
        n_cam = len(mine)
        poses = [ring_pose(j, wl["ring"], wl["cam_t"], dev) for j in mine]
        q, t = torch.cat([p[0] for p in poses]), torch.cat([p[1] for p in poses])
        total_rays, my_rays = W * H * wl["views"], W * H * n_cam
        scaling, views, how = "strong", wl["views"], f"{wl['views']} views dealt round-robin (parallel.shard_views)"
    else:                                               # one view per rank (weak)
        q, t = ring_pose(rank, world, wl["cam_t"], dev)
        total_rays, my_rays = W * H * world, W * H
        scaling, views, how = "weak", world, "one view per GPU"
    loop = RenderLoop(scene, num_cameras=n_cam, px_width=W, px_height=H, focal_length=PX * H,
                      sensor_width=PX * W, sensor_height=PX * H, normals_eps=EPS).to(dev)
    loop.cull = bool(args.cull)
    tape = tape_of(loop.scene)
    fpr = flops_per_ray(tape, wl)

    spec_on, spec_note = False, "tape interpreter"
    if args.specialise == "on" and not args.cull:
        from ray_marching_b200 import specialise as SPEC
        if tape.n_instr <= SPEC.MAX_SPEC_INSTR:
            try:
                loop.specialise(True)
                spec_on, spec_note = True, "tape-specialised (RenderLoop.specialise())"
            except Exception as exc:                                            # noqa: BLE001
                spec_note = f"tape interpreter (specialisation unavailable: {str(exc)[:120]})"

    def frame(qd, td, out, collective=True):
        """Kernels: [reduce_init], render, [finalize]; collective: one 8-float MAX all-reduce when a
        normalising shader runs on a sharded batch (shader.py:35-36,84: global min/max)."""
        out = loop.render_planes(qd, td, modes=modes, marching_steps=STEPS, bounces=bounces,
                                 finalize=False, out=out, **shard)
        if needs_reduce:
            if collective:
                parallel.allreduce_shader_range(out.reduce)
            ops.shade_finalize(out, mode_mask)
        return out

    launches_per_step = 3 if needs_reduce else 1
    rec = {}

    if wl["heavy"]:
        # ---- long frames: eager launches, L2 flush INSIDE the timed region -------------------------
        buf = [None]

        def step(i):
            buf[0] = frame(q, t, buf[0])
        for i in range(max(warmup, 3)):
            step(i)
        dev_ms, wall_ms = time_total(ctx, steps, step, flush=True)
        l2 = ("flushed before every step INSIDE the timed region (192 MB memset, ~0.05 ms, included in "
              "ms_per_step)")
        method = "one CUDA-event pair around all steps, eager launches"
        keep = buf
    else:
        # ---- short frames: M frames pipelined inside ONE CUDA graph ------------------------------
        # frame j's {8-float MAX all-reduce, finalize} run on a forked branch under the render of frame
        # j+1; outputs rotate through M buffers.  Host work per step = 1/M graph launch.
        M = max(m for m in (4, 2, 1) if steps % m == 0)
        bufs = [frame(q, t, None) for _ in range(M)]            # also initialises the communicator
        for b in range(M):
            bufs[b] = frame(q, t, bufs[b])
        torch.cuda.synchronize()
        side = torch.cuda.Stream(dev)
        graph, graph_note = None, None
        try:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, capture_error_mode="thread_local"):
                main = torch.cuda.current_stream(dev)
                for b in range(M):
                    loop.render_planes(q, t, modes=modes, marching_steps=STEPS, bounces=bounces,
                                       finalize=False, out=bufs[b], **shard)
                    if needs_reduce:
                        side.wait_stream(main)
                        with torch.cuda.stream(side):
                            parallel.allreduce_shader_range(bufs[b].reduce)
                            ops.shade_finalize(bufs[b], mode_mask)
                if needs_reduce:
                    main.wait_stream(side)
            graph = g
            graph_note = (f"{M} frames per CUDA-graph replay; frame j's all-reduce + finalize on a forked "
                          f"branch under the render of frame j+1")
        except Exception as exc:                                                 # noqa: BLE001
            torch.cuda.synchronize()
            graph_note = f"eager launches (graph capture failed: {str(exc)[:160]})"
        if graph is not None:
            def step(i):
                if i % M == 0:
                    graph.replay()
        else:
            def step(i):
                bufs[i % M] = frame(q, t, bufs[i % M])
        for i in range(max(warmup, 3) * M):
            step(i)
        dev_ms, wall_ms = time_total(ctx, steps, step)
        ring_mb = sum(x.numel() * 4 for x in bufs[0].images if x is not None) * M / 1e6
        l2 = (f"no flush: the path reads no per-ray input; outputs rotate through a ring of {M} buffers = "
              f"{ring_mb:.0f} MB ({'>' if ring_mb > 126 else '<'} 126 MB L2)")
        method = "one CUDA-event pair around all steps; " + graph_note
        keep = bufs
    rank_ms = ctx.gather(dev_ms / steps)
    rank_wall = ctx.gather(wall_ms / steps)

    # ---- dominant kernel alone (render_kernel) --------------------------------------------------
    res0 = keep[0]
    poses_k = ops.pose_records(q, t)
    fm = ops.fast_math_level(loop.fast_math, loop.cull)

    def launch_render():                               # `reduce` given: no init launch inside the pair
        ops.render(tape, loop.camera.struct(), poses_k, steps=STEPS, normals_eps=EPS, mode_mask=mode_mask,
                   bounces=bounces, early_exit=loop.early_exit, fast_math=fm, cmap=loop.shader.cyclic_cmap,
                   finalize=False, reduce=res0.reduce if needs_reduce else None, out=res0,
                   ray_begin=shard.get("ray_begin", 0), n_rays=shard.get("n_rays"),
                   strip_rays=shard.get("strip_rays", 0), strip_stride=shard.get("strip_stride", 0))
    reps = max(3, min(steps, 10 if wl["heavy"] else 20))
    kern_ms = kernel_alone(ctx, launch_render, reps)
    kern_all = ctx.gather(kern_ms)

    extras = {}
    # the same frame with bounding-sphere culling of smooth-union children
    if tape._cull_union is not None and not loop.cull and (headline or wl["name"] == "c5"):
        notes = {"culling_exact": "whole job, RenderLoop.cull = 'exact': a smooth union under a hard union is skipped "
                                  "where its bounding spheres prove, for every ray of the warp, that it loses the "
                                  "min -- pixels bit-identical to the headline (tests); executed < algorithmic "
                                  "work, so no roofline fraction is quoted",
                 "culling": "whole job, RenderLoop.cull = True: additionally smooth-union children whose term is "
                            "provably < 2^-24 of the largest are skipped (march steps and fast stencil samples); "
                            "executed < algorithmic work, so no roofline fraction is quoted"}
        for key, mode in (("culling_exact", "exact"), ("culling", True)):
            loop.cull = mode
            cbuf = [None]

            def cstep(i):
                cbuf[0] = frame(q, t, cbuf[0])
            for i in range(2):
                cstep(i)
            n_c = max(2, min(steps, 5))
            c_ms, _ = time_total(ctx, n_c, cstep, flush=True)
            loop.cull = False
            c_all = ctx.gather(c_ms / n_c)
            extras[key] = {"ms_per_step": max(c_all), "value": total_rays / (max(c_all) * 1e-3) / 1e6,
                           "unit": "Mrays/s", "per_rank_ms": Ctx.stats(c_all), "note": notes[key]}
            del cbuf
    # contiguous vs interleaved sharding of the single image (per-rank kernel time)
    if headline and wl["sharding"] == "strips" and world > 1:
        a, b = parallel.shard_range(W * H, rank, world)
        cont = ops.RenderResult()

        def launch_cont():
            ops.render(tape, loop.camera.struct(), poses_k, steps=STEPS, normals_eps=EPS, mode_mask=mode_mask,
                       bounces=bounces, fast_math=fm, cmap=loop.shader.cyclic_cmap, finalize=False, out=cont,
                       ray_begin=a, n_rays=b - a)
        launch_cont()
        c_all = ctx.gather(kernel_alone(ctx, launch_cont, 3))
        extras["sharding_compare"] = {"interleaved_kernel_ms": Ctx.stats(kern_all),
                                      "contiguous_kernel_ms": Ctx.stats(c_all)}
        del cont
    # the same frame through the general interpreter when the headline ran the specialised kernel
    if spec_on:
        loop.specialise(False)
        x_ms = kernel_alone(ctx, launch_render, reps)
        loop.specialise(True)
        extras["interpreter"] = {"kernel_ms": x_ms, "frac": fpr * my_rays / (x_ms * 1e-3) / 1e12 / ctx.peak["tflops"],
                                 "note": "per GPU, same frame through the general tape interpreter (the "
                                         "default of RenderLoop)"}

    e2e = measure_e2e(ctx, frame, wl, q, t, my_rays, steps) if headline else None
    e2e_all = ctx.gather(e2e["ms_total"]) if e2e else None

    if rank == 0:
        ms_step = max(rank_ms)
        value = total_rays / (ms_step * 1e-3) / 1e6
        kmax = max(kern_all)
        # strong scaling: every rank's kernel covers its share; the job's achieved rate = all flops / slowest
        ach = fpr * (total_rays if scaling == "strong" else my_rays) / (kmax * 1e-3) / 1e12
        peak_all = ctx.peak["tflops"] * (world if scaling == "strong" else 1)
        traffic, traffic_src = ncu_traffic("render_kernel", wl["name"])
        mufu = mufu_per_ray(tape, wl) * my_rays / (max(kern_all) * 1e-3)
        rec = {
            "metric": "Mrays/s fwd", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": steps,
            "warmup": max(warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": scaling,
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{wl['name']}: {wl['desc']}", "rays_per_step": total_rays,
                       "rays_per_gpu_per_step": my_rays, "marching_steps": STEPS, "views": views,
                       "sharding": how, "l2": l2, "timing": method, "cull": bool(args.cull),
                       "tape": tape.describe(), "render_kernel": spec_note,
                       "msteps_per_s": value * STEPS * (bounces + 1),
                       "per_rank_ms_per_step": Ctx.stats(rank_ms), "wall_ms_per_step": Ctx.stats(rank_wall)},
            "gpu_launches": launches_per_step * steps,
            "roofline": {"bound": "fp32", "kernel": "render_kernel", "achieved": ach, "peak": peak_all,
                         "unit": "TFLOP/s", "frac": ach / peak_all, "traffic": traffic,
                         "traffic_from": traffic_src, "traffic_unit": "bytes of DRAM per launch",
                         "kernel_ms": kmax, "kernel_ms_per_rank": Ctx.stats(kern_all),
                         "algorithmic_flops_per_ray": fpr,
                         "mufu_frac": mufu / (16 * 148 * 1.965e9),
                         "mufu_note": "special-function ops the formulas need (1 per sqrt/exp/log/div) x rays / "
                                      "kernel time, over 16 per clock per SM x 148 SMs x 1965 MHz",
                         "peak_source": PEAK_SOURCE + ("; x n_gpus for the sharded frame" if peak_all != ctx.peak["tflops"] else "")},
        }
        if e2e is not None:
            e_ms = max(e2e_all) / steps
            rec["e2e"] = {"value": total_rays / (e_ms * 1e-3) / 1e6, "unit": "Mrays/s",
                          "h2d_bytes_per_step": e2e["h2d"] * world, "d2h_bytes_per_step": e2e["d2h_all"],
                          "ms_per_step": e_ms,
                          "d2h_GBps": e2e["d2h_all"] / (e_ms * 1e-3) / 1e9,
                          "note": "pose from pinned host memory in, every image plane back to pinned host "
                                  "memory, double-buffered, wall clock, max over ranks"}
        rec.update(extras)
    del keep
    torch.cuda.empty_cache()
    return rec


def measure_e2e(ctx, frame, wl, q, t, my_rays, steps):
    """Public API with HOST buffers: every step copies the pose from pinned host memory, renders, and
    copies every image plane back to pinned host memory.  Two result buffers and a copy stream let
    the device->host copy of frame i overlap the render of frame i+1 (ordinary double buffering; all
    copies have landed when the timed region ends)."""
    from ray_marching_b200 import ops
    dev, modes = ctx.dev, wl["modes"]
    q_h, t_h = q.cpu().pin_memory(), t.cpu().pin_memory()
    host = [{m: torch.empty(my_rays, ops.MODE_CHANNELS[m]).pin_memory() for m in modes} for _ in range(2)]
    d2h = sum(h.numel() * 4 for h in host[0].values())
    h2d = (q_h.numel() + t_h.numel()) * 4
    main, side = torch.cuda.current_stream(dev), torch.cuda.Stream(dev)
    bufs, copied = [None, None], [None, None]

    def step(i):
        b = i & 1
        if copied[b] is not None:
            main.wait_event(copied[b])               # buffer b is free once its last copy finished
        qd, td = q_h.to(dev, non_blocking=True), t_h.to(dev, non_blocking=True)
        bufs[b] = frame(qd, td, bufs[b])
        rendered = torch.cuda.Event()
        rendered.record(main)
        with torch.cuda.stream(side):
            side.wait_event(rendered)
            for m in modes:
                host[b][m].copy_(bufs[b].images[m], non_blocking=True)
            copied[b] = torch.cuda.Event()
            copied[b].record(side)

    for i in range(3):
        step(i)
    torch.cuda.synchronize()
    ctx.barrier()
    t0 = time.perf_counter()
    for i in range(steps):
        step(i)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3
    ctx.barrier()
    all_d2h = ctx.gather(d2h)
    return {"ms_total": ms, "h2d": h2d, "d2h": d2h, "d2h_all": int(sum(all_d2h))}


# ---------------------------------------------------------------------------- GPU arm: fwd + bwd workloads
def run_backward(ctx, wl, steps, warmup, headline):
    """c4 / c4b: one Adam step of inverse rendering per step -- the fused loss kernel (camera, march
    replay, stencil, lambertian + raw depth, MSE residuals against a target rendered from perturbed
    parameters, reverse march, camera-pose contraction: ONE launch -- two for long tapes, whose replay is a
    kernel of its own -- RenderLoop.inverse_loss), the
    autograd scatter of d loss / d par to the nn.Parameters, the NCCL SUM all-reduce of the flat
    gradient (one view per rank), Adam -- captured as one CUDA graph."""
    from ray_marching_b200 import RenderLoop, ops, parallel
    from ray_marching_b200.frames import GraphedStep
    from ray_marching_b200.tape import tape_of
    args, dev, world, rank = ctx.args, ctx.dev, ctx.world, ctx.rank
    W, H = wl["width"], wl["height"]
    scene = wl["scene"]()
    loop = RenderLoop(scene, num_cameras=1, px_width=W, px_height=H, focal_length=PX * H,
                      sensor_width=PX * W, sensor_height=PX * H, normals_eps=EPS).to(dev)
    tape = tape_of(loop.scene)
    rays = W * H
    spec_on, spec_note = False, "tape interpreter"
    if args.specialise == "on":
        from ray_marching_b200 import specialise as SPEC
        if tape.n_instr <= SPEC.MAX_SPEC_INSTR_BWD:
            try:
                loop.specialise(True)
                spec_on, spec_note = True, "tape-specialised (RenderLoop.specialise())"
            except Exception as exc:                                            # noqa: BLE001
                spec_note = f"tape interpreter (specialisation unavailable: {str(exc)[:120]})"
    params = list(scene.parameters())
    q0, t0 = ring_pose(rank, world, wl["cam_t"], dev)
    q = q0.clone().requires_grad_(True)
    t = t0.clone().requires_grad_(True)

    # target: the same view rendered from a seeded perturbation of the parameters (p + 0.05 randn, seed 1)
    gen = torch.Generator().manual_seed(1)
    saved = [p.detach().clone() for p in params]
    with torch.no_grad():
        for p in params:
            p.add_(0.05 * torch.randn(p.shape, generator=gen).to(dev))
        res = loop.render_planes(q0, t0, gbuffer=True, modes=(0,), marching_steps=STEPS)
        lam_t = res.images[0].clone()
        depth_t = (res.pos - res.origin).norm(dim=-1, keepdim=True)
        for p, s0 in zip(params, saved):
            p.copy_(s0)
    del res
    try:                                   # one multi-tensor kernel per step instead of ~12 foreach launches
        opt = torch.optim.Adam(params + [q, t], lr=1e-2, fused=True, capturable=True)
        adam_note = "fused"
    except Exception:                                                            # noqa: BLE001
        opt = torch.optim.Adam(params + [q, t], lr=1e-2, foreach=True, capturable=True)
        adam_note = "foreach"

    def fused_fb():
        loss = loop.inverse_loss(q, t, lam_t, depth_t, marching_steps=STEPS)
        loss.backward()
        return loss

    def composed_fb():                 # the round-1 route: G-buffer planes + torch shader/loss algebra
        gb = loop.gbuffer(q, t, STEPS)
        lam = (-(gb["dirs"] * gb["normals"]).sum(-1, keepdim=True)).clamp(0, 1).reshape(-1, 1)
        depth = (gb["pos"] - gb["origin"]).norm(dim=-1, keepdim=True).reshape(-1, 1)
        loss = ((lam - lam_t) ** 2).mean() + ((depth - depth_t) ** 2).mean()
        loss.backward()
        return loss

    def allreduce():                   # one NCCL SUM over the flat gradient buffer (scene params + q, t)
        parallel.allreduce_gradients(params + [q, t], average=True)

    def fb_and_allreduce():
        loss = fused_fb()
        allreduce()
        return loss

    def eager_step(fb=fused_fb):
        opt.zero_grad(set_to_none=True)
        loss = fb()
        if world > 1:
            allreduce()
        opt.step()
        return loss

    state = {"graph": None, "note": "eager launches"}
    if args.graph:
        try:                                   # whole step incl. the NCCL all-reduce in ONE graph
            state["graph"] = GraphedStep(fb_and_allreduce if world > 1 else fused_fb, opt)
            state["note"] = ("whole step (fused loss+backward kernel, gradient scatter" +
                             (", NCCL all-reduce" if world > 1 else "") + ", Adam) = one CUDA-graph replay")
        except Exception as exc:                                                 # noqa: BLE001
            torch.cuda.synchronize()
            try:
                state["graph"] = GraphedStep(fused_fb, opt, between=allreduce if world > 1 else None)
                state["note"] = ("two CUDA-graph replays around one eager NCCL all-reduce (single-graph capture "
                                 f"failed: {str(exc)[:100]})")
            except Exception as exc2:                                            # noqa: BLE001
                torch.cuda.synchronize()
                state["graph"] = None
                state["note"] = f"eager launches (graph capture failed: {str(exc2)[:160]})"

    def step(i=0):
        state["loss"] = state["graph"]() if state["graph"] is not None else eager_step()
        return state["loss"]

    for i in range(max(warmup, 3)):
        step(i)
    dev_ms, wall_ms = time_each(ctx, steps, step)
    rank_ms = ctx.gather(dev_ms / steps)
    rank_wall = ctx.gather(wall_ms / steps)

    # dominant kernel alone: rm_render_loss_bwd with this loss
    par = tape.lowered_params(dev).detach()
    poses = ops.pose_records(q.detach(), t.detach())
    cam = loop.camera.struct()
    code = tape.code_tensor(dev)

    def launch_loss():
        ops._render_loss(code, par, poses, cam.n_views, cam.width, cam.height, cam.focal_length,
                         cam.sensor_width, cam.sensor_height, STEPS, EPS, lam_t, depth_t, 1.0 / rays, 1.0 / rays,
                         0, rays, tape.n_nodes, tape.max_depth, ops.BWD_CONVERGE_TOL, True)
    launch_loss()
    kern_ms = kernel_alone(ctx, launch_loss, 5)
    kern_all = ctx.gather(kern_ms)

    extras = {}
    # the composed route (forward G-buffer kernel + ~60 ATen launches + render_bwd kernel), eager vs fused
    if world == 1:
        for i in range(2):
            eager_step(composed_fb)
        c_ms, _ = time_each(ctx, max(3, min(steps, 10)), lambda i: eager_step(composed_fb))
        f_ms, _ = time_each(ctx, max(3, min(steps, 10)), lambda i: eager_step(fused_fb))
        n_x = max(3, min(steps, 10))
        extras["routes_eager"] = {"composed_ms_per_step": c_ms / n_x, "fused_ms_per_step": f_ms / n_x,
                                  "note": "eager launches, same optimiser: G-buffer planes + torch loss algebra + "
                                          "rm_render_bwd (round 1) vs RenderLoop.inverse_loss (one ABI call)"}

    def e2e_loop():
        host = torch.empty(1).pin_memory()
        for _ in range(2):
            host.copy_(step().detach().reshape(1), non_blocking=True)
            torch.cuda.synchronize()
        ctx.barrier()
        t0_ = time.perf_counter()
        for _ in range(steps):
            host.copy_(step().detach().reshape(1), non_blocking=True)      # loss read back every step
            torch.cuda.synchronize()
        ms = (time.perf_counter() - t0_) * 1e3
        ctx.barrier()
        return ms
    e2e_all = ctx.gather(e2e_loop()) if headline else None

    rec = {}
    if rank == 0:
        F = tape.flops_per_eval
        fwd_flops = 63 + STEPS * (F + 6) + F + 4 * F + 45 + 8 + 10
        bwd_flops = 2.0 * (63 + STEPS * (F + 6) + F + 4 * F + 45)   # 2x the differentiated forward region
        ms_step = max(rank_ms)
        value = rays * world / (ms_step * 1e-3) / 1e6
        kmax = max(kern_all)
        alg = (fwd_flops + bwd_flops) * rays          # the kernel does the forward AND the backward
        ach = alg / (kmax * 1e-3) / 1e12
        # which loss kernels run (csrc/rm_backward.cu: loss_route): long tapes replay in a kernel of their own
        split = not spec_on and tape.n_instr >= 24
        kname = "loss_replay_kernel + render_loss_reverse_kernel" if split else "render_loss_bwd_kernel"
        traffic, traffic_src = ncu_traffic(kname, wl["name"])
        rec = {
            "metric": "Mrays/s fwd+bwd", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": steps,
            "warmup": max(warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{wl['name']}: {wl['desc']}", "rays_per_gpu_per_step": rays,
                       "marching_steps": STEPS, "views": world, "sharding": "one view per GPU, replicated "
                       "parameters, NCCL SUM all-reduce of the flat gradient" if world > 1 else "one GPU",
                       "l2": "flushed between timed steps (192 MB memset, outside the per-step event pairs; "
                             "every kernel and collective of a step is launched on the timed stream)",
                       "timing": "per-step CUDA-event pairs", "step": state["note"], "tape": tape.describe(),
                       "kernels": spec_note, "adam": adam_note, "n_parameters": sum(p.numel() for p in params) + 7,
                       "per_rank_ms_per_step": Ctx.stats(rank_ms), "wall_ms_per_step": Ctx.stats(rank_wall),
                       "final_loss": float(state["loss"].detach())},
            "gpu_launches": (2 if split else 1) * steps,
            "roofline": {"bound": "fp32", "kernel": kname, "achieved": ach,
                         "peak": ctx.peak["tflops"], "unit": "TFLOP/s", "frac": ach / ctx.peak["tflops"],
                         "traffic": traffic, "traffic_from": traffic_src, "traffic_unit": "bytes of DRAM per launch",
                         "kernel_ms": kmax, "algorithmic_flops_per_ray": fwd_flops + bwd_flops,
                         "convention": "forward flops + 2 x the differentiated forward region (SURVEY 8(d)); the "
                                       "replay / recompute the kernel executes beyond that is not counted",
                         "step_algorithmic_tflops": (fwd_flops + bwd_flops) * rays / (ms_step * 1e-3) / 1e12,
                         "peak_source": PEAK_SOURCE},
        }
        if e2e_all is not None:
            e_ms = max(e2e_all) / steps
            rec["e2e"] = {"value": rays * world / (e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 0,
                          "d2h_bytes_per_step": 4 * world, "ms_per_step": e_ms,
                          "note": "optimisation state lives on the device; the loss is read back to pinned host "
                                  "memory every step (synchronising)"}
        rec.update(extras)
    state["graph"] = None
    torch.cuda.empty_cache()
    return rec


# ---------------------------------------------------------------------------- CPU arm
def _oracle_frame(wl, width, height, threads):
    """One frame of the workload on the CPU oracle port (float32, eager) at width x height."""
    from oracle import sdf_oracle as O
    from ray_marching_b200.rendering.shader import load_cmap
    torch.set_num_threads(threads)
    scene = wl["scene"]()
    orc = O.RenderLoopOracle(scene, 1, width, height, PX * wl["height"], PX * wl["width"],
                             PX * wl["height"], EPS, cmap=load_cmap())
    q = torch.tensor([[1.0, 0.0, 0.0, 0.0]])
    t = torch.tensor([list(wl["cam_t"])])

    def frame():
        if wl["backward"]:
            params = list(scene.parameters())
            qq, tt = q.clone().requires_grad_(True), t.clone().requires_grad_(True)
            gb = orc.gbuffer(qq, tt, STEPS)
            lam = O.shade_lambertian(gb["dirs"], gb["normals"])
            depth = (gb["pos"] - gb["px"]).norm(dim=-1, keepdim=True)
            loss = ((lam - 0.5) ** 2).mean() + ((depth - 0.3) ** 2).mean()
            torch.autograd.grad(loss, [qq, tt] + params, allow_unused=True)
            return
        with torch.no_grad():
            px, frames, pos0, dirs = O.camera_rays(*orc.base, q, t)
            if wl["bounces"]:
                pos, dirs, _, normals, lap = O.reflected_march(scene, pos0, dirs, STEPS, EPS, wl["bounces"])
                dists = O.sdf(scene, pos)
            else:
                pos = O.march(scene, pos0, dirs, STEPS)
                dists = O.sdf(scene, pos)
                normals, lap = O.normals_laplacian(scene, pos, EPS)
            for m in wl["modes"]:
                O.shade(px, q, frames, dirs, pos, normals, lap, dists, m, 1, orc.cmap)
    return frame


def cpu_sample_size(wl):
    """Bounded sample of the same workload: full frame for the small scenes, a centred sub-image
    (same camera, same per-ray work; rays are independent) for the large ones."""
    if wl["name"] in ("c1", "c2"):
        return wl["width"], wl["height"], "full frame"
    if wl["name"] == "c4":
        return 512, 512, "512x512 sub-sample of the 1024x1024 frame (same intrinsics per pixel), fwd+bwd"
    if wl["name"] == "c4b":
        return 96, 96, "96x96 sub-sample of the 1024x1024 frame (same intrinsics per pixel), fwd+bwd"
    if wl["name"] == "c3":
        return 192, 108, "192x108 sub-sample of the 4K frame, all 3 march segments"
    return 64, 36, "64x36 sub-sample of one 4K view"


def cpu_baseline(wl, budget_s=30.0, steps=1, warmup=1):
    threads = os.cpu_count() or 1
    w, h, sample = cpu_sample_size(wl)
    frame = _oracle_frame(wl, w, h, threads)
    for _ in range(warmup):
        frame()
    best, spent, n = float("inf"), 0.0, 0
    while n < max(steps, 1) or (n < 3 and spent < budget_s / 2):
        t0 = time.perf_counter()
        frame()
        dt = time.perf_counter() - t0
        best, spent, n = min(best, dt), spent + dt, n + 1
    return {"value": w * h / best / 1e6, "unit": "Mrays/s", "cores": threads, "kind": "port",
            "sample": f"{sample}; oracle/sdf_oracle.py (reference algorithm restated, torch CPU eager "
                      f"float32), best of {n}, {best:.3f} s per frame"}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (oracle port; the
    reference is Python/PyTorch and /root/reference does not exist on the GPU box)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = workload(args.workload)
    threads = os.cpu_count() or 1
    w, h, sample = cpu_sample_size(wl)
    frame = _oracle_frame(wl, w, h, threads)
    warm = min(max(args.warmup, 1), 2)
    t0 = time.perf_counter()
    for _ in range(warm):
        frame()
    per = (time.perf_counter() - t0) / warm
    steps = max(1, min(args.steps, 5, int(45.0 / max(per, 1e-3))))      # the whole run stays within ~a minute
    t0 = time.perf_counter()
    for _ in range(steps):
        frame()
    dt = (time.perf_counter() - t0) / steps
    value = w * h / dt / 1e6
    world = int(os.environ.get("WORLD_SIZE", "1"))
    line = {
        "impl": "reference", "metric": "Mrays/s fwd+bwd" if wl["backward"] else "Mrays/s fwd",
        "value": value, "unit": "Mrays/s", "n_gpus": world,
        "steps": steps, "warmup": warm, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "weak" if wl["sharding"] == "views" else "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{wl['name']}: {wl['desc']}", "sample": sample,
                   "marching_steps": STEPS, "note": "CPU arm: host cores only, rank 0; each step is the "
                   "bounded sample, value = sample rays / time (per-ray work is uniform)"},
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": threads, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------- driver
SUB_STEPS = {"c1": (40, 5), "c2": (40, 5), "c3": (4, 2), "c4": (30, 5), "c4b": (5, 2), "c5": (2, 1)}


def compact(rec):
    """Sub-record of other_configs: the figures the headline carries, without the prose."""
    if not rec:
        return rec
    keep = {k: rec[k] for k in ("metric", "value", "unit", "ms_per_step", "steps", "warmup", "scaling", "n_gpus",
                                "gpu_launches") if k in rec}
    keep["workload"] = rec["config"]["workload"]
    keep["timing"] = rec["config"]["timing"] + "; L2: " + rec["config"]["l2"]
    keep["per_rank_ms_per_step"] = rec["config"]["per_rank_ms_per_step"]
    for k in ("step", "kernels", "render_kernel", "sharding", "final_loss", "msteps_per_s"):
        if k in rec["config"]:
            keep[k] = rec["config"][k]
    r = rec["roofline"]
    keep["roofline"] = {k: r[k] for k in ("kernel", "frac", "achieved", "peak", "kernel_ms", "algorithmic_flops_per_ray",
                                          "mufu_frac") if k in r}
    for k in ("culling", "culling_exact", "interpreter", "routes_eager"):
        if k in rec:
            keep[k] = rec[k]
    return keep


def run_gpu(args):
    ctx = Ctx(args)
    wl = workload(args.workload)
    runner = run_backward if wl["backward"] else run_forward
    sampler = ClockSampler(ctx.local)
    if ctx.rank == 0:
        sampler.start()
    line = runner(ctx, wl, args.steps, args.warmup, True)
    # the sampler ran over the whole headline (warm-up, timed steps, kernel-alone loop, e2e: seconds of load)
    clocks = sampler.stop() if ctx.rank == 0 else None
    others = {}
    if args.others:
        for name in ("c1", "c2", "c4", "c4b", "c5", "c3"):
            if name == wl["name"]:
                continue
            sub = workload(name)
            k, w = SUB_STEPS[name]
            try:
                rec = (run_backward if sub["backward"] else run_forward)(ctx, sub, k, w, False)
                others[name] = compact(rec)
            except Exception as exc:                                             # noqa: BLE001
                # all ranks fail alike (same code, same shapes); report instead of losing the headline
                others[name] = {"error": f"{type(exc).__name__}: {str(exc)[:300]}"}
                torch.cuda.synchronize()
    if ctx.rank == 0:
        line["clocks"] = clocks
        line["config"]["numa"] = ctx.numa
        if others:
            line["other_configs"] = others
        # CPU baseline: rank 0 at N = 1 only (the host cores are shared by the ranks otherwise)
        line["cpu_baseline"] = cpu_baseline(wl, budget_s=args.cpu_budget) if ctx.world == 1 else None
        print(json.dumps(line), flush=True)
    if ctx.world > 1:
        ctx.dist.barrier()
        ctx.dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=["c1", "c2", "c3", "c4", "c4b", "c5"])
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--others", type=int, default=None,
                    help="1: also run the other BASELINE configs as short sub-records (default: on for the "
                         "default workload)")
    ap.add_argument("--specialise", default="on", choices=["off", "on"],
                    help="workloads with small tapes (c1, c2, c4) run the tape-specialised kernels "
                         "(RenderLoop.specialise()); the interpreter's kernel time is reported beside them")
    ap.add_argument("--graph", type=int, default=1, help="c4/c4b: 1 = the optimisation step is a CUDA-graph replay")
    ap.add_argument("--cull", type=int, default=0,
                    help="1: make bounding-sphere culling the measured configuration (default: reported "
                         "beside the fully evaluated headline for the smooth-union workloads)")
    ap.add_argument("--numa", type=int, default=1, help="1: bind each rank to its GPU's NUMA node (best effort)")
    ap.add_argument("--cpu-budget", type=float, default=20.0)
    args = ap.parse_args()
    if args.others is None:
        args.others = 1 if args.workload == DEFAULT_WORKLOAD else 0
    if args.impl == "reference":
        return run_reference(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus != world and args.gpus > 1:
        # convenience: python bench.py --gpus N re-launches itself under torchrun
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
               f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1", "--master-port", "29533",
               os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    run_gpu(args)


if __name__ == "__main__":
    main()
