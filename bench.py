#!/usr/bin/env python
"""Benchmark of the sphere-tracing hot path (contract: see the task statement / DESIGN.md §Measurement).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c1|c2|c3|c4|c4b|c5] [--impl reference]

A *step* is one pass of the hot path over one batch of synthetic input: one rendered frame (one
view per GPU; weak scaling), or, for c4, one forward+backward of the inverse-rendering loss.
``value`` is whole-job Mrays/s with inputs resident in HBM, ``e2e`` the same through the public
API from pinned host buffers (camera poses in, images out).  The dominant kernel is timed alone
with CUDA events for the roofline line.  ``--impl reference`` times the CPU oracle port
(oracle/sdf_oracle.py -- the reference's own PyTorch algorithm restated) on the host cores.
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

# shader flop counts of SURVEY.md §8(d)
SHADER_FLOPS = {0: 8, 1: 18, 2: 7, 3: 7, 4: 6, 5: 7, 6: 20, 7: 20}
STENCIL_MODES = {0, 4, 5, 6, 7}


def workload(name: str):
    """BASELINE.json configs as concrete inputs."""
    from ray_marching_b200 import scene as S
    if name == "c1":
        return dict(name=name, scene=S.make_test_scene2, desc="make_test_scene2 640x480 distance+proximity",
                    width=640, height=480, modes=(1, 2), bounces=0, cam_t=(0.0, 0.0, 1.0), backward=False)
    if name == "c2":
        return dict(name=name, scene=S.make_test_scene2,
                    desc="make_test_scene2 1920x1080 normal+laplacian+lambertian+vignette from one march",
                    width=1920, height=1080, modes=(4, 5, 0, 3), bounces=0, cam_t=(0.0, 0.0, 1.0),
                    backward=False)
    if name == "c3":
        return dict(name=name, scene=lambda: S.make_synthetic_scene(64, 0, True),
                    desc="synthetic-64 smooth union in shell, 3840x2160, 2 reflection bounces, lambertian",
                    width=3840, height=2160, modes=(0,), bounces=2, cam_t=(0.0, 0.0, -4.0), backward=False)
    if name == "c4":
        return dict(name=name, scene=S.make_test_scene2,
                    desc="inverse rendering fwd+bwd, make_test_scene2 1024x1024, lambertian+depth MSE",
                    width=1024, height=1024, modes=(0,), bounces=0, cam_t=(0.0, 0.1, 1.0), backward=True)
    if name == "c4b":
        return dict(name=name, scene=lambda: S.make_synthetic_scene(64, 0, True),
                    desc="inverse rendering fwd+bwd, synthetic-64 in shell (650 params) 1024x1024, lambertian+depth MSE",
                    width=1024, height=1024, modes=(0,), bounces=0, cam_t=(0.0, 0.0, -4.0), backward=True)
    if name == "c5":
        return dict(name=name, scene=lambda: S.make_synthetic_scene(512, 0, True),
                    desc="synthetic-512 smooth union in shell, 3840x2160 per view, lambertian",
                    width=3840, height=2160, modes=(0,), bounces=0, cam_t=(0.0, 0.0, -4.0), backward=False)
    raise SystemExit(f"unknown workload {name}")


def flops_per_ray(tape, wl, steps=STEPS) -> float:
    """Algorithmic forward flops per ray, SURVEY.md §8(d) counting rule."""
    F = tape.flops_per_eval
    stencil = bool(set(wl["modes"]) & STENCIL_MODES) or wl["bounces"] > 0
    seg = steps * (F + 6) + F + (4 * F + 45 if stencil else 0)
    total = 63 + seg * (wl["bounces"] + 1) + sum(SHADER_FLOPS[m] for m in wl["modes"])
    if wl["bounces"]:
        total += wl["bounces"] * 20            # hit test + reflection + lift
    return float(total)


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
    import ctypes as C
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


# ---------------------------------------------------------------------------- GPU arm
def run_gpu(args):
    import torch.distributed as dist
    from ray_marching_b200 import RenderLoop, ops, parallel
    from ray_marching_b200.tape import tape_of

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    wl = workload(args.workload)
    W, H = wl["width"], wl["height"]
    scene = wl["scene"]()
    loop = RenderLoop(scene, num_cameras=1, px_width=W, px_height=H, focal_length=PX * H,
                      sensor_width=PX * W, sensor_height=PX * H, normals_eps=EPS).to(dev)
    loop.early_exit = bool(args.early_exit)
    loop.cull = bool(args.cull)
    tape = tape_of(loop.scene)
    rays = W * H                                   # per GPU and step: one view per rank (weak scaling)
    fpr = flops_per_ray(tape, wl)
    modes = wl["modes"]
    needs_reduce = any(m in (1, 2, 5) for m in modes)
    # view `rank` of a `world`-view ring; N=1 is the reference camera of the config
    q, t = ring_pose(rank, world, wl["cam_t"], dev)
    res_holder = {}

    mode_mask = sum(1 << m for m in modes)

    def frame(qd, td, out):
        """One frame.  Kernels: [reduce_init], render, [finalize]; collective: one 8-float MAX
        all-reduce when a normalising shader runs on a sharded batch."""
        out = loop.render_planes(qd, td, modes=modes, marching_steps=STEPS, bounces=wl["bounces"],
                                 finalize=False, out=out)
        if needs_reduce:
            # global min/max over the whole camera batch (shader.py:35-36,84)
            parallel.allreduce_shader_range(out.reduce)
            ops.shade_finalize(out, mode_mask)
        return out

    # sharded job with a normalising shader: frame i's {8-float MAX all-reduce, finalize} run on a side
    # stream under the render of frame i+1 (parallel.PipelinedFrames); the timed region ends with every
    # frame finalised.  --pipeline 0 keeps the collective on the critical path.
    pipe = None
    if world > 1 and needs_reduce and args.pipeline and not wl["backward"]:
        pipe = parallel.PipelinedFrames(loop, modes, marching_steps=STEPS, bounces=wl["bounces"])

    def gpu_step():
        """One frame with everything resident in HBM."""
        if pipe is not None:
            res_holder["res"] = pipe.submit(q, t)
            return res_holder["res"]
        res_holder["res"] = frame(q, t, res_holder.get("res"))
        return res_holder["res"]

    # small tapes: the tape-specialised kernel (nvcc at scene-build time, cached in build/spec/);
    # any failure (no nvcc on the box) leaves the interpreter in place and says so
    spec_on, spec_note = False, "interpreter"
    if args.specialise == "on" and not args.cull:       # forward AND backward kernels (c1, c2, c4)
        from ray_marching_b200 import specialise as SPEC
        if tape.n_instr <= (SPEC.MAX_SPEC_INSTR_BWD if wl["backward"] else SPEC.MAX_SPEC_INSTR):
            try:
                loop.specialise(True)
                spec_on, spec_note = True, "tape-specialised (RenderLoop.specialise())"
            except Exception as exc:
                spec_note = f"interpreter (specialisation unavailable: {str(exc)[:120]})"

    if wl["backward"]:
        gpu_step, extra = make_backward_step(loop, scene, q, t, dev, world, wl["name"],
                                             graph=bool(args.graph) and tape.n_instr <= 40)
    launches_per_step = (3 if needs_reduce else 1) if not wl["backward"] else extra["launches"]

    flush = torch.empty(192 * 1024 * 1024 // 4, device=dev)      # 192 MB > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    peak = measure_fp32_peak(dev)
    for _ in range(max(args.warmup, 3)):
        gpu_step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    evs = []
    barrier()
    wall0 = time.perf_counter()
    for _ in range(args.steps):
        flush.zero_()                                # L2 flush between timed iterations (untimed)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        gpu_step()
        if pipe is not None and _ == args.steps - 1:
            pipe.wait()                              # the last frame's finalize is inside the timed region
        e1.record()
        evs.append((e0, e1))
    barrier()
    wall = time.perf_counter() - wall0
    total_ms = sum(a.elapsed_time(b) for a, b in evs)

    # dominant kernel alone (render_kernel), CUDA events on the launching stream
    kern_ms = None
    if not wl["backward"]:
        ks = []
        poses_k = ops.pose_records(q, t)
        mask_k = sum(1 << m for m in modes)
        for _ in range(max(3, min(args.steps, 20))):
            flush.zero_()
            if needs_reduce:
                ops.check(ops.lib().rm_reduce_init(res_holder["res"].reduce.data_ptr(),
                                                   torch.cuda.current_stream(dev).cuda_stream), "init")
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ops.render(tape, loop.camera.struct(), poses_k, steps=STEPS, normals_eps=EPS,
                       mode_mask=mask_k, bounces=wl["bounces"], early_exit=loop.early_exit,
                       fast_math=2 if (loop.fast_math and loop.cull) else int(loop.fast_math),
                       cmap=loop.shader.cyclic_cmap, finalize=False,
                       reduce=res_holder["res"].reduce if needs_reduce else None,
                       out=res_holder["res"])
            e1.record()
            e1.synchronize()
            ks.append(e0.elapsed_time(e1))
        kern_ms = sum(ks) / len(ks)

    # the same frame with bounding-sphere culling of smooth-union children:
    # executed work < algorithmic work, results equal to 1e-5 relative depth (tests)
    cull_info = None
    if not wl["backward"] and tape._cull_union is not None and not loop.cull:
        loop.cull = True
        for _ in range(3):
            gpu_step()
        torch.cuda.synchronize()
        ce0, ce1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n_c = max(3, min(args.steps, 10))
        ce0.record()
        for _ in range(n_c):
            gpu_step()
        ce1.record()
        ce1.synchronize()
        cull_ms = ce0.elapsed_time(ce1) / n_c
        loop.cull = False
        cull_info = {"ms_per_step": cull_ms, "value": rays / (cull_ms * 1e-3) / 1e6, "unit": "Mrays/s",
                     "note": "per-GPU, bounding-sphere culling of smooth-union children inside march "
                             "steps; executed < algorithmic work, so no roofline fraction is quoted"}

    # the same frame through the general tape INTERPRETER when the headline ran the tape-specialised
    # kernel (specialise.py): reported beside it
    interp_info = None
    if spec_on:
        loop.specialise(False)
        # backward workloads: the headline step may be a captured graph (which keeps the kernels it was
        # captured with), so the interpreter comparison runs the eager step
        x_step = extra["eager_step"] if wl["backward"] else gpu_step
        for _ in range(3):
            x_step()
        torch.cuda.synchronize()
        xs = []
        for _ in range(max(3, min(args.steps, 20))):
            flush.zero_()
            xe0, xe1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            xe0.record()
            x_step()
            xe1.record()
            xe1.synchronize()
            xs.append(xe0.elapsed_time(xe1))
        x_ms = sum(xs) / len(xs)
        interp_info = {"ms_per_step": x_ms, "value": rays / (x_ms * 1e-3) / 1e6, "unit": "Mrays/s",
                       "frac_of_fp32_peak_whole_step": (None if wl["backward"] else
                                                        fpr * rays / (x_ms * 1e-3) / 1e12 / peak["tflops"]),
                       "note": "per-GPU, same frame through the general tape interpreter (the default of "
                               "RenderLoop; the headline uses RenderLoop.specialise())"}
        loop.specialise(True)

    # the same frame with the march loop's bit-identical fixed-point exit (a ray whose position stops
    # changing reproduces itself; the loop ends when the whole warp is there): executed < algorithmic
    # work, so it is reported beside the fully marched headline, without a roofline fraction
    exit_info = None
    if not wl["backward"] and not loop.early_exit:
        loop.early_exit = True
        for _ in range(3):
            gpu_step()
        torch.cuda.synchronize()
        ys = []
        for _ in range(max(3, min(args.steps, 20))):
            flush.zero_()
            ye0, ye1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ye0.record()
            gpu_step()
            ye1.record()
            ye1.synchronize()
            ys.append(ye0.elapsed_time(ye1))
        loop.early_exit = False
        y_ms = sum(ys) / len(ys)
        exit_info = {"ms_per_step": y_ms, "value": rays / (y_ms * 1e-3) / 1e6, "unit": "Mrays/s",
                     "note": "per-GPU, RenderLoop.early_exit = True: output bit-identical to the fully "
                             "marched frame (tests); executed < algorithmic work, no roofline fraction"}

    # the same frame as ONE CUDA-graph replay (reduce init + render + finalize captured once, the way
    # ray_marching_b200.frames.FrameLoop runs a display loop): c1 is launch-latency scale
    graph_info = None
    if not wl["backward"] and world == 1:
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, capture_error_mode="relaxed"):
            gpu_step()
        n_g = 200 if rays <= (1 << 22) else 5
        for _ in range(3):
            g.replay()
        ge0, ge1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ge0.record()
        for _ in range(n_g):
            g.replay()
        ge1.record()
        ge1.synchronize()
        g_ms = ge0.elapsed_time(ge1) / n_g
        graph_info = {"ms_per_frame": g_ms, "value": rays / (g_ms * 1e-3) / 1e6, "unit": "Mrays/s",
                      "replays": n_g,
                      "note": "back-to-back CUDA-graph replays of the whole frame, no L2 flush (the path "
                              "reads nothing per ray; outputs are write-only)"}
        del g

    # end-to-end: pinned host pose in, images out to pinned host memory, every step
    e2e = measure_e2e(frame, wl, q, t, dev, args, barrier) if not wl["backward"] else \
        extra["e2e"](args, barrier)

    # sustained repeat of the same step while the clock sampler runs (>= 1.5 s under load)
    t_end = time.perf_counter() + 1.5
    while time.perf_counter() < t_end:
        for _ in range(10):
            gpu_step()
        torch.cuda.synchronize()
    clocks = sampler.stop() if rank == 0 else None

    tmax = torch.tensor([total_ms, e2e["ms_total"], kern_ms or 0.0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms, kern_max = (float(x) for x in tmax.tolist())

    if rank == 0:
        ms_step = total_ms / args.steps
        value = rays * world / (ms_step * 1e-3) / 1e6
        line = {
            "metric": "Mrays/s fwd+bwd" if wl["backward"] else "Mrays/s fwd",
            "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{wl['name']}: {wl['desc']}", "rays_per_gpu_per_step": rays,
                       "marching_steps": STEPS, "views": world, "sharding": "one view per GPU",
                       "l2": "flushed between timed iterations (192 MB memset, untimed)",
                       "early_exit": bool(args.early_exit), "cull": bool(args.cull), "tape": tape.describe(),
                       "render_kernel": spec_note,
                       **({"collective": "8-float MAX all-reduce + finalize of frame i on a side stream under "
                                         "the render of frame i+1 (parallel.PipelinedFrames)"}
                          if pipe is not None else {}),
                       **({"step": extra["note"]()} if wl["backward"] else {}),
                       "msteps_per_s": value * STEPS * (wl["bounces"] + 1),
                       "wall_s_timed_region": wall},
            "e2e": {"value": rays * world / (e2e_ms / args.steps * 1e-3) / 1e6, "unit": "Mrays/s",
                    "h2d_bytes_per_step": e2e["h2d"], "d2h_bytes_per_step": e2e["d2h"]},
            "gpu_launches": launches_per_step * args.steps,
            "clocks": clocks,
        }
        if "f16_ms_total" in e2e:
            line["e2e_f16_planes"] = {
                "value": rays / (e2e["f16_ms_total"] / args.steps * 1e-3) / 1e6, "unit": "Mrays/s",
                "d2h_bytes_per_step": e2e["f16_d2h"],
                "note": "rank 0's own loop (per GPU): the e2e loop with binary16 image planes on the wire "
                        "(rm_planes_to_f16, float32 arithmetic, pixels exact to 2.4e-4); the headline "
                        "e2e keeps the reference's float32 planes"}
        if exit_info is not None:
            line["early_exit"] = exit_info
        if interp_info is not None:
            line["interpreter"] = interp_info
        if graph_info is not None:
            line["frame_graph"] = graph_info
        if cull_info is not None:
            line["culling"] = cull_info
        if kern_ms is not None:
            ach = fpr * rays / (kern_max * 1e-3) / 1e12
            traffic, traffic_src = ncu_traffic("render_kernel", wl["name"])
            line["roofline"] = {
                "bound": "fp32", "kernel": "render_kernel", "achieved": ach, "peak": peak["tflops"],
                "unit": "TFLOP/s", "frac": ach / peak["tflops"], "traffic": traffic,
                "traffic_source": traffic_src, "traffic_unit": "bytes of DRAM per launch",
                "kernel_ms": kern_max, "algorithmic_flops_per_ray": fpr,
                "peak_source": "measured: FFMA microbenchmark (rm_probe_ffma) on this GPU; "
                               "nominal 74.5 TFLOP/s at 1965 MHz",
            }
        else:
            line["roofline"] = extra["roofline"](total_ms / args.steps, peak)
        # CPU baseline: rank 0 at N = 1 only (the host cores are shared by the ranks otherwise)
        line["cpu_baseline"] = cpu_baseline(wl, budget_s=args.cpu_budget) if world == 1 else None
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def measure_e2e(frame, wl, q, t, dev, args, barrier):
    """Public API with HOST buffers: every step copies the pose from pinned host memory, renders, and
    copies every image plane back to pinned host memory.  Two result buffers and a copy stream let
    the device->host copy of frame i overlap the render of frame i+1 (ordinary double buffering; all
    copies have landed when the timed region ends)."""
    from ray_marching_b200 import ops
    modes = wl["modes"]
    q_h, t_h = q.cpu().pin_memory(), t.cpu().pin_memory()
    rays = wl["width"] * wl["height"]
    host = [{m: torch.empty(rays, ops.MODE_CHANNELS[m]).pin_memory() for m in modes} for _ in range(2)]
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

    for i in range(4):
        step(i)
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        step(i)
    torch.cuda.synchronize()
    barrier()
    out = {"ms_total": (time.perf_counter() - t0) * 1e3, "h2d": h2d, "d2h": d2h}

    # the same loop with binary16 image planes on the wire (rm_planes_to_f16): float32
    # arithmetic, pixels exact to 2.4e-4 -- reported beside the float32 headline, never instead of it
    mask = sum(1 << m for m in modes)
    host16 = [{m: torch.empty(rays, ops.MODE_CHANNELS[m], dtype=torch.float16).pin_memory() for m in modes}
              for _ in range(2)]
    half, copied16 = [None, None], [None, None]

    def step16(i):
        b = i & 1
        if copied16[b] is not None:
            main.wait_event(copied16[b])
        qd, td = q_h.to(dev, non_blocking=True), t_h.to(dev, non_blocking=True)
        bufs[b] = frame(qd, td, bufs[b])
        half[b] = ops.planes_to_half(bufs[b], mask, out=half[b])
        rendered = torch.cuda.Event()
        rendered.record(main)
        with torch.cuda.stream(side):
            side.wait_event(rendered)
            for m in modes:
                host16[b][m].copy_(half[b][m], non_blocking=True)
            copied16[b] = torch.cuda.Event()
            copied16[b].record(side)

    for i in range(4):
        step16(i)
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        step16(i)
    torch.cuda.synchronize()
    barrier()
    out["f16_ms_total"] = (time.perf_counter() - t0) * 1e3
    out["f16_d2h"] = d2h // 2
    return out


def ncu_traffic(kernel: str, workload: str):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel (bytes per launch) from the
    newest committed `ncu --set full` summary of this workload (profiles/rNN/*_<kernel>_<wl>_summary.txt);
    None when no capture of this workload is committed.  (bench.py never runs under a profiler.)"""
    import glob
    import re
    tag = {"render_kernel": "render", "render_bwd_kernel": "render_bwd"}[kernel]
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
    return (total, os.path.relpath(files[-1], ROOT)) if seen == 2 else (None, None)


def make_backward_step(loop, scene, q, t, dev, world, wl_name="c4", graph=True):
    """c4: one Adam step of inverse rendering -- forward (camera, march, stencil, lambertian + raw
    depth), MSE loss against a target rendered from perturbed parameters, backward through the
    replay kernels, all-reduce of the flat gradient over ranks (one view per rank), Adam update."""
    import torch.distributed as dist
    from ray_marching_b200 import ops, parallel
    from ray_marching_b200.tape import tape_of
    tape = tape_of(loop.scene)
    params = list(scene.parameters())
    q = q.clone().requires_grad_(True)
    t = t.clone().requires_grad_(True)

    def gbuf():
        gb = loop.gbuffer(q, t, STEPS)          # fused forward kernel; backward = rm_render_bwd
        lam = (-(gb["dirs"] * gb["normals"]).sum(-1, keepdim=True)).clamp(0, 1)
        depth = (gb["pos"] - gb["origin"]).norm(dim=-1, keepdim=True)
        return lam, depth, None

    gen = torch.Generator().manual_seed(1)
    saved = [p.detach().clone() for p in params]
    with torch.no_grad():
        for p in params:
            p.add_(0.05 * torch.randn(p.shape, generator=gen).to(dev))
    lam_t, depth_t, _ = [x.detach() if isinstance(x, torch.Tensor) else x for x in gbuf()]
    with torch.no_grad():
        for p, s0 in zip(params, saved):
            p.copy_(s0)
    opt = torch.optim.Adam(params + [q, t], lr=1e-2, foreach=True, capturable=True)
    state = {"graph": None, "graph_note": "eager launches"}

    def forward_backward():
        lam, depth, rays = gbuf()
        loss = ((lam - lam_t) ** 2).mean() + ((depth - depth_t) ** 2).mean()
        loss.backward()
        return loss

    def allreduce():           # one NCCL SUM over the flat gradient buffer (scene params + q, t)
        parallel.allreduce_gradients(params + [q, t], average=True)

    def eager_step():
        opt.zero_grad(set_to_none=True)
        loss = forward_backward()
        if world > 1:
            allreduce()
        opt.step()
        return loss

    # the step as CUDA-graph replays (frames.GraphedStep) when the kernels are short enough for launch
    # overhead to matter (registry scenes): one graph at N = 1; at N > 1 two graphs (forward+backward,
    # Adam) with the NCCL all-reduce launched eagerly between them
    if graph:
        try:
            from ray_marching_b200.frames import GraphedStep
            state["graph"] = GraphedStep(forward_backward, opt, between=allreduce if world > 1 else None)
            state["graph_note"] = ("whole step (forward, loss, backward, Adam) = one CUDA-graph replay"
                                   if world == 1 else
                                   "two CUDA-graph replays (forward+loss+backward | Adam) around one eager "
                                   "NCCL all-reduce of the flat gradient buffer")
        except Exception as exc:        # capture refused: keep the eager step and say so
            torch.cuda.synchronize()
            state["graph"] = None
            state["graph_note"] = f"eager launches (graph capture failed: {str(exc)[:160]})"

    def step():
        loss = state["graph"]() if state["graph"] is not None else eager_step()
        state["loss"] = loss
        return loss

    rays_n = loop.px_width * loop.px_height
    F = tape.flops_per_eval
    fwd_flops = 63 + STEPS * (F + 6) + F + 4 * F + 45 + 8 + 10

    def roofline(ms_step, peak):
        # dominant kernel: render_bwd_kernel, timed alone with the upstream gradients of this loss
        par = tape.lowered_params(dev).detach()
        poses = ops.pose_records(q.detach(), t.detach())
        cam = loop.camera.struct()
        g3 = torch.randn(rays_n, 3, device=dev)
        code = tape.code_tensor(dev)
        ks = []
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ops._render_bwd(code, par, poses, cam.n_views, cam.width, cam.height, cam.focal_length,
                            cam.sensor_width, cam.sensor_height, STEPS, EPS, g3, g3, g3, None, g3, None,
                            tape.n_nodes, tape.max_depth, ops.BWD_CONVERGE_TOL)
            e1.record()
            e1.synchronize()
            ks.append(e0.elapsed_time(e1))
        kms = sum(ks[1:]) / len(ks[1:])
        bwd_flops = 2.0 * (63 + STEPS * (F + 6) + F + 4 * F + 45)   # 2x the differentiated forward region
        alg = bwd_flops * rays_n
        ach = alg / (kms * 1e-3) / 1e12
        traffic, traffic_src = ncu_traffic("render_bwd_kernel", wl_name)
        return {"bound": "fp32", "kernel": "render_bwd_kernel", "achieved": ach, "peak": peak["tflops"],
                "unit": "TFLOP/s", "frac": ach / peak["tflops"], "traffic": traffic,
                "traffic_source": traffic_src, "traffic_unit": "bytes of DRAM per launch", "kernel_ms": kms,
                "algorithmic_flops_per_ray": bwd_flops,
                "step_algorithmic_tflops": 3.0 * fwd_flops * rays_n / (ms_step * 1e-3) / 1e12,
                "peak_source": "measured: FFMA microbenchmark (rm_probe_ffma) on this GPU; "
                               "nominal 74.5 TFLOP/s at 1965 MHz"}

    def e2e(args, barrier):
        host = torch.empty(1).pin_memory()
        for _ in range(2):
            host.copy_(step().detach().reshape(1), non_blocking=True)
            torch.cuda.synchronize()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            host.copy_(step().detach().reshape(1), non_blocking=True)      # loss read back every step
            torch.cuda.synchronize()
        barrier()
        return {"ms_total": (time.perf_counter() - t0) * 1e3, "h2d": 0, "d2h": 4}

    # my kernels per step: render_kernel (G-buffer forward), render_bwd_kernel
    return step, {"launches": 2, "roofline": roofline, "e2e": e2e, "eager_step": eager_step,
                  "note": lambda: state["graph_note"]}


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
    for _ in range(min(args.warmup, 2)):
        frame()
    t0 = time.perf_counter()
    steps = max(1, min(args.steps, 5))
    for _ in range(steps):
        frame()
    dt = (time.perf_counter() - t0) / steps
    value = w * h / dt / 1e6
    line = {
        "impl": "reference", "metric": "Mrays/s fwd+bwd" if wl["backward"] else "Mrays/s fwd",
        "value": value, "unit": "Mrays/s", "n_gpus": int(os.environ.get("WORLD_SIZE", "1")),
        "steps": steps, "warmup": min(args.warmup, 2), "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": f"{wl['name']}: {wl['desc']}", "sample": sample,
                   "marching_steps": STEPS, "note": "CPU arm: host cores only, rank 0"},
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": threads, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--specialise", default="on", choices=["off", "on"],
                    help="workloads with small tapes (c1, c2, c4) run the tape-specialised kernels "
                         "(RenderLoop.specialise()); the interpreter timing is reported as an extra key")
    ap.add_argument("--graph", type=int, default=1,
                    help="c4 on a registry scene: 1 = the whole optimisation step is one CUDA-graph replay "
                         "(frames.GraphedStep), 0 = eager launches")
    ap.add_argument("--pipeline", type=int, default=1,
                    help="N > 1 with a normalising shader: 1 = all-reduce + finalize of frame i overlap the "
                         "render of frame i+1 (parallel.PipelinedFrames), 0 = in sequence")
    ap.add_argument("--workload", default="c2", choices=["c1", "c2", "c3", "c4", "c4b", "c5"])
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--early-exit", type=int, default=0,
                    help="1: bit-identical fixed-point exit of the march loop (executed < algorithmic work)")
    ap.add_argument("--cull", type=int, default=0,
                    help="1: make bounding-sphere culling the measured configuration (default: reported "
                         "beside the fully evaluated headline for the smooth-union workloads)")
    ap.add_argument("--cpu-budget", type=float, default=20.0)
    args = ap.parse_args()
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
