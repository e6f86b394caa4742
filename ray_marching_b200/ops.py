"""``torch.library`` custom ops over the C ABI (``include/rm_b200.h``).

Forward ops take the lowered tape as two tensors (``code``: HOST int32 instruction words that the
library copies into the launch parameters; ``par``: device float32) so that autograd sees ``par``
as an ordinary differentiable input; ``tape.lower_theta`` connects it to the scene's
``nn.Parameter`` objects.  Backward ops are registered with ``register_autograd``
(see the bottom of this file).  Nothing here computes on the CPU: non-CUDA tensors raise.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Tuple

import torch
from torch import Tensor

from . import _lib, specialise as _spec
from ._lib import (RmCamera, RmLossArgs, RmRenderArgs, RmRenderGrads, RmRenderOut, RmTape, check, lib, ptr,
                   stream_ptr)

# Backward kernels: step-length tolerance (relative to 1 + |p|_inf) below which a ray counts as
# settled and its remaining march steps are reversed in closed form (include/rm_b200.h,
# rm_march_bwd).  0 reverses every step of the unrolled graph.
BWD_CONVERGE_TOL = 1e-7


def set_backward_converge_tol(tol: float) -> float:
    """Sets the settled-ray tolerance of the backward kernels; returns the previous value."""
    global BWD_CONVERGE_TOL
    old, BWD_CONVERGE_TOL = BWD_CONVERGE_TOL, float(tol)
    return old


MODES = ["lambertian", "distance", "proximity", "vignette", "normal", "laplacian",
         "tangent", "spin"]
MODE_CHANNELS = [1, 1, 1, 1, 3, 1, 3, 3]
NORMALISED_MODES = (1 << 1) | (1 << 2) | (1 << 5)


def fast_math_level(fast_math, cull) -> int:
    """``RenderLoop.fast_math`` / ``.cull`` -> ``RmRenderArgs.fast_math`` (include/rm_b200.h): culling needs
    the approximate-math kernels; ``cull="exact"`` selects the skips that leave every pixel bit-identical."""
    if not fast_math:
        return 0
    if isinstance(cull, str):
        if cull != "exact":
            raise ValueError(f"cull must be False, True or 'exact', not {cull!r}")
        return 3
    return 2 if cull else 1


def _tape_struct(code: Tensor, par: Tensor, n_nodes: int, max_depth: int) -> RmTape:
    if code.is_cuda or code.dtype != torch.int32 or not code.is_contiguous():
        raise _lib.RmError("tape code must be a contiguous int32 HOST tensor (tape.code_tensor())")
    return RmTape(ptr(code), ptr(par), code.shape[0], par.numel(), max_depth, n_nodes)


def _flat3(t: Tensor) -> Tensor:
    return t.reshape(-1, 3).contiguous()


# ------------------------------------------------------------------------------ sdf_eval
@torch.library.custom_op("rm_b200::sdf_eval", mutates_args=())
def _sdf_eval(code: Tensor, par: Tensor, points: Tensor, n_nodes: int, max_depth: int) -> Tensor:
    dev = _lib.require_cuda(par, points)
    pts = _flat3(points)
    out = torch.empty(pts.shape[0], device=dev, dtype=torch.float32)
    tape = _tape_struct(code, par, n_nodes, max_depth)
    check(lib().rm_sdf_eval(C.byref(tape), ptr(pts), pts.shape[0], ptr(out), stream_ptr(dev)),
          "rm_sdf_eval")
    return out.reshape(*points.shape[:-1], 1)


@_sdf_eval.register_fake
def _(code, par, points, n_nodes, max_depth):
    return points.new_empty(*points.shape[:-1], 1)


def sdf_eval(tape, points: Tensor) -> Tensor:
    """``scene(points (...,3)) -> (...,1)`` through the tape interpreter."""
    if points.shape[-1] != 3:
        raise ValueError(f"query positions must have a trailing dimension of 3, got {tuple(points.shape)}")
    _lib.require_cuda(points)
    return _sdf_eval(tape.code_tensor(points.device), tape.lowered_params(points.device), points,
                     tape.n_nodes, tape.max_depth)


# ------------------------------------------------------------------------------ march
@torch.library.custom_op("rm_b200::march", mutates_args=())
def _march(code: Tensor, par: Tensor, origin: Tensor, dirs: Tensor, steps: int, fast: bool,
           n_nodes: int, max_depth: int) -> Tensor:
    dev = _lib.require_cuda(par, origin, dirs)
    o, v = _flat3(origin), _flat3(dirs)
    out = torch.empty_like(o)
    tape = _tape_struct(code, par, n_nodes, max_depth)
    check(lib().rm_march(C.byref(tape), ptr(o), ptr(v), o.shape[0], steps, int(fast), ptr(out),
                         stream_ptr(dev)), "rm_march")
    return out.reshape(origin.shape)


@_march.register_fake
def _(code, par, origin, dirs, steps, fast, n_nodes, max_depth):
    return torch.empty_like(origin)


def march(tape, origin: Tensor, dirs: Tensor, steps: int = 32, fast: bool = False) -> Tensor:
    """``SDFMarcher.forward``: ``p <- f(p) v + p`` repeated ``steps`` times."""
    origin, dirs = torch.broadcast_tensors(origin, dirs)
    _lib.require_cuda(origin, dirs)
    par = tape.lowered_params(origin.device)
    if torch.is_grad_enabled() and (par.requires_grad or origin.requires_grad or dirs.requires_grad):
        fast = False                    # the backward kernel replays the IEEE-math trajectory
    return _march(tape.code_tensor(origin.device), par, origin, dirs,
                  int(steps), bool(fast), tape.n_nodes, tape.max_depth)


# ------------------------------------------------------------------------------ normals
@torch.library.custom_op("rm_b200::normals", mutates_args=())
def _normals(code: Tensor, par: Tensor, points: Tensor, eps: float, n_nodes: int,
             max_depth: int) -> Tuple[Tensor, Tensor]:
    dev = _lib.require_cuda(par, points)
    pts = _flat3(points)
    n = torch.empty_like(pts)
    lap = torch.empty(pts.shape[0], device=dev, dtype=torch.float32)
    tape = _tape_struct(code, par, n_nodes, max_depth)
    check(lib().rm_normals(C.byref(tape), ptr(pts), pts.shape[0], float(eps), ptr(n), ptr(lap),
                           stream_ptr(dev)), "rm_normals")
    return n.reshape(points.shape), lap.reshape(*points.shape[:-1], 1)


@_normals.register_fake
def _(code, par, points, eps, n_nodes, max_depth):
    return torch.empty_like(points), points.new_empty(*points.shape[:-1], 1)


def normals(tape, points: Tensor, eps: float) -> Tuple[Tensor, Tensor]:
    """Fused ``SDFNormals.forward`` (no autograd; the differentiable route composes sdf_eval)."""
    _lib.require_cuda(points)
    return _normals(tape.code_tensor(points.device), tape.lowered_params(points.device), points, float(eps),
                    tape.n_nodes, tape.max_depth)


# ------------------------------------------------------------------------------ camera
def camera_struct(n_views, width, height, focal, sensor_w, sensor_h) -> RmCamera:
    return RmCamera(int(n_views), int(width), int(height), float(focal), float(sensor_w),
                    float(sensor_h))


def camera_rays(cam: RmCamera, poses: Tensor, ray_begin: int = 0,
                n_rays: Optional[int] = None) -> Tuple[Tensor, Tensor]:
    """World-space ray origins/directions ``(n,3)`` for a contiguous range of rays."""
    dev = _lib.require_cuda(poses)
    total = cam.n_views * cam.width * cam.height
    n = total - ray_begin if n_rays is None else n_rays
    origin = torch.empty(n, 3, device=dev, dtype=torch.float32)
    dirs = torch.empty(n, 3, device=dev, dtype=torch.float32)
    poses = poses.contiguous()
    check(lib().rm_camera_rays(C.byref(cam), ptr(poses), ray_begin, n, ptr(origin), ptr(dirs),
                               stream_ptr(dev)), "rm_camera_rays")
    return origin, dirs


# ------------------------------------------------------------------------------ fused render
class RenderResult:
    """Planes produced by one fused render call (all on the GPU, local-ray indexed)."""

    def __init__(self):
        self.origin = self.dirs = self.pos = self.sdf = self.normals = self.laplacian = None
        self.images: List[Optional[Tensor]] = [None] * 8
        self.rgba = None
        self.reduce = None


def render(tape, cam: RmCamera, poses: Tensor, *, steps: int = 32, normals_eps: float = 5e-2,
           mode_mask: int = 0, gbuffer: bool = False, display_mode: int = -1, degree: int = 1,
           bounces: int = 0, hit_tau: float = 1e-3, lift_delta: Optional[float] = None,
           early_exit: bool = False, fast_math: int = 1, cmap: Optional[Tensor] = None,
           ray_begin: int = 0, n_rays: Optional[int] = None, strip_rays: int = 0, strip_stride: int = 0,
           finalize: bool = True, reduce: Optional[Tensor] = None,
           out: Optional[RenderResult] = None) -> RenderResult:
    """One launch of the fused camera->march->stencil->shade kernel (no autograd).

    ``fast_math``: 0 IEEE everywhere, 1 approximate sqrt/exp inside march steps, 2 additionally
    bounding-sphere culling of smooth-union children, 3 = 1 + only the exact whole-union skips of 2.
    ``strip_rays`` / ``strip_stride``: interleaved shard (``parallel.shard_strips``; ``n_rays`` is then
    required): local ray l = global ray ``ray_begin + (l // strip_rays) * strip_stride + l % strip_rays``.
    ``finalize=False`` leaves modes 1/2/5 un-normalised (so a sharded job can all-reduce
    ``result.reduce`` first and then call :func:`shade_finalize`)."""
    dev = _lib.require_cuda(poses)
    with torch.no_grad():
        par = tape.lowered_params(dev)
        total = cam.n_views * cam.width * cam.height
        if strip_rays and n_rays is None:
            raise _lib.RmError("render: an interleaved shard needs n_rays (parallel.shard_strips)")
        n = total - ray_begin if n_rays is None else n_rays
        res = out or RenderResult()
        new = lambda *shape: torch.empty(*shape, device=dev, dtype=torch.float32)
        if out is not None:
            # planes reused from an earlier call must fit THIS ray range (the C ABI sees no sizes)
            for name in ("origin", "dirs", "pos", "sdf", "normals", "laplacian", "rgba"):
                t = getattr(res, name)
                if t is not None and (t.shape[0] != n or t.device != dev):
                    setattr(res, name, None)
            for m in range(8):
                t = res.images[m]
                if t is not None and (t.shape[0] != n or t.device != dev):
                    res.images[m] = None
            if res.reduce is not None and res.reduce.device != dev:
                res.reduce = None
        if gbuffer:
            res.origin = res.origin if res.origin is not None else new(n, 3)
            res.dirs = res.dirs if res.dirs is not None else new(n, 3)
            res.pos = res.pos if res.pos is not None else new(n, 3)
            res.sdf = res.sdf if res.sdf is not None else new(n)
            res.normals = res.normals if res.normals is not None else new(n, 3)
            res.laplacian = res.laplacian if res.laplacian is not None else new(n)
        for m in range(8):
            if mode_mask >> m & 1 and res.images[m] is None:
                res.images[m] = new(n, MODE_CHANNELS[m])
        if display_mode >= 0 and res.rgba is None:
            res.rgba = new(n, 4)
        full_mask = mode_mask | ((1 << display_mode) if display_mode >= 0 else 0)
        if full_mask & NORMALISED_MODES:
            if reduce is not None:
                res.reduce = reduce
            elif res.reduce is None:
                res.reduce = new(_lib.REDUCE_FLOATS)
            if reduce is None:
                check(lib().rm_reduce_init(ptr(res.reduce), stream_ptr(dev)), "rm_reduce_init")
        o = RmRenderOut()
        o.origin, o.dirs, o.pos = ptr(res.origin), ptr(res.dirs), ptr(res.pos)
        o.sdf, o.normals, o.laplacian = ptr(res.sdf), ptr(res.normals), ptr(res.laplacian)
        for m in range(8):
            o.image[m] = ptr(res.images[m]) if mode_mask >> m & 1 else 0
        o.rgba, o.reduce = ptr(res.rgba), ptr(res.reduce)
        poses = poses.contiguous()
        a = RmRenderArgs(ray_begin, n, int(steps), int(bounces), float(normals_eps), float(hit_tau),
                         float(2 * normals_eps if lift_delta is None else lift_delta),
                         int(mode_mask), int(display_mode), int(degree), int(early_exit),
                         int(fast_math), ptr(cmap), 0.0, int(strip_rays), int(strip_stride))
        t = _tape_struct(tape.code_tensor(dev), par, tape.n_nodes, tape.max_depth)
        # a tape-specialised library (specialise.py) exports rm_render with the same ABI
        spec = getattr(tape, "spec_lib", None) if fast_math < 2 else None
        status = (spec or lib()).rm_render(C.byref(t), C.byref(cam), ptr(poses), C.byref(a), C.byref(o),
                                           stream_ptr(dev))
        if status != 0 and spec is not None:
            raise _lib.RmError(f"rm_render (specialised) failed with status {status}: "
                               f"{spec.rm_last_error().decode(errors='replace')}")
        check(status, "rm_render")
        res._keep = (par, poses, cmap)       # keep inputs alive until the stream has consumed them
        if finalize and (full_mask & NORMALISED_MODES):
            shade_finalize(res, mode_mask, display_mode)
    return res


def shade_finalize(res: RenderResult, mode_mask: int, display_mode: int = -1) -> None:
    """Second phase of the normalising shaders (modes 1, 2, 5), in place."""
    dev = res.reduce.device
    n = next(t for t in (res.images + [res.rgba]) if t is not None).shape[0]
    imgs = (C.c_void_p * 8)(*[ptr(t) for t in res.images])
    check(lib().rm_shade_finalize(ptr(res.reduce), int(mode_mask), n, imgs, ptr(res.rgba),
                                  int(display_mode), stream_ptr(dev)), "rm_shade_finalize")


def planes_to_half(res: RenderResult, mode_mask: int, out: Optional[List[Optional[Tensor]]] = None
                   ) -> List[Optional[Tensor]]:
    """float16 copies of the (finalised) image planes of ``mode_mask`` for transport -- the
    reference's own CUDA run dtype (main.py:21); arithmetic stays float32.  One launch
    (``rm_planes_to_f16``); ``out`` reuses buffers of an earlier call."""
    first = next(t for t in res.images if t is not None)
    dev, n = first.device, first.shape[0]
    half = list(out) if out is not None else [None] * 8
    for m in range(8):
        if mode_mask >> m & 1:
            if res.images[m] is None:
                raise _lib.RmError(f"planes_to_half: mode {m} was not rendered")
            if half[m] is None:
                half[m] = torch.empty(n, MODE_CHANNELS[m], device=dev, dtype=torch.float16)
    src = (C.c_void_p * 8)(*[ptr(t) for t in res.images])
    dst = (C.c_void_p * 8)(*[ptr(t) if (mode_mask >> m & 1) else None for m, t in enumerate(half)])
    check(lib().rm_planes_to_f16(int(mode_mask), n, src, dst, stream_ptr(dev)), "rm_planes_to_f16")
    return half


def pose_records(orientations: Tensor, translations: Tensor) -> Tensor:
    """``(N,4),(N,3) -> (N,8)`` pose records ``(q, t, 0)`` (differentiable: one concatenate).  The
    kernels expand them to the rotation polynomial of ``Q.rotation`` (ray_marching.py:58-63) and
    the forward axis of ``QuaternionToSO3`` (quaternion.py:114-124)."""
    _lib.require_cuda(orientations, translations)
    pad = torch.zeros_like(orientations[:, :1])
    return torch.cat([orientations, translations, pad], dim=-1).contiguous()


# ------------------------------------------------------------------------------ backward ops
@torch.library.custom_op("rm_b200::sdf_vjp", mutates_args=())
def _sdf_vjp(code: Tensor, par: Tensor, points: Tensor, grad: Tensor, n_nodes: int,
             max_depth: int) -> Tuple[Tensor, Tensor]:
    dev = _lib.require_cuda(par, points, grad)
    pts = _flat3(points)
    g = grad.reshape(-1).contiguous()
    dpts = torch.empty_like(pts)
    dpar = torch.empty_like(par)
    tape = _tape_struct(code, par, n_nodes, max_depth)
    spec = _spec.selected_for(code)
    if spec is not None:
        _spec.spec_check(spec, spec.rm_sdf_vjp(C.byref(tape), ptr(pts), ptr(g), pts.shape[0], ptr(dpts),
                                               ptr(dpar), stream_ptr(dev)), "rm_sdf_vjp")
    else:
        check(lib().rm_sdf_vjp(C.byref(tape), ptr(pts), ptr(g), pts.shape[0], ptr(dpts), ptr(dpar),
                               stream_ptr(dev)), "rm_sdf_vjp")
    return dpar, dpts.reshape(points.shape)


@_sdf_vjp.register_fake
def _(code, par, points, grad, n_nodes, max_depth):
    return torch.empty_like(par), torch.empty_like(points)


def _sdf_eval_setup(ctx, inputs, output):
    code, par, points, n_nodes, max_depth = inputs
    ctx.save_for_backward(code, par, points)
    ctx.meta = (n_nodes, max_depth)


def _sdf_eval_backward(ctx, grad):
    code, par, points = ctx.saved_tensors
    dpar, dpts = _sdf_vjp(code, par, points, grad.contiguous(), *ctx.meta)
    return None, dpar, dpts, None, None


_sdf_eval.register_autograd(_sdf_eval_backward, setup_context=_sdf_eval_setup)


@torch.library.custom_op("rm_b200::march_bwd", mutates_args=())
def _march_bwd(code: Tensor, par: Tensor, origin: Tensor, dirs: Tensor, grad_pos: Tensor,
               steps: int, n_nodes: int, max_depth: int, tol: float) -> Tuple[Tensor, Tensor, Tensor]:
    dev = _lib.require_cuda(par, origin, dirs, grad_pos)
    o, v, g = _flat3(origin), _flat3(dirs), _flat3(grad_pos)
    n = o.shape[0]
    spec = _spec.selected_for(code)
    traj = torch.empty(max(int((spec or lib()).rm_bwd_scratch_floats(n, steps)), 1), device=dev,
                       dtype=torch.float32)
    do, dv = torch.empty_like(o), torch.empty_like(v)
    dpar = torch.empty_like(par)
    tape = _tape_struct(code, par, n_nodes, max_depth)
    if spec is not None:
        _spec.spec_check(spec, spec.rm_march_bwd(C.byref(tape), ptr(o), ptr(v), ptr(g), n, steps, float(tol),
                                                 ptr(traj), ptr(do), ptr(dv), ptr(dpar), stream_ptr(dev)),
                         "rm_march_bwd")
    else:
        check(lib().rm_march_bwd(C.byref(tape), ptr(o), ptr(v), ptr(g), n, steps, float(tol), ptr(traj),
                                 ptr(do), ptr(dv), ptr(dpar), stream_ptr(dev)), "rm_march_bwd")
    return dpar, do.reshape(origin.shape), dv.reshape(dirs.shape)


@_march_bwd.register_fake
def _(code, par, origin, dirs, grad_pos, steps, n_nodes, max_depth, tol):
    return torch.empty_like(par), torch.empty_like(origin), torch.empty_like(dirs)


def _march_setup(ctx, inputs, output):
    code, par, origin, dirs, steps, fast, n_nodes, max_depth = inputs
    if fast:
        raise _lib.RmError("march(fast=True) is not differentiable; autograd uses the IEEE-math march")
    ctx.save_for_backward(code, par, origin, dirs)
    ctx.meta = (steps, n_nodes, max_depth)


def _march_backward(ctx, grad):
    code, par, origin, dirs = ctx.saved_tensors
    steps, n_nodes, max_depth = ctx.meta
    dpar, do, dv = _march_bwd(code, par, origin, dirs, grad.contiguous(), steps, n_nodes, max_depth,
                              BWD_CONVERGE_TOL)
    return None, dpar, do, dv, None, None, None, None


_march.register_autograd(_march_backward, setup_context=_march_setup)


# ------------------------------------------------------------------------------ fused G-buffer
@torch.library.custom_op("rm_b200::render_gbuffer", mutates_args=())
def _render_gbuffer(code: Tensor, par: Tensor, poses: Tensor, n_views: int, width: int, height: int,
                    focal: float, sensor_w: float, sensor_h: float, steps: int, eps: float,
                    n_nodes: int, max_depth: int) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor, Tensor]:
    dev = _lib.require_cuda(par, poses)
    n = n_views * width * height
    new = lambda *shape: torch.empty(*shape, device=dev, dtype=torch.float32)
    origin, dirs, pos, nrm = new(n, 3), new(n, 3), new(n, 3), new(n, 3)
    sdf, lap = new(n), new(n)
    o = RmRenderOut()
    o.origin, o.dirs, o.pos, o.sdf, o.normals, o.laplacian = (ptr(origin), ptr(dirs), ptr(pos), ptr(sdf),
                                                              ptr(nrm), ptr(lap))
    poses = poses.contiguous()
    a = RmRenderArgs(0, n, int(steps), 0, float(eps), 0.0, 0.0, 0, -1, 1, 0, 0, 0, 0.0)   # IEEE math
    cam = camera_struct(n_views, width, height, focal, sensor_w, sensor_h)
    t = _tape_struct(code, par, n_nodes, max_depth)
    spec = _spec.selected_for(code)
    if spec is not None and spec.has_forward:
        _spec.spec_check(spec, spec.rm_render(C.byref(t), C.byref(cam), ptr(poses), C.byref(a), C.byref(o),
                                              stream_ptr(dev)), "rm_render")
    else:
        check(lib().rm_render(C.byref(t), C.byref(cam), ptr(poses), C.byref(a), C.byref(o),
                              stream_ptr(dev)), "rm_render")
    return origin, dirs, pos, sdf, nrm, lap


@_render_gbuffer.register_fake
def _(code, par, poses, n_views, width, height, focal, sensor_w, sensor_h, steps, eps, n_nodes, max_depth):
    n = n_views * width * height
    v3 = lambda: poses.new_empty(n, 3)
    return v3(), v3(), v3(), poses.new_empty(n), v3(), poses.new_empty(n)


@torch.library.custom_op("rm_b200::render_bwd", mutates_args=())
def _render_bwd(code: Tensor, par: Tensor, poses: Tensor, n_views: int, width: int, height: int,
                focal: float, sensor_w: float, sensor_h: float, steps: int, eps: float,
                g_origin: Optional[Tensor], g_dirs: Optional[Tensor], g_pos: Optional[Tensor],
                g_sdf: Optional[Tensor], g_normals: Optional[Tensor], g_lap: Optional[Tensor],
                n_nodes: int, max_depth: int, tol: float) -> Tuple[Tensor, Tensor]:
    dev = _lib.require_cuda(par, poses, g_origin, g_dirs, g_pos, g_sdf, g_normals, g_lap)
    n = n_views * width * height
    cont = lambda g: None if g is None else g.contiguous()
    gs = [cont(g) for g in (g_origin, g_dirs, g_pos, g_sdf, g_normals, g_lap)]
    grads = RmRenderGrads(*[ptr(g) for g in gs])
    spec = _spec.selected_for(code)                 # the scratch size belongs to the library that runs
    traj = torch.empty(max(int((spec or lib()).rm_bwd_scratch_floats(n, steps)), 1), device=dev,
                       dtype=torch.float32)
    dpar = torch.empty_like(par)
    poses = poses.contiguous()
    dposes = torch.empty_like(poses)
    a = RmRenderArgs(0, n, int(steps), 0, float(eps), 0.0, 0.0, 0, -1, 1, 0, 0, 0, float(tol))
    cam = camera_struct(n_views, width, height, focal, sensor_w, sensor_h)
    t = _tape_struct(code, par, n_nodes, max_depth)
    if spec is not None:
        _spec.spec_check(spec, spec.rm_render_bwd(C.byref(t), C.byref(cam), ptr(poses), C.byref(a),
                                                  C.byref(grads), ptr(traj), ptr(dpar), ptr(dposes),
                                                  stream_ptr(dev)), "rm_render_bwd")
    else:
        check(lib().rm_render_bwd(C.byref(t), C.byref(cam), ptr(poses), C.byref(a), C.byref(grads),
                                  ptr(traj), ptr(dpar), ptr(dposes), stream_ptr(dev)), "rm_render_bwd")
    return dpar, dposes


@_render_bwd.register_fake
def _(code, par, poses, n_views, width, height, focal, sensor_w, sensor_h, steps, eps, g_origin, g_dirs,
      g_pos, g_sdf, g_normals, g_lap, n_nodes, max_depth, tol):
    return torch.empty_like(par), torch.empty_like(poses)


def _gbuffer_setup(ctx, inputs, output):
    ctx.set_materialize_grads(False)            # unused planes arrive as None -> skipped in the kernel
    ctx.save_for_backward(inputs[0], inputs[1], inputs[2])
    ctx.meta = inputs[3:]


def _gbuffer_backward(ctx, g_origin, g_dirs, g_pos, g_sdf, g_normals, g_lap):
    code, par, poses = ctx.saved_tensors
    (n_views, width, height, focal, sw, sh, steps, eps, n_nodes, max_depth) = ctx.meta
    dpar, dposes = _render_bwd(code, par, poses, n_views, width, height, focal, sw, sh, steps, eps,
                               g_origin, g_dirs, g_pos, g_sdf, g_normals, g_lap, n_nodes, max_depth,
                               BWD_CONVERGE_TOL)
    return (None, dpar, dposes) + (None,) * 10


_render_gbuffer.register_autograd(_gbuffer_backward, setup_context=_gbuffer_setup)


def render_gbuffer(tape, cam: RmCamera, orientations: Tensor, translations: Tensor, steps: int = 32,
                   normals_eps: float = 5e-2):
    """Differentiable fused G-buffer: ``(origin, dirs, pos, sdf, normals, laplacian)``, flat ray-major.
    One forward kernel (IEEE math) and one backward kernel (``rm_render_bwd``) that returns the
    gradients of the scene parameters and of the camera quaternion / translation."""
    dev = _lib.require_cuda(orientations, translations)
    poses = pose_records(orientations, translations)
    return _render_gbuffer(tape.code_tensor(dev), tape.lowered_params(dev), poses, cam.n_views, cam.width,
                           cam.height, cam.focal_length, cam.sensor_width, cam.sensor_height, int(steps),
                           float(normals_eps), tape.n_nodes, tape.max_depth)


# ------------------------------------------------------------------------------ fused loss + backward
@torch.library.custom_op("rm_b200::render_loss", mutates_args=())
def _render_loss(code: Tensor, par: Tensor, poses: Tensor, n_views: int, width: int, height: int,
                 focal: float, sensor_w: float, sensor_h: float, steps: int, eps: float,
                 target_lam: Optional[Tensor], target_depth: Optional[Tensor], w_lam: float, w_depth: float,
                 ray_begin: int, n_rays: int, n_nodes: int, max_depth: int, tol: float,
                 want_grad: bool, strip_rays: int = 0, strip_stride: int = 0) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """One kernel: ``(loss, terms[2], d loss/d par, d loss/d poses)`` of the config-4 loss."""
    dev = _lib.require_cuda(par, poses, target_lam, target_depth)
    cont = lambda x: None if x is None else x.reshape(-1).contiguous()
    tl, td = cont(target_lam), cont(target_depth)
    for x in (tl, td):
        if x is not None and x.numel() != n_rays:
            raise _lib.RmError(f"render_loss: a target plane has {x.numel()} elements for {n_rays} rays")
    spec = _spec.selected_for(code)
    t = _tape_struct(code, par, n_nodes, max_depth)
    n_scratch = max(int((spec or lib()).rm_loss_scratch_floats(C.byref(t), n_rays, steps)), 1)
    traj = torch.empty(n_scratch, device=dev, dtype=torch.float32)
    terms = torch.empty(2, device=dev, dtype=torch.float32)
    poses = poses.contiguous()
    dpar = torch.empty_like(par) if want_grad else torch.zeros_like(par)
    dposes = torch.empty_like(poses) if want_grad else torch.zeros_like(poses)
    a = RmRenderArgs(int(ray_begin), int(n_rays), int(steps), 0, float(eps), 0.0, 0.0, 0, -1, 1, 0, 0, 0, float(tol),
                     int(strip_rays), int(strip_stride))
    la = RmLossArgs(ptr(tl), ptr(td), float(w_lam), float(w_depth), ptr(terms), n_scratch)
    cam = camera_struct(n_views, width, height, focal, sensor_w, sensor_h)
    args = (C.byref(t), C.byref(cam), ptr(poses), C.byref(a), C.byref(la), ptr(traj),
            ptr(dpar) if want_grad else None, ptr(dposes) if want_grad else None, stream_ptr(dev))
    if spec is not None:
        _spec.spec_check(spec, spec.rm_render_loss_bwd(*args), "rm_render_loss_bwd")
    else:
        check(lib().rm_render_loss_bwd(*args), "rm_render_loss_bwd")
    return terms.sum(), terms, dpar, dposes


@_render_loss.register_fake
def _(code, par, poses, n_views, width, height, focal, sensor_w, sensor_h, steps, eps, target_lam, target_depth,
      w_lam, w_depth, ray_begin, n_rays, n_nodes, max_depth, tol, want_grad, strip_rays=0, strip_stride=0):
    return par.new_empty(()), par.new_empty(2), torch.empty_like(par), torch.empty_like(poses)


def _render_loss_setup(ctx, inputs, output):
    loss, terms, dpar, dposes = output
    ctx.mark_non_differentiable(terms, dpar, dposes)
    if not inputs[20]:
        raise _lib.RmError("render_loss was evaluated with want_grad=False but autograd is recording")
    ctx.save_for_backward(dpar, dposes)


def _render_loss_backward(ctx, g_loss, g_terms, g_dpar, g_dposes):
    dpar, dposes = ctx.saved_tensors
    return (None, g_loss * dpar, g_loss * dposes) + (None,) * 20


_render_loss.register_autograd(_render_loss_backward, setup_context=_render_loss_setup)


def render_loss(tape, cam: RmCamera, orientations: Tensor, translations: Tensor, *, target_lambertian=None,
                target_depth=None, weight_lambertian: Optional[float] = None,
                weight_depth: Optional[float] = None, steps: int = 32, normals_eps: float = 5e-2,
                ray_begin: int = 0, n_rays: Optional[int] = None, strip_rays: int = 0, strip_stride: int = 0):
    """``(loss, terms)``: the inverse-rendering loss of SURVEY.md 8(d) -- ``w_l sum (lambertian - target)^2 +
    w_d sum (depth - target)^2`` over rays ``[ray_begin, ray_begin + n_rays)`` -- through ONE kernel that
    also produces its gradients (``rm_render_loss_bwd``); ``loss.backward()`` only scales them.  Weights
    default to ``1 / n_rays_total`` (the ``.mean()`` of the reference formulation)."""
    dev = _lib.require_cuda(orientations, translations)
    total = cam.n_views * cam.width * cam.height
    n = total - ray_begin if n_rays is None else int(n_rays)
    wl = 1.0 / total if weight_lambertian is None else float(weight_lambertian)
    wd = 1.0 / total if weight_depth is None else float(weight_depth)
    poses = pose_records(orientations, translations)
    par = tape.lowered_params(dev)
    want = torch.is_grad_enabled() and (par.requires_grad or poses.requires_grad)
    loss, terms, _, _ = _render_loss(tape.code_tensor(dev), par, poses, cam.n_views, cam.width, cam.height,
                                     cam.focal_length, cam.sensor_width, cam.sensor_height, int(steps),
                                     float(normals_eps), target_lambertian, target_depth, wl, wd, int(ray_begin),
                                     n, tape.n_nodes, tape.max_depth, BWD_CONVERGE_TOL, bool(want),
                                     int(strip_rays), int(strip_stride))
    return loss, terms


BACKWARD_READY = True
