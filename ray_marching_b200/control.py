"""Drop-in ``RenderLoop`` and ``make_reflection_directions`` (reference ``control.py:179-258``).

``EventAggregator`` (keyboard / mouse input through pynput, ``control.py:18-176``) is host
input glue that needs an X session and is out of scope; its ``get_state()`` tuple is what
``RenderLoop.forward`` consumes, unchanged.
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch
import torch.nn as nn
from torch import Tensor

from . import _lib, ops
from .rendering.ray_marching import PinholeCamera, SDFMarcher, SDFNormals, _grad_mode
from .rendering.shader import Shader
from .tape import tape_of


def make_reflection_directions(surface_normals: Tensor, ray_dirs: Tensor) -> Tensor:
    """``v - 2 (v.n) n`` (reference ``control.py:179-194``)."""
    return ray_dirs - 2 * (ray_dirs * surface_normals).sum(dim=-1, keepdim=True) * surface_normals


class RenderLoop(nn.Module):
    """camera -> march -> f(p) -> stencil -> shader, ``(N,H,W,3)`` (reference ``control.py:197-258``).

    Without autograd one fused kernel produces the requested image.  With autograd recording
    (scene parameters or camera pose require grad) the same pipeline is composed from the
    differentiable custom ops.  Extra, non-reference arguments are keyword-only."""

    def __init__(self, scene, num_cameras: int = 1, px_width: int = 800, px_height: int = 800,
                 focal_length: float = 17e-3, sensor_width: float = 17e-3,
                 sensor_height: float = 17e-3, normals_eps: float = 5e-2):
        super().__init__()
        self.scene = scene
        self.px_width = px_width
        self.px_height = px_height
        self.camera = PinholeCamera(num_cameras=num_cameras, px_width=px_width,
                                    px_height=px_height, focal_length=focal_length,
                                    sensor_width=sensor_width, sensor_height=sensor_height)
        self.marcher = SDFMarcher(sdf_scene=self.scene)
        self.normals = SDFNormals(sdf_scene=self.scene, normals_eps=normals_eps)
        self.shader = Shader()
        self.early_exit = False          # bit-identical fixed-point exit of the march loop
        self.fast_math = True            # approximate sqrt/exp inside march steps only
        # bounding-sphere culling: False; "exact" = whole smooth unions skipped where they provably lose the
        # enclosing min (bit-identical pixels); True = additionally children below 2^-24 of the largest term
        self.cull = False

    def specialise(self, enable: bool = True) -> "RenderLoop":
        """Compiles (nvcc, once per scene structure, cached under build/spec/) and selects a render
        kernel with this scene's tape unrolled into straight-line code instead of interpreted
        (``specialise.py``); parameters stay run-time data.  Small tapes only; ``enable=False``
        returns to the interpreter.  Affects the fused no-autograd path and the autograd path
        (G-buffer forward and the fused backward kernel)."""
        from . import specialise as S
        tape = tape_of(self.scene)
        handle = S.specialised_lib(tape) if enable else None
        tape.spec_lib = handle if (handle is not None and handle.has_forward) else None
        S.select(tape, handle)                   # the autograd path (render_gbuffer / render_bwd) follows
        return self

    # ---- fused, no autograd -------------------------------------------------------------
    def render_planes(self, orientations: Tensor, translations: Tensor, *, modes: Sequence[int] = (),
                      gbuffer: bool = False, display_mode: int = -1, degree: int = 1,
                      marching_steps: int = 32, bounces: int = 0, hit_tau: float = 1e-3,
                      lift_delta: Optional[float] = None, ray_begin: int = 0,
                      n_rays: Optional[int] = None, strip_rays: int = 0, strip_stride: int = 0,
                      finalize: bool = True, out: Optional[ops.RenderResult] = None) -> ops.RenderResult:
        """One launch for several shader modes and/or the G-buffer (planes are flat, ray-major)."""
        _lib.require_cuda(orientations, translations)
        mask = 0
        for m in modes:
            mask |= 1 << (m % 8)
        poses = ops.pose_records(orientations, translations)
        return ops.render(tape_of(self.scene), self.camera.struct(), poses, steps=marching_steps,
                          normals_eps=self.normals.normals_eps, mode_mask=mask, gbuffer=gbuffer,
                          display_mode=display_mode, degree=degree, bounces=bounces,
                          hit_tau=hit_tau, lift_delta=lift_delta, early_exit=self.early_exit,
                          fast_math=ops.fast_math_level(self.fast_math, self.cull),
                          cmap=self.shader.cyclic_cmap,
                          ray_begin=ray_begin, n_rays=n_rays, strip_rays=strip_rays,
                          strip_stride=strip_stride, finalize=finalize, out=out)

    def display_frame(self, orientations: Tensor, translations: Tensor, mode: int = 0,
                      degree: int = 1, marching_steps: int = 32) -> Tensor:
        """Headless replacement of ``Window.draw`` input (``main.py:78-84``): contiguous
        ``(H,W,4)`` float32 RGBA, alpha 1 (single camera)."""
        if self.camera.num_cameras != 1:
            raise ValueError("display_frame renders one camera; average a batch on the caller side")
        res = self.render_planes(orientations, translations, display_mode=mode % 8, degree=degree,
                                 marching_steps=marching_steps)
        return res.rgba.reshape(self.px_height, self.px_width, 4)

    def supersampled(self, orientations: Tensor, translations: Tensor, mode: int = 0, factor: int = 2,
                     degree: int = 1, marching_steps: int = 32) -> Tensor:
        """``factor`` x ``factor`` rays per pixel, box-filtered (the reference's TODO "supersampling",
        README.md:246).  The sub-pixel rays are the pixel centres of a sensor of the
        same size with ``factor`` times the resolution -- the closed form of ``PinholeCamera`` at
        ``(factor W, factor H)`` -- marched and shaded by the same fused kernel (normalising shaders take
        their range over all sub-pixel rays), then averaged per pixel.  ``(N,H,W,3)``, no autograd."""
        if factor < 1:
            raise ValueError("factor must be >= 1")
        N, H, W = self.camera.num_cameras, self.px_height, self.px_width
        cam = ops.camera_struct(N, W * factor, H * factor, self.camera.focal_length,
                                self.camera.sensor_width, self.camera.sensor_height)
        m = mode % 8
        poses = ops.pose_records(orientations, translations)
        res = ops.render(tape_of(self.scene), cam, poses, steps=marching_steps,
                         normals_eps=self.normals.normals_eps, mode_mask=1 << m, degree=degree,
                         early_exit=self.early_exit,
                         fast_math=ops.fast_math_level(self.fast_math, self.cull),
                         cmap=self.shader.cyclic_cmap)
        fine = res.images[m].reshape(N, H, factor, W, factor, -1)
        return fine.mean(dim=(2, 4)).expand(-1, H, W, 3)

    # ---- reference signature --------------------------------------------------------------
    def forward(self, orientations: Tensor, translations: Tensor, mode: int = 0, degree: int = 1,
                marching_steps: int = 32):
        N, H, W = self.camera.num_cameras, self.px_height, self.px_width
        tape = tape_of(self.scene)
        if not (_grad_mode(orientations, translations) or _grad_mode(*tape.params)):
            m = mode % 8
            res = self.render_planes(orientations, translations, modes=(m,), degree=degree,
                                     marching_steps=marching_steps)
            return res.images[m].reshape(N, H, W, -1).expand(-1, H, W, 3)
        # autograd recording: fused differentiable G-buffer (one forward + one backward kernel),
        # then the reference's per-pixel shader algebra on the planes
        gb = self.gbuffer(orientations, translations, marching_steps)
        pixel_frames = self.camera.quaternion_to_so3(orientations[:, None, None, :])
        image = self.shader(gb["origin"], orientations, pixel_frames, gb["dirs"], gb["pos"], gb["normals"],
                            gb["laplacian"], gb["sdf"], mode=mode, degree=degree)
        return image.expand(-1, H, W, 3)

    def inverse_loss(self, orientations: Tensor, translations: Tensor, target_lambertian: Optional[Tensor],
                     target_depth: Optional[Tensor], *, marching_steps: int = 32,
                     weight_lambertian: Optional[float] = None, weight_depth: Optional[float] = None,
                     ray_begin: int = 0, n_rays: Optional[int] = None, strip_rays: int = 0,
                     strip_stride: int = 0, return_terms: bool = False):
        """The inverse-rendering loss of BASELINE config 4,
        ``mean((lambertian - target_lambertian)^2) + mean((depth - target_depth)^2)`` with
        ``depth = |marched position - ray origin|``, evaluated AND differentiated by one fused kernel
        (``rm_render_loss_bwd``): the forward replay, the shading, the residuals, the adjoint seeds, the
        reverse march and the camera-pose contraction never leave registers.  Equivalent to composing
        :meth:`gbuffer` with the reference's shader algebra (``shader.py:16-20``) and ``.mean()`` -- and
        tested against it -- at a third of the GPU time.  ``ray_begin`` / ``n_rays`` select a shard of the
        batch (targets are indexed by the local ray); weights default to ``1 / total rays``."""
        loss, terms = ops.render_loss(tape_of(self.scene), self.camera.struct(), orientations, translations,
                                      target_lambertian=target_lambertian, target_depth=target_depth,
                                      weight_lambertian=weight_lambertian, weight_depth=weight_depth,
                                      steps=marching_steps, normals_eps=self.normals.normals_eps,
                                      ray_begin=ray_begin, n_rays=n_rays, strip_rays=strip_rays,
                                      strip_stride=strip_stride)
        return (loss, terms) if return_terms else loss

    def gbuffer(self, orientations: Tensor, translations: Tensor, marching_steps: int = 32):
        """Differentiable G-buffer of control.py:239-245 as ``(N,H,W,c)`` tensors: ``origin`` (= pixel
        positions), ``dirs``, ``pos``, ``sdf``, ``normals``, ``laplacian``."""
        N, H, W = self.camera.num_cameras, self.px_height, self.px_width
        o, v, p, f, n, lap = ops.render_gbuffer(tape_of(self.scene), self.camera.struct(), orientations,
                                                translations, marching_steps, self.normals.normals_eps)
        r3 = lambda x: x.reshape(N, H, W, 3)
        r1 = lambda x: x.reshape(N, H, W, 1)
        return {"origin": r3(o), "dirs": r3(v), "pos": r3(p), "sdf": r1(f), "normals": r3(n),
                "laplacian": r1(lap)}
