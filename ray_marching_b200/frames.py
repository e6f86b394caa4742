"""Headless frame loop, scripted camera paths and image writers.

Replaces the interactive tail of the reference's ``main.py:52-88`` -- ``EventAggregator.get_state``
-> ``RenderLoop`` -> ``Window.draw`` -- which needs an X session, pynput and an OpenGL window
(unavailable on this box).  What remains is its data path:

* :class:`CameraPath`   the camera-state integration of ``get_state`` (``control.py:148-165``) as a
                        device kernel over a batch of views, optionally over a whole scripted
                        sequence of inputs that lives on the GPU;
* :class:`FrameLoop`    camera update + fused render + normalisation captured ONCE as a CUDA graph;
                        a frame is one graph replay that leaves the ``(H,W,4)`` float32 RGBA tensor
                        ``Window.draw`` consumes (``torchwindow/window.py:146-174``) in a static
                        device buffer -- no host work, allocation or launch overhead per frame
                        (config 1 is launch-latency scale: SURVEY.md §8(d));
* :class:`GraphedStep` an inverse-rendering step (forward, loss, backward, Adam) captured as one CUDA
                        graph -- the training-side counterpart of :class:`FrameLoop`;
* :func:`write_png` / :func:`write_npy`  the ``save_frame`` branch the reference left commented out
                        (``main.py:72-77``).
"""
from __future__ import annotations

import ctypes as C
import struct
import zlib
from typing import Optional

import numpy as np
import torch
from torch import Tensor

from . import _lib, ops
from ._lib import check, lib, ptr, stream_ptr
from .tape import tape_of


class CameraPath:
    """Camera state ``(orientation (N,4), position (N,3))`` integrated on the device.

    ``step(translation_input, orientation_input)`` applies one ``get_state`` update
    (``control.py:148-165``: ``position += Q.rotation(translation_input * translation_sensitivity,
    orientation)``; ``orientation = normalize(Q.multiply(orientation, Q.to_versor(orientation_input
    * rotation_sensitivity)))``).  With ``script`` -- a ``(frames, N, 6)`` tensor of
    ``(translation_input, orientation_input)`` rows -- every :meth:`advance` applies the next row of
    the script, indexed by a device-side counter, so it can be replayed from inside a CUDA graph.
    The sensitivities default to ``EventAggregator``'s (``control.py:30-31``: 0.1 and 0.25)."""

    def __init__(self, initial_position, initial_orientation, *, translation_sensitivity: float = 0.1,
                 rotation_sensitivity: float = 0.25, script: Optional[Tensor] = None, device="cuda"):
        device = torch.device(device)
        if device.type != "cuda":
            raise _lib.RmError("CameraPath integrates on the GPU; there is no CPU path")
        pos = torch.as_tensor(initial_position, dtype=torch.float32).reshape(-1, 3)
        ori = torch.as_tensor(initial_orientation, dtype=torch.float32).reshape(-1, 4)
        if pos.shape[0] != ori.shape[0]:
            raise ValueError("one position and one orientation per camera")
        self.num_cameras = pos.shape[0]
        self.translation_sensitivity = float(translation_sensitivity)
        self.rotation_sensitivity = float(rotation_sensitivity)
        self.poses = torch.cat([ori, pos, torch.zeros(self.num_cameras, 1)], dim=-1).to(device).contiguous()
        self.counter = torch.zeros(1, dtype=torch.int32, device=device)
        self._one = torch.zeros(1, self.num_cameras, 6, device=device)
        self.script = None
        if script is not None:
            script = torch.as_tensor(script, dtype=torch.float32)
            if script.dim() != 3 or script.shape[1:] != (self.num_cameras, 6):
                raise ValueError(f"script must be (frames, {self.num_cameras}, 6), got {tuple(script.shape)}")
            self.script = script.to(device).contiguous()

    @property
    def orientation(self) -> Tensor:
        return self.poses[:, 0:4]

    @property
    def position(self) -> Tensor:
        return self.poses[:, 4:7]

    def step(self, translation_input, orientation_input) -> None:
        """One update from explicit inputs, each ``(3,)`` or ``(N,3)`` (host or device)."""
        dev = self.poses.device
        ti = torch.as_tensor(translation_input, dtype=torch.float32).to(dev).expand(self.num_cameras, 3)
        oi = torch.as_tensor(orientation_input, dtype=torch.float32).to(dev).expand(self.num_cameras, 3)
        self._one[0].copy_(torch.cat([ti, oi], dim=-1))
        check(lib().rm_camera_update(ptr(self.poses), ptr(self._one), self.num_cameras, 1,
                                     self.translation_sensitivity, self.rotation_sensitivity, None,
                                     stream_ptr(dev)), "rm_camera_update")

    def advance(self) -> None:
        """Applies the next row of the script (wraps around); capturable in a CUDA graph."""
        if self.script is None:
            raise ValueError("advance() needs a script; use step() for explicit inputs")
        check(lib().rm_camera_update(ptr(self.poses), ptr(self.script), self.num_cameras,
                                     self.script.shape[0], self.translation_sensitivity,
                                     self.rotation_sensitivity, ptr(self.counter),
                                     stream_ptr(self.poses.device)), "rm_camera_update")


class FrameLoop:
    """``while running: state = get_state(); image = render_loop(...); window.draw(image)`` without
    the window: every :meth:`frame` leaves the display tensor in :attr:`rgba`.

    ``graph=True`` captures {camera update (when the path is scripted), reduction init, fused render,
    normalisation} once; the capture is redone when the lowered scene parameters change (a static
    scene under a moving camera never recaptures)."""

    def __init__(self, render_loop, path: CameraPath, *, mode: int = 0, degree: int = 1,
                 marching_steps: int = 32, graph: bool = True):
        if render_loop.camera.num_cameras != 1 or path.num_cameras != 1:
            raise ValueError("the display loop shows one camera (main.py:78-84 averages a batch; "
                             "render batches with RenderLoop.render_planes instead)")
        self.loop, self.path = render_loop, path
        self.mode, self.degree, self.marching_steps = int(mode) % 8, int(degree), int(marching_steps)
        self.use_graph = bool(graph)
        self._res = ops.RenderResult()
        self._graph = None
        self._key = None
        self.frames = 0

    @property
    def rgba(self) -> Tensor:
        """``(H,W,4)`` float32, alpha 1: the tensor ``Window.draw`` takes."""
        return self._res.rgba.reshape(self.loop.px_height, self.loop.px_width, 4)

    def _launch(self, scripted: bool) -> None:
        if scripted:
            self.path.advance()
        loop = self.loop
        ops.render(tape_of(loop.scene), loop.camera.struct(), self.path.poses, steps=self.marching_steps,
                   normals_eps=loop.normals.normals_eps, display_mode=self.mode, degree=self.degree,
                   early_exit=loop.early_exit,
                   fast_math=ops.fast_math_level(loop.fast_math, loop.cull),
                   cmap=loop.shader.cyclic_cmap, out=self._res)

    def frame(self, translation_input=None, orientation_input=None) -> Tensor:
        """Advances the camera (next script row, or the given explicit inputs) and renders."""
        scripted = translation_input is None and orientation_input is None
        if not scripted:
            zero = (0.0, 0.0, 0.0)
            self.path.step(zero if translation_input is None else translation_input,
                           zero if orientation_input is None else orientation_input)
        elif self.path.script is None:
            scripted = False                      # static camera
        if not self.use_graph:
            self._launch(scripted)
        else:
            with torch.no_grad():                 # the cached lowering (no autograd graph), as ops.render uses
                par = tape_of(self.loop.scene).lowered_params(self.path.poses.device)
            # everything _launch bakes into the captured kernel arguments
            loop, cam = self.loop, self.loop.camera
            key = (par.data_ptr(), par._version, scripted, self.mode, self.degree, self.marching_steps,
                   bool(loop.early_exit), bool(loop.fast_math), str(loop.cull),
                   id(getattr(tape_of(loop.scene), "spec_lib", None)), float(loop.normals.normals_eps),
                   cam.num_cameras, cam.px_width, cam.px_height, float(cam.focal_length),
                   float(cam.sensor_width), float(cam.sensor_height))
            if self._graph is None or key != self._key:
                self._capture(scripted)
                self._key, self._par = key, par        # holding `par` keeps its address from being reused
            self._graph.replay()
        self.frames += 1
        return self.rgba

    def _capture(self, scripted: bool) -> None:
        dev = self.path.poses.device
        saved = (self.path.poses.clone(), self.path.counter.clone())
        torch.cuda.synchronize(dev)
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):            # warm-up outside capture: allocations, smem attributes
            self._launch(scripted)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, capture_error_mode="relaxed"):
            self._launch(scripted)
        # the warm-up advanced the path once: restore it so that capture has no side effect
        self.path.poses.copy_(saved[0])
        self.path.counter.copy_(saved[1])
        self._graph = g


class GraphedStep:
    """An optimisation step (forward kernels, loss, ``backward()``, optimiser update) captured ONCE
    as a CUDA graph; :meth:`__call__` replays it and returns the static loss tensor.

    The inverse-rendering loop of config 4 on a registry scene is launch-bound -- the G-buffer and
    backward kernels take ~0.55 ms of GPU time per step, the ~60 launches of the loss algebra, the
    parameter lowering and Adam cost more than that in host time -- so, like a display frame
    (:class:`FrameLoop`), the step is replayed from a graph.

    ``forward_backward()`` runs the forward, the loss and ``loss.backward()`` and returns the loss; it
    may not synchronise, read tensors back, or call ``zero_grad`` / ``optimizer.step()`` (this class
    does).  ``between`` -- e.g. the NCCL all-reduce of the gradients of a view-sharded job -- runs
    eagerly between the two halves; with it the step is two graphs (forward+backward, optimiser
    update), without it one.  The optimiser must be created with ``capturable=True``.  Gradients live
    in the graph's memory pool and are overwritten by every replay."""

    def __init__(self, forward_backward, optimizer, *, between=None, warmup: int = 3):
        self.optimizer, self.between = optimizer, between
        params = [p for group in optimizer.param_groups for p in group["params"]]
        saved = [p.detach().clone() for p in params]
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(max(warmup, 1)):          # allocator warm-up + optimiser state allocation
                optimizer.zero_grad(set_to_none=True)
                forward_backward()
                if between is not None:
                    between()
                optimizer.step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        # the warm-up steps must not count: restore the parameters and return the optimiser state
        # (allocated now, as graph capture needs) to its initial zeros
        with torch.no_grad():
            for p, p0 in zip(params, saved):
                p.copy_(p0)
            for st in optimizer.state.values():
                for v in st.values():
                    if torch.is_tensor(v):
                        v.zero_()
        optimizer.zero_grad(set_to_none=True)
        self.graph = torch.cuda.CUDAGraph()
        self.update_graph = None
        with torch.cuda.graph(self.graph):
            self.loss = forward_backward()
            if between is None:
                optimizer.step()
        if between is not None:
            between()
            self.update_graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.update_graph, pool=self.graph.pool()):
                optimizer.step()
            # the capture above ran one real update: undo it
            torch.cuda.synchronize()
            with torch.no_grad():
                for p, p0 in zip(params, saved):
                    p.copy_(p0)
                for st in optimizer.state.values():
                    for v in st.values():
                        if torch.is_tensor(v):
                            v.zero_()

    def __call__(self) -> Tensor:
        self.graph.replay()
        if self.update_graph is not None:
            self.between()
            self.update_graph.replay()
        return self.loss


def to_rgba8(rgba: Tensor) -> np.ndarray:
    """``(H,W,4)`` float RGBA in [0,1] (NaN -> 0, as an 8-bit framebuffer would show it) -> uint8 on
    the host."""
    x = torch.nan_to_num(rgba.detach().float(), nan=0.0).clamp(0, 1)
    return (x * 255.0 + 0.5).to(torch.uint8).cpu().numpy()


def write_png(path, rgba: Tensor) -> None:
    """Writes ``(H,W,4)`` (or ``(H,W,3)``) float images in [0,1] as an 8-bit PNG (zlib only)."""
    img = to_rgba8(rgba)
    h, w, c = img.shape
    if c not in (3, 4):
        raise ValueError("write_png takes (H,W,3) or (H,W,4)")
    raw = np.concatenate([np.zeros((h, 1), np.uint8), img.reshape(h, w * c)], axis=1).tobytes()

    def chunk(tag: bytes, data: bytes) -> bytes:
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n")
        f.write(chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 6 if c == 4 else 2, 0, 0, 0)))
        f.write(chunk(b"IDAT", zlib.compress(raw, 6)))
        f.write(chunk(b"IEND", b""))


def read_png(path) -> np.ndarray:
    """Decodes what :func:`write_png` wrote (8-bit RGB/RGBA, filter 0) -- for tests and tools."""
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w = 8, b"", 0
    while pos < len(data):
        n, tag = struct.unpack(">I", data[pos:pos + 4])[0], data[pos + 4:pos + 8]
        body = data[pos + 8:pos + 8 + n]
        if tag == b"IHDR":
            w, h, depth, ctype = struct.unpack(">IIBB", body[:10])
            c = 4 if ctype == 6 else 3
        elif tag == b"IDAT":
            idat += body
        pos += 12 + n
    rows = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + w * c)
    assert not rows[:, 0].any(), "only filter type 0 is supported"
    return rows[:, 1:].reshape(h, w, c).copy()


def write_npy(path, image: Tensor) -> None:
    """Lossless float32 dump of any plane / image."""
    np.save(path, image.detach().float().cpu().numpy())
