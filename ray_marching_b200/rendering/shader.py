"""Drop-in ``Shader`` (reference ``rendering/shader.py:174-263``).

The fused render kernel evaluates these shaders in its epilogue (``csrc/rm_shade.cuh``,
``csrc/rm_render.cuh``); this module keeps the reference's stand-alone call signature for
callers that hold a G-buffer, and is what autograd differentiates in the inverse-rendering
configuration.  It only accepts CUDA tensors.
"""
from __future__ import annotations

import math
import os

import numpy as np
import torch
import torch.nn as nn
from torch import Tensor

from .. import _lib
from .. import quaternion as Q

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "data", "cyclic_cmap_f32.npy")
MODES = ['lambertian', 'distance', 'proximity', 'vignette', 'normal', 'laplacian', 'tangent', 'spin']
GAMMA = 1 / 2.33


def load_cmap() -> Tensor:
    """The (4096,3) cyclic colour LUT of the tangent / spin shaders (input data of the
    reference, ``data/cyclic_cmap.pt`` loaded at ``shader.py:177``), as float32."""
    return torch.from_numpy(np.load(_DATA))


def _log_normalise(x: Tensor) -> Tensor:
    logs = x.clamp(1e-2, float('inf')).log()
    lo, hi = logs.min(), logs.max()
    return ((logs - lo) / (hi - lo)).pow(GAMMA)


def _colour_wheel(re: Tensor, im: Tensor, cmap: Tensor, degree: int) -> Tensor:
    size = cmap.shape[0]
    idx = ((torch.atan2(im, re) / math.tau + 0.5) * degree * size).floor().long().remainder(size)
    return torch.hypot(re, im)[..., None] * cmap[idx]


class Shader(nn.Module):
    def __init__(self):
        super().__init__()
        self.register_buffer("cyclic_cmap", load_cmap())

    def forward(self, px_coords: Tensor, camera_orientation: Tensor, pixel_frames: Tensor,
                ray_directions: Tensor, surface_coords: Tensor, surface_normals: Tensor,
                surface_laplacian: Tensor, surface_distances: Tensor, mode: int,
                degree: int) -> Tensor:
        _lib.require_cuda(ray_directions)
        name = MODES[mode % len(MODES)]
        v, n = ray_directions, surface_normals
        if name == 'lambertian':                                     # shader.py:16-20
            return (-(v * n).sum(-1, keepdim=True)).clamp(0, 1)
        if name == 'distance':                                       # shader.py:27-38
            return _log_normalise((px_coords - surface_coords).norm(dim=-1, keepdim=True))
        if name == 'proximity':                                      # shader.py:45-55
            return _log_normalise(surface_distances)
        if name == 'vignette':                                       # shader.py:62-66
            fwd = pixel_frames[:, :, 2]                               # third column, per view
            return (v * fwd[:, None, None, :]).sum(-1, keepdim=True).pow(3)
        if name == 'normal':                                         # shader.py:73-74
            return n.abs().clamp(0, 1)
        if name == 'laplacian':                                      # shader.py:81-89
            L = surface_laplacian
            return ((1 - L / L.abs().max()) / 2).clamp(0, 1).pow(GAMMA)
        conj = Q.conjugate(camera_orientation)[:, None, None, :]
        if name == 'tangent':                                        # shader.py:125-150
            across = n - (n * v).sum(-1, keepdim=True) * v
            s = Q.rotation(across, conj)
            return _colour_wheel(s[..., 0], s[..., 1], self.cyclic_cmap, degree)
        # spin                                                        # shader.py:157-171
        val = Q.multiply(torch.nn.functional.pad(n, [1, 0]), conj.expand(*n.shape[:-1], 4))
        a, bcd = val[..., 0], val[..., 1:]
        re = a * a - (bcd * bcd).sum(-1)
        im = bcd.norm(dim=-1) * a * 2
        return _colour_wheel(im, re, self.cyclic_cmap, degree)
