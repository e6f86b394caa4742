"""Seeded synthetic multi-primitive scenes for BASELINE configs 3-5.

Child ``i`` has type ``i % 6`` in [sphere, box, torus, capsule, disk, rounded box],
wrapped in an affine transform with a random unit quaternion and a translation in
``[-2, 2]^3``; the root is a smooth union (k = 22).  ``shell=True`` wraps the result
as ``SDFUnion([SDFOnion(SDFBox((5,5,5)), 0.1), root])`` (the ``make_test_scene2``
pattern, ``scene_registry.py:55-60``) so that every ray hits something.

Only reference-API classes are used, so the same generator builds the scene from the
reference's own classes when ``classes`` is given (golden-vector generation).
"""
from __future__ import annotations

import torch

from .scene_registry import _default_classes


def _u(g, lo, hi, n=None):
    shape = () if n is None else (n,)
    return (torch.rand(shape, generator=g) * (hi - lo) + lo)


def make_synthetic_scene(num_primitives: int, seed: int = 0, shell: bool = True,
                         blend_k: float = 22.0, classes=None):
    c = classes or _default_classes()
    g = torch.Generator().manual_seed(seed)
    children = []
    for i in range(num_primitives):
        kind = i % 6
        if kind == 0:
            prim = c.SDFSphere(radius=float(_u(g, 0.1, 0.3)))
        elif kind == 1:
            prim = c.SDFBox(halfsides=tuple(_u(g, 0.05, 0.25, 3).tolist()))
        elif kind == 2:
            prim = c.SDFTorus(radius1=float(_u(g, 0.2, 0.4)), radius2=float(_u(g, 0.03, 0.08)))
        elif kind == 3:
            prim = c.SDFLine(start=tuple(_u(g, -0.5, 0.5, 3).tolist()),
                             end=tuple(_u(g, -0.5, 0.5, 3).tolist()),
                             radius=float(_u(g, 0.02, 0.07)))
        elif kind == 4:
            prim = c.SDFDisk(radius=float(_u(g, 0.1, 0.3)))
        else:
            prim = c.SDFRounding(c.SDFBox(halfsides=tuple(_u(g, 0.05, 0.25, 3).tolist())),
                                 rounding=0.02)
        q = torch.randn(4, generator=g)
        q = q / q.norm()
        t = _u(g, -2.0, 2.0, 3)
        children.append(c.SDFAffineTransformation(
            sdf=prim, orientation=q.tolist(), translation=t.tolist()))
    root = c.SDFSmoothUnion(sdfs=children, blend_k=blend_k)
    if not shell:
        return root
    return c.SDFUnion([c.SDFOnion(c.SDFBox(halfsides=(5.0, 5.0, 5.0)), radius=0.1), root])
