"""The reference's two registry scenes, built from the drop-in node classes.

Same factory names and the same numbers as ``scene/scene_registry.py:18-51`` and
``:54-79`` — the scenes are *inputs* of the hot path, so they are restated as data.
``classes`` lets the golden-vector generator build the identical tree out of the
reference's own classes.
"""
from __future__ import annotations

import types

from . import primitives as _p
from . import transformations as _t


def _default_classes():
    ns = types.SimpleNamespace()
    for mod in (_p, _t):
        for name in dir(mod):
            if name.startswith("SDF") and name != "SDFNode":
                setattr(ns, name, getattr(mod, name))
    return ns


def make_test_scene(classes=None):
    """Smooth union (k=22) of an onion box, a sphere, a capsule and a torus."""
    c = classes or _default_classes()
    r2 = 0.5 ** 0.5
    shell_box = c.SDFAffineTransformation(
        sdf=c.SDFOnion(c.SDFBox(halfsides=(0.1, 0.2, 0.05)), radius=0.1),
        orientation=[0.9014, 0.25, 0.25, 0.25],
        translation=[0.0, 0.25, 0.25],
    )
    ball = c.SDFAffineTransformation(
        sdf=c.SDFSphere(radius=0.5),
        orientation=[1.0, 0.0, 0.0, 0.0],
        translation=[0.0, 0.0, 1.0],
    )
    capsule = c.SDFLine(start=(-1.0, 1.0, 2.0), end=(1.0, 1.0, 0.0), radius=0.1)
    ring = c.SDFAffineTransformation(
        sdf=c.SDFTorus(radius1=0.5, radius2=0.1),
        orientation=[0.0, r2, r2, 0.0],
        translation=[0.0, 0.5, 1.0],
    )
    return c.SDFSmoothUnion(sdfs=[shell_box, ball, capsule, ring], blend_k=22.0)


def make_test_scene2(classes=None):
    """Hard union of an enclosing onion-box shell with {sphere, torus, capsule}."""
    c = classes or _default_classes()
    shell = c.SDFOnion(c.SDFBox(halfsides=(5.0, 5.0, 5.0)), radius=0.1)
    inner = c.SDFUnion(sdfs=[
        c.SDFSphere(radius=0.5),
        c.SDFTorus(radius1=1.0, radius2=0.25),
        c.SDFLine(start=(1.0, 0.0, 0.0), end=(-1.0, 0.0, 0.0), radius=0.1),
    ])
    return c.SDFUnion([shell, inner])
