"""Seeded scene/point cases shared by the golden generator and the tests.  TEST INFRASTRUCTURE ONLY.

Every builder takes a ``classes`` namespace so the same case can be built from the
reference's classes (``oracle/make_golden.py``, build container only) or from the drop-in
classes (tests)."""
from __future__ import annotations

import torch


def node_cases(c):
    """name -> scene node; covers every primitive and combinator."""
    r2 = 0.5 ** 0.5
    box = lambda: c.SDFBox(halfsides=(0.1, 0.2, 0.05))
    cases = {
        "sphere": c.SDFSphere(0.5),
        "box": box(),
        "plane": c.SDFPlane(),
        "line": c.SDFLine((-1.0, 1.0, 2.0), (1.0, 1.0, 0.0), 0.1),
        "disk": c.SDFDisk(0.5),
        "torus": c.SDFTorus(0.5, 0.1),
        "rounding_box": c.SDFRounding(box(), 0.05),
        "onion_box": c.SDFOnion(box(), 0.1),
        "affine_torus": c.SDFAffineTransformation(c.SDFTorus(0.5, 0.1), (0.0, r2, r2, 0.0),
                                                  (0.0, 0.5, 1.0)),
        "affine_plane_nonunit": c.SDFAffineTransformation(c.SDFPlane(), (0.9014, 0.25, 0.25, 0.25),
                                                          (0.3, -0.2, 0.1)),
        "affine_nested": c.SDFAffineTransformation(
            c.SDFUnion([
                c.SDFAffineTransformation(c.SDFDisk(0.4), (0.8, 0.1, -0.5, 0.3), (0.5, 0.0, 0.0)),
                c.SDFBox((0.3, 0.1, 0.2)),
                c.SDFRounding(c.SDFOnion(c.SDFSphere(0.6), 0.05), 0.01),
            ]),
            (0.7, -0.2, 0.6, 0.1), (-0.2, 0.4, 0.3)),
        "smooth_nested": c.SDFSmoothUnion([
            c.SDFSphere(0.4),
            c.SDFOnion(c.SDFSmoothUnion([
                c.SDFAffineTransformation(c.SDFBox((0.2, 0.2, 0.2)), (1.0, 0.0, 0.0, 0.0), (0.6, 0.0, 0.0)),
                c.SDFTorus(0.7, 0.1)], 9.0), 0.03),
            c.SDFUnion([c.SDFLine((0.0, -1.0, 0.0), (0.0, 1.0, 0.5), 0.05), c.SDFPlane()]),
        ], 6.0),
    }
    return cases


def eval_points(seed: int = 0, n: int = 192) -> torch.Tensor:
    """Random points in [-2.5,2.5]^3 plus adversarial ones (origin, axes, on-surface, far)."""
    g = torch.Generator().manual_seed(seed)
    rnd = (torch.rand(n, 3, generator=g) - 0.5) * 5.0
    special = torch.tensor([
        [0.3, -0.2, 0.7], [0.0, 0.0, 0.0], [-1.5, 2.0, 0.25],
        [0.5, 0.0, 0.0], [0.0, 0.5, 0.0], [0.0, 0.0, 0.5], [-0.5, 0.0, 0.0],
        [0.1, 0.2, 0.05], [0.1, 0.1, 0.1], [0.05, 0.05, 0.05], [0.0, 0.0, 1.0],
        [1.0, 1.0, 0.0], [-1.0, 1.0, 2.0], [0.0, 1.0, 1.0],
        [1e3, -2e3, 5e2], [1e9, 1e9, -1e9],
    ])
    return torch.cat([special, rnd], dim=0)


def config1_camera(width: int, height: int, px: float = 3.45e-6):
    """Intrinsics of main.py:25,39-41 for an image of the given size."""
    return dict(num_cameras=1, px_width=width, px_height=height, focal_length=px * height,
                sensor_width=px * width, sensor_height=px * height)
