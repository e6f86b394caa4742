"""CPU oracle for the differentiable sphere-tracing hot path.  TEST INFRASTRUCTURE ONLY.

This file restates, in plain PyTorch CPU float32 arithmetic, the algorithm of
kyle-rosa/ray_marching for the path BASELINE.json names.  It exists to *check* the
CUDA kernels: only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
CPU-baseline / ``--impl reference`` legs may import it.  The product package
(``ray_marching_b200``) never imports it and has no CPU fallback.

Parity status: PINNED.  ``oracle/make_golden.py`` imports the *unmodified* reference
modules from ``/root/reference`` (possible only in the build container), runs them on
seeded inputs and commits the results under ``tests/golden/``;
``tests/test_oracle_golden.py`` asserts that this restatement reproduces them
(values, NaN masks and autograd gradients).  The reference itself ships no tests or
golden vectors, so those fixtures are the pin.

Every function cites the reference lines it follows.  The same torch operators are
used wherever autograd's sub-gradient conventions matter (``norm`` at 0, ``max(dim)``
ties, ``clamp`` bounds, ``where`` masks, ``logsumexp``) so that ``torch.autograd`` on
this oracle *is* the reference gradient.

Scene trees are walked by class name, so the oracle accepts trees built from the
drop-in classes (``ray_marching_b200.scene``) or from the reference's own classes.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F
from torch import Tensor

TAU = math.tau


# --------------------------------------------------------------------------- L0
# quaternion.py -- scalar-first (w, x, y, z) on the trailing axis


def q_cross(u: Tensor, v: Tensor) -> Tensor:
    """3-vector cross product (quaternion.py:6-21)."""
    ux, uy, uz = u.unbind(-1)
    vx, vy, vz = v.unbind(-1)
    return torch.stack(
        [uy * vz - uz * vy, uz * vx - ux * vz, ux * vy - uy * vx], dim=-1
    )


def q_multiply(p: Tensor, q: Tensor) -> Tensor:
    """Hamilton product (quaternion.py:24-46); summation order kept as in the reference."""
    pw, px, py, pz = p.unbind(-1)
    qw, qx, qy, qz = q.unbind(-1)
    return torch.stack(
        [
            pw * qw - (px * qx + py * qy + pz * qz),
            (pw * qx + px * qw + py * qz) - pz * qy,
            (pw * qy + py * qw + pz * qx) - px * qz,
            (pw * qz + px * qy + pz * qw) - py * qx,
        ],
        dim=-1,
    )


def q_conjugate(q: Tensor) -> Tensor:
    """(w, -x, -y, -z) (quaternion.py:49-52)."""
    sign = torch.tensor([1.0, -1.0, -1.0, -1.0], dtype=q.dtype, device=q.device)
    return q * sign


def q_rotation(v: Tensor, q: Tensor) -> Tensor:
    """``v + w t + qv x t`` with ``t = 2 qv x v`` (quaternion.py:55-72).

    A rotation only for unit ``q``; the reference applies it to un-normalised scene
    orientations, so the polynomial is kept as is."""
    qv = q[..., 1:]
    t = q_cross(qv, v) * 2
    return q_cross(qv, t) + q[..., :1] * t + v


def q_to_versor(v: Tensor) -> Tensor:
    """(sqrt(1-|v|^2), v) (quaternion.py:75-91)."""
    w = (1 - (v * v).sum(-1, keepdim=True)).pow(0.5)
    return torch.cat([w, v], dim=-1)


def camera_state_update(position: Tensor, orientation: Tensor, translation_input: Tensor,
                        orientation_input: Tensor, translation_sensitivity: float,
                        rotation_sensitivity: float):
    """One camera-state integration step of EventAggregator.get_state (control.py:148-165):
    position (N,3), orientation (N,4); inputs (1,3) or (N,3).  Returns the new (position, orientation);
    the position update uses the OLD orientation."""
    new_position = q_rotation((translation_input * translation_sensitivity).expand_as(position),
                              orientation) + position
    new_orientation = torch.nn.functional.normalize(
        q_multiply(orientation, q_to_versor(orientation_input * rotation_sensitivity).expand_as(orientation)),
        p=2, dim=-1, eps=0)
    return new_position, new_orientation


def q_to_so3(q: Tensor) -> Tensor:
    """Row-major 3x3 from pair products, flattened to (-1,3,3) (quaternion.py:114-124)."""
    w, x, y, z = q.unbind(-1)
    ww, wx, wy, wz = w * w, w * x, w * y, w * z
    xx, xy, xz = x * x, x * y, x * z
    yy, yz, zz = y * y, y * z, z * z
    rows = [
        ww + xx - yy - zz, 2 * (xy - wz), 2 * (wy + xz),
        2 * (xy + wz), ww - xx + yy - zz, 2 * (yz - wx),
        2 * (xz - wy), 2 * (wx + yz), ww - xx - yy + zz,
    ]
    return torch.stack(rows, dim=-1).view(-1, 3, 3)


# --------------------------------------------------------------------------- L1
# scene/primitives.py and scene/transformations.py, dispatched on the class name


def _norm(x: Tensor) -> Tensor:
    return x.norm(dim=-1, p=2, keepdim=True)


def _sphere(n, p):  # primitives.py:11-17
    return _norm(p) - n.radius


def _box(n, p):  # primitives.py:25-33
    a = p.abs() - n.halfsides
    amax = a.max(dim=-1, keepdim=True).values
    outside = _norm(a.where(a > 0.0, 0.0))
    return outside + amax.where(amax < 0.0, 0.0)


def _plane(n, p):  # primitives.py:40-41
    return p[..., [0]]


def _line(n, p):  # primitives.py:51-61
    ab = n.end - n.start
    ab_unit2 = ab / (ab * ab).sum(dim=-1, keepdim=True)
    ap = p - n.start
    h = (ap * ab_unit2).sum(dim=-1, keepdim=True).clamp(0.0, 1.0)
    return _norm(h * ab - ap) - n.radius


def _disk(n, p):  # primitives.py:69-82
    rim = _norm(p[..., [1, 2]]) - n.radius
    return torch.cat([p[..., [0]], rim.where(rim > 0.0, 0.0)], dim=-1).norm(dim=-1, keepdim=True)


def _torus(n, p):  # primitives.py:91-102
    ring = _norm(p[..., [0, 2]]) - n.radius1
    return _norm(torch.cat([ring, p[..., [1]]], dim=-1)) - n.radius2


def _affine(n, p):  # transformations.py:33-42
    lead = p.shape[:-1]
    local = q_rotation(
        p - n.translation.expand(*lead, 3),
        q_conjugate(n.orientation).expand(*lead, 4),
    )
    return sdf(n.sdf, local)


def _smooth_union(n, p):  # transformations.py:67-71
    stacked = torch.stack([sdf(c, p) for c in n.sdfs], dim=-2)
    return (stacked * (-n.blend_k)).logsumexp(dim=-2) / (-n.blend_k)


def _union(n, p):  # transformations.py:90-94
    return torch.stack([sdf(c, p) for c in n.sdfs], dim=-2).min(dim=-2).values


def _rounding(n, p):  # transformations.py:117-118
    return sdf(n.sdf, p) - n.rounding


def _onion(n, p):  # transformations.py:131-132
    return sdf(n.sdf, p).abs() - n.radius


_DISPATCH = {
    "SDFSphere": _sphere, "SDFBox": _box, "SDFPlane": _plane, "SDFLine": _line,
    "SDFDisk": _disk, "SDFTorus": _torus, "SDFAffineTransformation": _affine,
    "SDFSmoothUnion": _smooth_union, "SDFUnion": _union, "SDFRounding": _rounding,
    "SDFOnion": _onion,
}


def sdf(node, points: Tensor) -> Tensor:
    """Signed distance of ``points (...,3)`` to the tree rooted at ``node`` -> ``(...,1)``."""
    return _DISPATCH[type(node).__name__](node, points)


# --------------------------------------------------------------------------- L2
# rendering/ray_marching.py


def camera_base_rays(num_cameras, px_width, px_height, focal_length, sensor_width, sensor_height):
    """Sensor-plane ray origins and unit directions, camera frame (ray_marching.py:24-50)."""
    theta = torch.tensor(
        [[[sensor_width / 2, 0.0, 0.0], [0.0, -sensor_height / 2, 0.0]]]
    ).expand(num_cameras, 2, 3)
    grid = F.affine_grid(theta=theta.float(), size=(num_cameras, 1, px_height, px_width),
                         align_corners=False)
    origins = F.pad(grid, pad=[0, 1], value=0.0)
    focus = torch.tensor([[[[0.0, 0.0, -focal_length]]]]).expand(num_cameras, 1, 1, 3)
    dirs = F.normalize(origins - focus, p=2, dim=-1, eps=0)
    return origins, dirs


def camera_rays(base_origins, base_dirs, orientation: Tensor, translation: Tensor):
    """World-space rays (ray_marching.py:57-64): returns (pos, frames(N,3,3), pos, dirs)."""
    q = orientation[:, None, None, :]
    pos = q_rotation(base_origins, q) + translation[:, None, None, :]
    dirs = q_rotation(base_dirs, q)
    return pos, q_to_so3(q), pos, dirs


def march(scene, pos: Tensor, dirs: Tensor, steps: int = 32) -> Tensor:
    """Fixed-count sphere trace ``p <- f(p) v + p`` (ray_marching.py:72-84)."""
    for _ in range(steps):
        pos = sdf(scene, pos) * dirs + pos
    return pos


def stencil_offsets(eps: float):
    """Tetrahedral offsets, their differences and the inverse (ray_marching.py:94-113)."""
    s = 0.5 ** 0.5
    off = torch.tensor([[1.0, 0.0, -s], [-1.0, 0.0, -s], [0.0, 1.0, s], [0.0, -1.0, s]])
    off = F.normalize(off, dim=-1, p=2, eps=0.0) * eps
    rel = off[[1, 2, 3], :] - off[[0], :]
    return off, rel, rel.inverse()


def normals_laplacian(scene, coords: Tensor, eps: float):
    """Surface normal and Laplacian from 4+1 SDF samples (ray_marching.py:115-125)."""
    off, _, inv = stencil_offsets(eps)
    samples = sdf(scene, coords[..., None, :] + off)            # (...,4,1)
    delta = samples[..., [1, 2, 3], :] - samples[..., [0], :]      # (...,3,1)
    normals = (inv * delta[..., None, :, 0]).sum(dim=-1)
    normals = F.normalize(normals, dim=-1, p=2, eps=0.0)
    lap = (sdf(scene, coords) - samples.mean(dim=-2)) * (6 / eps ** 2)
    return normals, lap


# --------------------------------------------------------------------------- shaders
# rendering/shader.py


def shade_lambertian(dirs, normals):  # shader.py:16-20
    return (dirs * normals).sum(dim=-1, keepdim=True).neg().clamp(0, 1)


def _log_normalised(x):  # shared tail of shader.py:28-38 and :45-55
    logs = x.clamp(1e-2, float("inf")).log()
    return ((logs - logs.min()) / (logs.max() - logs.min())).pow(1 / 2.33)


def shade_distance(px_coords, surface_coords):  # shader.py:27-38
    return _log_normalised(_norm(px_coords - surface_coords))


def shade_proximity(surface_distances):  # shader.py:45-55
    return _log_normalised(surface_distances)


def shade_vignette(dirs, pixel_frames):  # shader.py:62-66
    return (dirs * pixel_frames[..., 2]).sum(dim=-1, keepdim=True).pow(3)


def shade_normal(normals):  # shader.py:73-74
    return normals.abs().clamp(0, 1)


def shade_laplacian(lap):  # shader.py:81-89
    return ((lap / lap.abs().max()) * -1 + 1).div(2).clamp(0, 1).pow(1 / 2.33)


def _angle_colour(re, im, cmap, degree):  # shader.py:92-104
    size = cmap.shape[0]
    idx = ((torch.atan2(im, re) / TAU + 0.5) * degree * size).floor().long().remainder(size)
    return cmap[idx, :]


def _domain_colour(re, im, cmap, degree):  # shader.py:107-118
    bright = torch.stack([re, im], dim=-1).pow(2).sum(dim=-1, keepdim=True).pow(1 / 2)
    return bright * _angle_colour(re, im, cmap, degree)


def shade_tangent(cam_q_conj, dirs, normals, cmap, degree=1):  # shader.py:125-150
    across = normals - (normals * dirs).sum(dim=-1, keepdim=True) * dirs
    screen = q_rotation(across, cam_q_conj)
    return _domain_colour(screen[..., 0], screen[..., 1], cmap, degree)


def shade_spin(cam_q_conj, normals, cmap, degree=1):  # shader.py:157-171
    v = q_multiply(F.pad(normals, [1, 0], value=0.0), cam_q_conj)
    a, bcd = v[..., 0], v[..., 1:]
    re = a.pow(2) - bcd.pow(2).sum(dim=-1)
    im = bcd.norm(dim=-1, p=2) * a * 2
    return _domain_colour(im, re, cmap, degree)     # (im, re) order as in shader.py:171


MODES = ["lambertian", "distance", "proximity", "vignette", "normal", "laplacian",
         "tangent", "spin"]


def shade(px_coords, cam_q, pixel_frames, dirs, coords, normals, lap, dists, mode, degree, cmap):
    """Mode dispatcher (shader.py:190-263); ``mode`` is taken modulo 8."""
    name = MODES[mode % len(MODES)]
    if name == "lambertian":
        return shade_lambertian(dirs, normals)
    if name == "distance":
        return shade_distance(px_coords, coords)
    if name == "proximity":
        return shade_proximity(dists)
    if name == "vignette":
        return shade_vignette(dirs, pixel_frames)
    if name == "normal":
        return shade_normal(normals)
    if name == "laplacian":
        return shade_laplacian(lap)
    conj = q_conjugate(cam_q)[:, None, None, :]
    if name == "tangent":
        return shade_tangent(conj, dirs, normals, cmap, degree)
    return shade_spin(conj, normals, cmap, degree)


# --------------------------------------------------------------------------- L3
# control.py (cannot be imported: needs pyautogui/pynput) -- restated


def reflect(normals, dirs):
    """``v - 2 (v.n) n`` written as the reference does (control.py:179-194)."""
    return (normals * (dirs * -1)).sum(dim=-1, keepdim=True) * normals * 2 + dirs


class RenderLoopOracle:
    """camera -> march -> f(p) -> stencil -> shader (control.py:197-258)."""

    def __init__(self, scene, num_cameras=1, px_width=800, px_height=800, focal_length=17e-3,
                 sensor_width=17e-3, sensor_height=17e-3, normals_eps=5e-2, cmap=None):
        self.scene = scene
        self.N, self.W, self.H = num_cameras, px_width, px_height
        self.eps = normals_eps
        self.base = camera_base_rays(num_cameras, px_width, px_height, focal_length,
                                     sensor_width, sensor_height)
        self.cmap = cmap

    def gbuffer(self, orientations, translations, marching_steps=32):
        px, frames, pos0, dirs = camera_rays(*self.base, orientations, translations)
        pos = march(self.scene, pos0, dirs, marching_steps)
        dists = sdf(self.scene, pos)
        normals, lap = normals_laplacian(self.scene, pos, self.eps)
        return dict(px=px, frames=frames, dirs=dirs, pos=pos, dists=dists,
                    normals=normals, lap=lap)

    def __call__(self, orientations, translations, mode=0, degree=1, marching_steps=32):
        g = self.gbuffer(orientations, translations, marching_steps)
        img = shade(g["px"], orientations, g["frames"], g["dirs"], g["pos"], g["normals"],
                    g["lap"], g["dists"], mode, degree, self.cmap)
        return img.expand(-1, self.H, self.W, 3)


def reflected_march(scene, pos, dirs, steps, eps, bounces, tau=1e-3, delta=None):
    """Multi-bounce composition for BASELINE config 3.

    The reference ships no multi-bounce loop (its marcher has no reflected branch and
    ``make_reflection_directions`` is never called), so the semantics are *defined* here
    out of reference parts: march; stencil normal; a ray "hits" when ``|f(p)| < tau``; hit
    rays get ``v <- reflect(n, v)`` and are lifted off the surface by ``delta * n``; then
    the next segment marches.  Parity for this branch is against this composition.
    Returns the final (pos, dirs, hit mask of the last tested segment, normals, lap).
    """
    if delta is None:
        delta = 2 * eps
    hit = None
    for b in range(bounces + 1):
        pos = march(scene, pos, dirs, steps)
        normals, lap = normals_laplacian(scene, pos, eps)
        if b == bounces:
            break
        hit = sdf(scene, pos).abs() < tau
        dirs = torch.where(hit, reflect(normals, dirs), dirs)
        pos = torch.where(hit, pos + delta * normals, pos)
    return pos, dirs, hit, normals, lap


def config4_loss(g, target):
    """Inverse-rendering loss of SURVEY.md §8(d): Lambertian MSE + raw depth MSE."""
    lam = shade_lambertian(g["dirs"], g["normals"])
    depth = _norm(g["pos"] - g["px"])
    return ((lam - target["lam"]) ** 2).mean() + ((depth - target["depth"]) ** 2).mean()
