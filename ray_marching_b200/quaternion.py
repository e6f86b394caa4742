"""Scalar-first quaternion helpers with the reference's names (``quaternion.py``).

These free functions are host-side glue (camera-pose updates, ``control.py:150-165``) and
act on tensors of any device.  On the render path the same algebra is evaluated inside the
CUDA kernels: the camera rotation and every scene affine transform are lowered to a 3x3
frame matrix (:func:`rotation_matrix`, the polynomial ``I + 2w[v]x + 2[v]x^2`` that
``rotation`` computes, valid for non-unit quaternions as well) and the Hamilton product of
the spin shader is fused in ``csrc/rm_shade.cuh``.
"""
from __future__ import annotations

import torch
import torch.nn as nn
from torch import Tensor


def cross_product(U: Tensor, V: Tensor) -> Tensor:
    """``U x V`` on the trailing axis (reference ``quaternion.py:6-21``)."""
    return torch.linalg.cross(U, V, dim=-1) if U.shape == V.shape else torch.linalg.cross(
        *torch.broadcast_tensors(U, V), dim=-1)


def multiply(p: Tensor, q: Tensor) -> Tensor:
    """Hamilton product ``p (x) q`` (reference ``quaternion.py:24-46``)."""
    pw, pv = p[..., :1], p[..., 1:]
    qw, qv = q[..., :1], q[..., 1:]
    w = pw * qw - (pv * qv).sum(-1, keepdim=True)
    v = pw * qv + qw * pv + cross_product(pv, qv)
    return torch.cat([w, v], dim=-1)


def conjugate(q: Tensor) -> Tensor:
    """``(w, -x, -y, -z)`` (reference ``quaternion.py:49-52``)."""
    return torch.cat([q[..., :1], -q[..., 1:]], dim=-1)


def rotation_matrix(q: Tensor) -> Tensor:
    """``R(q) (...,3,3)`` with ``rotation(V, q) == R(q) @ V`` for *any* q (unit or not).

    ``rotation`` is ``V + 2w (v x V) + 2 v x (v x V)``; collecting terms gives
    ``R = I + 2w [v]x + 2 (v v^T - |v|^2 I)``.  Used by the tape lowering, so gradients with
    respect to un-normalised scene orientations match the reference's autograd."""
    w, x, y, z = q.unbind(-1)
    xx, yy, zz = x * x, y * y, z * z
    xy, xz, yz = x * y, x * z, y * z
    wx, wy, wz = w * x, w * y, w * z
    one = torch.ones_like(w)
    rows = [
        one - 2 * (yy + zz), 2 * (xy - wz), 2 * (xz + wy),
        2 * (xy + wz), one - 2 * (xx + zz), 2 * (yz - wx),
        2 * (xz - wy), 2 * (yz + wx), one - 2 * (xx + yy),
    ]
    return torch.stack(rows, dim=-1).reshape(*q.shape[:-1], 3, 3)


def rotation(V: Tensor, q: Tensor) -> Tensor:
    """Rotate ``V (...,3)`` by ``q (...,4)`` (reference ``quaternion.py:55-72``)."""
    return (rotation_matrix(q) @ V[..., None])[..., 0]


def to_versor(V: Tensor) -> Tensor:
    """``(sqrt(1 - |V|^2), V)`` (reference ``quaternion.py:75-91``); NaN for ``|V| > 1``."""
    return torch.cat([(1 - V.square().sum(-1, keepdim=True)).sqrt(), V], dim=-1)


class QuaternionToSO3(nn.Module):
    """``q (...,4) -> (-1,3,3)`` row-major (reference ``quaternion.py:94-124``).

    Differs from :func:`rotation_matrix` for non-unit q (diagonal uses ``ww+xx-yy-zz``)."""

    def __init__(self):
        super().__init__()
        # the reference's index buffer (quaternion.py:107-112); kept so state_dict keys match
        self.register_buffer("pairs", torch.tensor([(i, j) for i in range(4) for j in range(i, 4)]))

    def forward(self, q: Tensor) -> Tensor:
        w, x, y, z = q.unbind(-1)
        m = torch.stack([
            w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (w * y + x * z),
            2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x),
            2 * (x * z - w * y), 2 * (w * x + y * z), w * w - x * x - y * y + z * z,
        ], dim=-1)
        return m.view(-1, 3, 3)
