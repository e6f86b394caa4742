"""Pure-torch executor of the lowered tape -- TEST INFRASTRUCTURE.

Mirrors ``csrc/rm_eval.cuh`` instruction by instruction (frames, post-ops, online log-sum-exp)
so that the *lowering* (``ray_marching_b200/tape.py``: encoding, frame composition, parameter
gather and its autograd) can be validated on a CPU-only machine against the oracle.
"""
import torch

from ray_marching_b200 import tape as T


def _len(*comps):
    """Euclidean length with torch.norm's zero sub-gradient at the origin (as the reference)."""
    return torch.stack(comps, dim=-1).norm(dim=-1)


def run(tape, par, pts):
    """par (n_par,), pts (n,3) -> (n,1).  Differentiable w.r.t. both."""
    code = tape.code
    f4 = par.reshape(-1, 4)
    x, y, z = pts.unbind(-1)
    stack, acc = [], None
    d = None
    for w0, w1 in code.tolist():
        op, has_frame, fold, _root, posts, _ordinal, off, _node, _cull = T.decode(w0, w1)
        P = off
        if op < T.OP_OPEN_MIN:
            qx, qy, qz = x, y, z
            if has_frame:
                r0, r1, r2 = f4[P], f4[P + 1], f4[P + 2]
                P += 3
                qx = r0[0] * x + r0[1] * y + r0[2] * z + r0[3]
                qy = r1[0] * x + r1[1] * y + r1[2] * z + r1[3]
                qz = r2[0] * x + r2[1] * y + r2[2] * z + r2[3]
            if op == T.OP_SPHERE:
                d = _len(qx, qy, qz) - f4[P][0]
            elif op == T.OP_BOX:
                h = f4[P]
                ax, ay, az = qx.abs() - h[0], qy.abs() - h[1], qz.abs() - h[2]
                a = torch.stack([ax, ay, az], dim=-1)
                mx = a.max(dim=-1).values                       # ties -> first index, as torch.max(dim)
                d = a.where(a > 0, 0.0).norm(dim=-1) + mx.where(mx < 0, 0.0)
            elif op == T.OP_PLANE:
                d = qx
            elif op == T.OP_LINE:
                p0, p1, p2 = f4[P], f4[P + 1], f4[P + 2]
                apx, apy, apz = qx - p0[0], qy - p0[1], qz - p0[2]
                h = (apx * p2[0] + apy * p2[1] + apz * p2[2]).clamp(0, 1)
                wx, wy, wz = h * p1[0] - apx, h * p1[1] - apy, h * p1[2] - apz
                d = _len(wx, wy, wz) - p0[3]
            elif op == T.OP_DISK:
                rim = _len(qy, qz) - f4[P][0]
                d = _len(qx, rim.where(rim > 0, 0.0))
            elif op == T.OP_TORUS:
                ring = _len(qx, qz) - f4[P][0]
                d = _len(ring, qy) - f4[P][1]
            else:
                raise AssertionError(op)
            P += {T.OP_LINE: 3, T.OP_PLANE: 0}.get(op, 1)
        elif op in (T.OP_OPEN_MIN, T.OP_OPEN_SMOOTH):
            stack.append(acc)
            if op == T.OP_OPEN_MIN:
                acc = ["min", None]
            else:
                acc = ["smooth", None, None, f4[P][0]]       # running max, sum, c = -k
            continue
        else:
            if op == T.OP_CLOSE_MIN:
                d = acc[1]
            else:
                d = (acc[1] + torch.log(acc[2])) / acc[3]
            acc = stack.pop()
        for i, kind in enumerate(posts):
            d = (d.abs() if kind == T.POST_ONION else d) - f4[P][i]
        if fold == T.FOLD_MIN:
            acc[1] = d if acc[1] is None else torch.where(d < acc[1], d, acc[1])   # first child wins ties
        elif fold == T.FOLD_SMOOTH:
            xk = acc[3] * d
            if acc[1] is None:
                acc[1], acc[2] = xk, torch.ones_like(xk)
            else:
                m = torch.maximum(acc[1], xk)
                acc[2] = acc[2] * torch.exp(acc[1] - m) + torch.exp(xk - m)
                acc[1] = m
    return d[:, None]
