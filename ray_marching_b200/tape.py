"""Scene-build-time lowering of an SDF module tree to a flat instruction tape.

The reference evaluates a scene by recursive ``nn.Module`` calls that materialise one
``(N,H,W,k)`` tensor per torch op (``scene/transformations.py:33-41,67-71,90-94``).  Here the
tree is walked once and turned into

* ``code``  int32 ``(n_instr, 2)``: ``(head | ordinal << 16, param_offset/4 | node_id << 16)``
* ``par``   float32 ``(n_par,)``: per-instruction parameter records, 16-byte aligned.

``code`` is static host data: the C ABI copies it into the kernel-parameter block of every
launch (constant bank), so the per-ray CUDA interpreter (``csrc/rm_eval.cuh``) decodes it on the
warp-uniform datapath; ``par`` changes under an optimiser, lives on the device and is staged
into shared memory by each block.

Lowering rules
--------------
* **Frames.** Every affine transform chain above a leaf is composed at lowering time into one
  world->local frame ``q = A p + b`` (``A = R(conj(q_n)) ... R(conj(q_1))``).  ``R`` is the same
  polynomial in the (un-normalised) quaternion as ``quaternion.rotation``
  (``quaternion.py:55-72``), so values *and* d/dq agree with the reference for non-unit q.
  Every leaf below an affine carries its own copy of the composed frame (3 float4 rows), so an
  instruction is self-contained; autograd sums the copies' gradients through the gather.
* **Unary value ops** (rounding ``f - rho``, onion ``|f| - r``; ``transformations.py:117-118``,
  ``:131-132``) are attached as post-ops to the leaf / union-close instruction they act on.
* **Unions** become ``OPEN`` ... children ... ``CLOSE``; every child instruction carries the fold
  kind of its parent (hard ``min`` or online ``logsumexp``) and its ordinal among the parent's
  children (``min`` ties go to the first child, as ``torch.min(dim)`` does).
* **Parameters are read live on every call**: ``lowered_params()`` gathers the current
  ``nn.Parameter`` values into ``par`` with a handful of batched, autograd-tracked torch ops
  (a concatenate, the frame composition, one permutation gather), so the backward kernel only
  has to return ``d loss / d par`` and autograd scatters it to ``.grad`` of every parameter.

``flops_per_eval`` follows the algorithmic counting rule of SURVEY.md §8(d).
"""
from __future__ import annotations

import dataclasses
from typing import List, Optional

import numpy as np
import torch
import torch.nn as nn

from . import quaternion as Q

# ---- instruction encoding (mirrored in csrc/rm_tape.h) -------------------------------------
OP_SPHERE, OP_BOX, OP_PLANE, OP_LINE, OP_DISK, OP_TORUS = 1, 2, 3, 4, 5, 6
OP_OPEN_MIN, OP_OPEN_SMOOTH, OP_CLOSE_MIN, OP_CLOSE_SMOOTH = 8, 9, 10, 11
FOLD_NONE, FOLD_MIN, FOLD_SMOOTH = 0, 1, 2
POST_ROUND, POST_ONION = 1, 2
MAX_POST = 3
HEAD_FRAME = 1 << 4     # leaf: a world->local frame record (3 float4) precedes the primitive record
HEAD_ROOT = 1 << 7      # OPEN/CLOSE of the outermost union: no accumulator save/restore
MAX_DEPTH = 8           # union nesting depth supported by the interpreter's spill stack
MAX_INSTR = 1024        # largest kernel-parameter code block (csrc/rm_tape.h)
MAX_ORDINAL = 0xFFFF    # children per union
MAX_PAR_F4 = 0xFFFF     # parameter records addressable by an instruction (float4 units)
WORD1_CULL = 1 << 31    # leaf: it has an entry with a finite radius in the culled union's sphere table; OPEN/CLOSE: that union
CULL_MIN_CHILDREN = 8   # smooth unions smaller than this are not worth bounding-sphere tests

_PRIM_OP = {"SDFSphere": OP_SPHERE, "SDFBox": OP_BOX, "SDFPlane": OP_PLANE,
            "SDFLine": OP_LINE, "SDFDisk": OP_DISK, "SDFTorus": OP_TORUS}
_PRIM_F4 = {OP_SPHERE: 1, OP_BOX: 1, OP_PLANE: 0, OP_LINE: 3, OP_DISK: 1, OP_TORUS: 1}
# algorithmic flops per node, SURVEY.md §8(d)
_PRIM_FLOPS = {OP_SPHERE: 7, OP_BOX: 19, OP_PLANE: 0, OP_LINE: 23, OP_DISK: 10, OP_TORUS: 10}
AFFINE_FLOPS, ROUND_FLOPS, ONION_FLOPS = 33, 1, 2


def encode_head(op, frame=False, fold=0, posts=()):
    """Low 16 bits of word 0: op[0:4] frame[4] fold[5:7] root[7] npost[8:10] post kinds[10:16]."""
    head = op | (HEAD_FRAME if frame else 0) | (fold << 5) | (len(posts) << 8)
    for i, kind in enumerate(posts):
        head |= kind << (10 + 2 * i)
    return head


def decode(word0: int, word1: int):
    """(op, has_frame, fold, root, post kinds, ordinal, par offset in float4, node id, cull flag)."""
    word0 &= 0xFFFFFFFF
    word1 &= 0xFFFFFFFF
    npost = (word0 >> 8) & 3
    posts = [(word0 >> (10 + 2 * i)) & 3 for i in range(npost)]
    return (word0 & 15, bool(word0 & HEAD_FRAME), (word0 >> 5) & 3, bool(word0 & HEAD_ROOT), posts,
            word0 >> 16, word1 & 0xFFFF, (word1 >> 16) & 0x7FFF, bool(word1 & WORD1_CULL))


class _Src:
    """Where a lowered float comes from: piece name + index inside that piece."""
    ZERO = ("zero", 0)


@dataclasses.dataclass
class _Affine:
    q_slot: int            # start of the quaternion in theta
    t_slot: int            # start of the translation in theta
    parent: int            # index of the enclosing affine (-1: world)
    level: int             # chain depth (0 for a top-level affine)


def _inverse_3x3(A: torch.Tensor) -> torch.Tensor:
    """Inverse of (n, 3, 3) matrices by the adjugate (rows r0, r1, r2: columns of the inverse are
    r1 x r2, r2 x r0, r0 x r1 over the determinant); elementwise torch only."""
    r0, r1, r2 = A[:, 0], A[:, 1], A[:, 2]
    c0, c1, c2 = torch.linalg.cross(r1, r2), torch.linalg.cross(r2, r0), torch.linalg.cross(r0, r1)
    det = (r0 * c0).sum(-1)
    return torch.stack([c0, c1, c2], dim=-1) / det[:, None, None]


def _singular_range_3x3(A: torch.Tensor):
    """(upper bound of the largest, lower bound of the smallest) singular value of (n, 3, 3) matrices:
    the trigonometric closed form for the eigenvalues of the symmetric A^T A, moved outwards by 1e-3
    relative -- far more than its float32 error, far less than anything the bounds are used for
    (culling margins).  Elementwise torch only."""
    M = A.transpose(-1, -2) @ A
    m00, m11, m22 = M[:, 0, 0], M[:, 1, 1], M[:, 2, 2]
    m01, m02, m12 = M[:, 0, 1], M[:, 0, 2], M[:, 1, 2]
    q = (m00 + m11 + m22) / 3.0
    p1 = m01 * m01 + m02 * m02 + m12 * m12
    p2 = (m00 - q) ** 2 + (m11 - q) ** 2 + (m22 - q) ** 2 + 2.0 * p1
    p = torch.sqrt(p2 / 6.0)
    ps = torch.where(p > 0, p, torch.ones_like(p))
    b00, b11, b22 = (m00 - q) / ps, (m11 - q) / ps, (m22 - q) / ps
    b01, b02, b12 = m01 / ps, m02 / ps, m12 / ps
    det_b = (b00 * (b11 * b22 - b12 * b12) - b01 * (b01 * b22 - b12 * b02) + b02 * (b01 * b12 - b11 * b02))
    phi = torch.acos(torch.clamp(det_b / 2.0, -1.0, 1.0)) / 3.0
    big = q + 2.0 * p * torch.cos(phi)
    small = q + 2.0 * p * torch.cos(phi + 2.0943951023931953)          # + 2 pi / 3
    # a scaled rotation (p ~ 0) has a triple eigenvalue q; the margin also covers the cancellation in `small`
    tol = 1e-3 * big + 1e-6
    return torch.sqrt(big + tol), torch.sqrt(torch.clamp(small - tol, min=0.0))


class Tape:
    """Lowered scene: static code + the recipe that gathers live parameters into ``par``."""

    def __init__(self, root: nn.Module):
        self.root = root
        self.params: List[nn.Parameter] = []
        self._slot_of = {}
        off = 0
        for p in root.parameters():
            if id(p) in self._slot_of:
                continue
            self._slot_of[id(p)] = off
            self.params.append(p)
            off += p.numel()
        self.n_theta = off

        self._code: List[List[int]] = []
        self._src: List[tuple] = []        # per par float: (piece, index)
        self._affines: List[_Affine] = []
        self._lines: List[tuple] = []      # (start_slot, end_slot)
        self.n_nodes = 0
        self.n_leaves = 0
        self.max_depth = 0
        self.flops_per_eval = 0
        self.leaf_kinds: List[int] = []

        # forward culling: the largest smooth union gets bounding spheres for its
        # leaf children, so the march can skip children whose softmax term is provably negligible
        smooth = [m for m in root.modules() if type(m).__name__ == "SDFSmoothUnion"]
        big = max(smooth, key=lambda m: len(m.sdfs), default=None)
        self._cull_union = big if big is not None and len(big.sdfs) >= CULL_MIN_CHILDREN else None
        self._cull_leaves: List[dict] = []
        self._cull_table = 0
        self._in_cull = False

        self._ord = [0]                     # next child ordinal of every open union (root: dummy)
        self._emit(root, frame=-1, fold=FOLD_NONE, posts=[], depth=0)

        self.n_instr = len(self._code)
        self.n_par = len(self._src)
        assert self.n_par % 4 == 0
        if self.n_instr > MAX_INSTR:
            raise NotImplementedError(
                f"scene lowers to {self.n_instr} instructions; the kernels take at most {MAX_INSTR}")
        if self.n_par // 4 > MAX_PAR_F4:
            raise NotImplementedError("scene parameter block exceeds the 16-bit record offsets")
        words = []
        for head, off4, ordinal, node in self._code:
            if ordinal > MAX_ORDINAL:
                raise NotImplementedError(f"a union with more than {MAX_ORDINAL} children")
            words.append([head | (ordinal << 16), off4 | ((node & 0x7FFF) << 16) | (node & WORD1_CULL)])
        # uint32 bit patterns stored as int32 (two's complement), contiguous (n_instr, 2)
        self.code = np.ascontiguousarray(np.asarray(words, dtype=np.uint32).reshape(-1, 2)
                                         .view(np.int32))
        if self.max_depth > MAX_DEPTH:
            raise NotImplementedError(
                f"union nesting depth {self.max_depth} exceeds the interpreter limit {MAX_DEPTH}")
        self._finalise_recipe()
        self._dev_cache = {}
        self._par_cache = None
        self.spec_lib = None       # ctypes handle of a tape-specialised library (specialise.py), opt-in
        self._code_host = None

    # ------------------------------------------------------------------ tree walk
    def _slot(self, p: nn.Parameter) -> int:
        return self._slot_of[id(p)]

    def _f4(self, entries):
        """Append one 16-byte record; returns its offset in float4 units."""
        assert len(entries) <= 4
        off4 = len(self._src) // 4
        self._src.extend(list(entries) + [_Src.ZERO] * (4 - len(entries)))
        return off4

    def _theta(self, p, i=0):
        return ("theta", self._slot(p) + i)

    def _post_record(self, posts):
        kinds, entries = [], []
        for kind, p in reversed(posts):          # innermost op is applied first
            kinds.append(kind)
            entries.append(self._theta(p))
        if len(kinds) > MAX_POST:
            raise NotImplementedError(
                f"more than {MAX_POST} chained rounding/onion ops on one node are not supported")
        return kinds, entries

    def _next_ordinal(self) -> int:
        o = self._ord[-1]
        self._ord[-1] += 1
        return o

    def _emit(self, node, frame, fold, posts, depth):
        kind = type(node).__name__
        if kind == "SDFAffineTransformation":
            lvl = 0 if frame < 0 else self._affines[frame].level + 1
            self._affines.append(_Affine(self._slot(node.orientation),
                                         self._slot(node.translation), frame, lvl))
            self.flops_per_eval += AFFINE_FLOPS
            return self._emit(node.sdf, len(self._affines) - 1, fold, posts, depth)
        if kind == "SDFRounding":
            self.flops_per_eval += ROUND_FLOPS
            return self._emit(node.sdf, frame, fold, posts + [(POST_ROUND, node.rounding)], depth)
        if kind == "SDFOnion":
            self.flops_per_eval += ONION_FLOPS
            return self._emit(node.sdf, frame, fold, posts + [(POST_ONION, node.radius)], depth)
        if kind in ("SDFUnion", "SDFSmoothUnion"):
            smooth = kind == "SDFSmoothUnion"
            n = len(node.sdfs)
            if n == 0:
                raise ValueError(f"{kind} without children")
            self.flops_per_eval += (5 * n + 1) if smooth else (n - 1)
            if not smooth and fold == FOLD_MIN and not posts:
                # a hard union directly inside a hard union: min is associative and the first-index
                # tie rule is preserved by keeping the order, so the children are folded straight
                # into the parent's accumulator (no OPEN/CLOSE, no extra nesting level)
                for child in node.sdfs:
                    self._emit(child, frame, FOLD_MIN, [], depth)
                return
            node_id = self.n_nodes
            self.n_nodes += 1
            self.max_depth = max(self.max_depth, depth + 1)
            ordinal = self._next_ordinal()
            root = HEAD_ROOT if depth == 0 else 0      # nothing to save / restore around the root union
            culled = node is self._cull_union
            flag = WORD1_CULL if culled else 0
            open_index = len(self._code)
            if smooth:
                # (-k, s_min, L_max, log2 n): the last three are 0 unless this is the culled union
                off4 = self._f4([("neg_theta", self._slot(node.blend_k))] +
                                ([("cullu", 0), ("cullu", 1), ("cullu", 2)] if culled else []))
                self._code.append([encode_head(OP_OPEN_SMOOTH) | root, off4, 0, node_id | flag])
            else:
                self._code.append([encode_head(OP_OPEN_MIN) | root, 0, 0, node_id])
            child_fold = FOLD_SMOOTH if smooth else FOLD_MIN
            self._ord.append(0)
            for child in node.sdfs:
                self._in_cull = culled                  # only DIRECT children carry a bounding sphere
                self._emit(child, frame, child_fold, [], depth + 1)
            self._in_cull = False
            self._ord.pop()
            if culled:
                # Sphere table of the culled union: one float4 (world centre, radius / s_min) per instruction of
                # its body, in instruction order, contiguous -- lane L of a warp reads entry L of a chunk
                # (csrc/rm_eval.cuh).  Instructions without a bounding sphere (planes, nested unions and their
                # leaves) get an infinite radius: never skipped, and they void the whole-union bound.
                sphere_of = {c["code_index"]: ci for ci, c in enumerate(self._cull_leaves)}
                self._cull_table = len(self._src) // 4
                for j in range(open_index + 1, len(self._code)):
                    ci = sphere_of.get(j)
                    self._f4([("cull", ci * 4 + i) for i in range(4)] if ci is not None
                             else [_Src.ZERO, _Src.ZERO, _Src.ZERO, ("big", 0)])
            if culled:
                self._code[open_index][2] = self._cull_table     # ordinal field of the OPEN = table offset (float4)
            kinds, entries = self._post_record(posts)
            off4 = self._f4(entries) if kinds else 0
            op = OP_CLOSE_SMOOTH if smooth else OP_CLOSE_MIN
            self._code.append([encode_head(op, False, fold, kinds) | root, off4, ordinal, node_id | flag])
            return
        if kind not in _PRIM_OP:
            raise NotImplementedError(f"cannot lower scene node of type {kind}")
        op = _PRIM_OP[kind]
        self.flops_per_eval += _PRIM_FLOPS[op]
        self.n_leaves += 1
        self.leaf_kinds.append(op)
        cull = self._in_cull and op != OP_PLANE          # a half-space has no bounding sphere
        self._in_cull = False
        if cull:
            ci = len(self._cull_leaves)
            pz = self.n_theta                                      # index of the padding zero
            info = dict(frame=frame, add=[pz] * 4, box=[pz] * 3, ls=[pz] * 3, le=[pz] * 3,
                        code_index=len(self._code))                # the instruction this sphere bounds
            adds = [self._slot(p) for _, p in posts]
            if op in (OP_SPHERE, OP_DISK):
                adds.append(self._slot(node.radius))
            elif op == OP_TORUS:
                adds += [self._slot(node.radius1), self._slot(node.radius2)]
            elif op == OP_BOX:
                info["box"] = [self._slot(node.halfsides) + i for i in range(3)]
            elif op == OP_LINE:
                adds.append(self._slot(node.radius))
                info["ls"] = [self._slot(node.start) + i for i in range(3)]
                info["le"] = [self._slot(node.end) + i for i in range(3)]
            if len(adds) > 4:
                raise NotImplementedError("too many radius terms for a bounding sphere")
            info["add"] = adds + [pz] * (4 - len(adds))
            self._cull_leaves.append(info)
        first = len(self._src) // 4
        if frame >= 0:
            for r in range(3):
                self._f4([("A", frame * 9 + r * 3 + c) for c in range(3)] + [("b", frame * 3 + r)])
        if op == OP_SPHERE:
            self._f4([self._theta(node.radius)])
        elif op == OP_BOX:
            self._f4([self._theta(node.halfsides, i) for i in range(3)])
        elif op == OP_LINE:
            li = len(self._lines)
            self._lines.append((self._slot(node.start), self._slot(node.end)))
            self._f4([self._theta(node.start, i) for i in range(3)] + [self._theta(node.radius)])
            self._f4([("AB", li * 3 + i) for i in range(3)])
            self._f4([("ABn", li * 3 + i) for i in range(3)])
        elif op == OP_DISK:
            self._f4([self._theta(node.radius)])
        elif op == OP_TORUS:
            self._f4([self._theta(node.radius1), self._theta(node.radius2)])
        kinds, entries = self._post_record(posts)
        if kinds:
            self._f4(entries)
        self._code.append([encode_head(op, frame >= 0, fold, kinds), first, self._next_ordinal(),
                           WORD1_CULL if cull else 0])

    # ------------------------------------------------------------------ lowering recipe
    def _finalise_recipe(self):
        nA, nL, P = len(self._affines), len(self._lines), self.n_theta
        nC = len(self._cull_leaves)
        base = {"zero": 0, "theta": 1, "neg_theta": 1 + P, "AB": 1 + 2 * P,
                "ABn": 1 + 2 * P + 3 * nL, "A": 1 + 2 * P + 6 * nL,
                "b": 1 + 2 * P + 6 * nL + 9 * nA, "cull": 1 + 2 * P + 6 * nL + 12 * nA,
                "cullu": 1 + 2 * P + 6 * nL + 12 * nA + 4 * nC,
                "big": 1 + 2 * P + 6 * nL + 12 * nA + 4 * nC + (3 if self._cull_union is not None else 0)}
        i64 = lambda rows: torch.tensor(rows, dtype=torch.int64).reshape(len(rows), -1)
        self._cull_idx = None
        if nC:
            self._cull_idx = dict(
                frame=torch.tensor([c["frame"] for c in self._cull_leaves], dtype=torch.int64),
                add=i64([c["add"] for c in self._cull_leaves]), box=i64([c["box"] for c in self._cull_leaves]),
                ls=i64([c["ls"] for c in self._cull_leaves]), le=i64([c["le"] for c in self._cull_leaves]))
        self._cull_n = len(self._cull_union.sdfs) if self._cull_union is not None else 0
        self._perm = torch.tensor([base[piece] + idx for piece, idx in self._src],
                                  dtype=torch.int64)
        ar3, ar4 = np.arange(3), np.arange(4)
        self._q_idx = torch.tensor(np.array([a.q_slot + ar4 for a in self._affines]).reshape(-1, 4),
                                   dtype=torch.int64)
        self._t_idx = torch.tensor(np.array([a.t_slot + ar3 for a in self._affines]).reshape(-1, 3),
                                   dtype=torch.int64)
        self._ls_idx = torch.tensor(np.array([s + ar3 for s, _ in self._lines]).reshape(-1, 3),
                                    dtype=torch.int64)
        self._le_idx = torch.tensor(np.array([e + ar3 for _, e in self._lines]).reshape(-1, 3),
                                    dtype=torch.int64)
        # affine chains are composed level by level (batched); level 0 = directly under world
        levels = {}
        for i, a in enumerate(self._affines):
            levels.setdefault(a.level, []).append(i)
        self._levels = [
            (torch.tensor(idx, dtype=torch.int64),
             torch.tensor([self._affines[i].parent for i in idx], dtype=torch.int64))
            for lvl, idx in sorted(levels.items()) if lvl > 0
        ]

    def _on(self, device):
        key = str(device)
        if key not in self._dev_cache:
            mv = lambda t: t.to(device)
            self._dev_cache[key] = dict(
                perm=mv(self._perm), q_idx=mv(self._q_idx), t_idx=mv(self._t_idx),
                ls_idx=mv(self._ls_idx), le_idx=mv(self._le_idx),
                levels=[(mv(a), mv(b)) for a, b in self._levels],
                cull=None if self._cull_idx is None else {k: mv(v) for k, v in self._cull_idx.items()},
            )
        return self._dev_cache[key]

    @property
    def device(self):
        return self.params[0].device if self.params else torch.device("cpu")

    def code_tensor(self, device=None) -> torch.Tensor:
        """The instruction words as a HOST int32 tensor ``(n_instr, 2)`` (``device`` is ignored: the
        C ABI reads the code on the host and passes it to the kernels by value)."""
        if self._code_host is None:
            self._code_host = torch.from_numpy(self.code)
        return self._code_host

    def theta(self, device=None) -> torch.Tensor:
        """Flat view of the current raw parameters (autograd-connected)."""
        if not self.params:
            return torch.zeros(0, device=device or self.device)
        return torch.cat([p.reshape(-1) for p in self.params])

    def lower_theta(self, theta: torch.Tensor) -> torch.Tensor:
        """``theta (P,) -> par (n_par,)``; differentiable, device-agnostic (used on CPU in tests)."""
        d = self._on(theta.device)
        pieces = [theta.new_zeros(1), theta, -theta]
        if len(self._lines):
            s, e = theta[d["ls_idx"]], theta[d["le_idx"]]
            ab = e - s                                        # primitives.py:52
            pieces += [ab.reshape(-1), (ab / (ab * ab).sum(-1, keepdim=True)).reshape(-1)]  # :53-54
        if len(self._affines):
            q, t = theta[d["q_idx"]], theta[d["t_idx"]]
            A = Q.rotation_matrix(Q.conjugate(q))             # transformations.py:35-41
            b = -(A @ t[..., None])[..., 0]
            for idx, parent in d["levels"]:                   # nested chains: inner o outer
                Ai, Ap = A[idx], A[parent]
                b = b.index_put((idx,), (Ai @ b[parent][..., None])[..., 0] + b[idx])
                A = A.index_put((idx,), Ai @ Ap)
            pieces += [A.reshape(-1), b.reshape(-1)]
        if self._cull_union is not None:
            pieces += self._cull_pieces(theta, d, A if len(self._affines) else None,
                                        b if len(self._affines) else None)
            pieces += [theta.new_full((1,), 3.0e38)]          # "big": the radius of an instruction without a sphere
        return torch.cat(pieces)[d["perm"]]

    def _cull_pieces(self, theta, d, A, b):
        """Bounding spheres of the culled union's leaf children (not differentiated): world centre and
        radius / s_min per leaf, and (s_min, L_max, log2 n) for the union record, where every frame
        satisfies s_min |dx| <= |A dx| <= L_max |dx|.  Plain elementwise torch only (closed-form
        singular values and inverse of the 3x3 frames): the lowering runs inside captured CUDA graphs
        (the optimisation step of config 4b), where the LAPACK-backed ``torch.linalg`` calls cannot.
        The forward culling kernels and the reverse sweep of the fused loss kernel (dead children of
        the union, csrc/rm_sweep.cuh) read them."""
        import math
        nC = len(self._cull_leaves)
        with torch.no_grad():
            theta = theta.detach()
            c = d["cull"]
            tp = torch.cat([theta, theta.new_zeros(1)])
            radius = tp[c["add"]].sum(-1) + tp[c["box"]].norm(dim=-1)
            s, e = tp[c["ls"]], tp[c["le"]]
            radius = radius + 0.5 * (e - s).norm(dim=-1)
            centre = 0.5 * (s + e)                                   # local centre (capsule midpoint)
            smin = theta.new_ones(nC)
            lmax = theta.new_ones(())
            if A is not None:
                A = A.detach()
                smax_a, smin_a = _singular_range_3x3(A)
                lmax = torch.clamp(smax_a.max(), min=1.0)
                has = c["frame"] >= 0
                f = c["frame"].clamp(min=0)
                world = (_inverse_3x3(A)[f] @ (centre - b.detach()[f])[..., None])[..., 0]
                centre = torch.where(has[:, None], world, centre)
                smin = torch.where(has, smin_a[f], smin)
            smin_all = torch.clamp(smin.min(), max=1.0) if nC else theta.new_ones(())
            rec = torch.cat([centre, (radius / smin)[:, None]], dim=-1).reshape(-1)
            union = torch.stack([smin_all, lmax, theta.new_tensor(math.log2(max(self._cull_n, 1)))])
            return [rec, union]

    def invalidate(self) -> None:
        """Drops the cached lowering.  Needed only after writes autograd cannot see -- in-place
        writes through ``p.data`` (``p.data.fill_(..)``, ``p.data.clamp_(..)``) share the storage but
        not the version counter.  Optimiser steps, ``load_state_dict``, ``p.copy_()`` under
        ``no_grad`` and ``p.data = x`` are detected without this."""
        self._par_cache = None

    def lowered_params(self, device=None) -> torch.Tensor:
        """Current ``par`` on the parameters' device (``device`` only matters for a scene
        without parameters).  Cached while no parameter changes and autograd is off (a static
        scene under a moving camera costs nothing per frame).  The cache key is every parameter's
        version counter AND storage address; see :meth:`invalidate` for the one kind of write it
        cannot see."""
        if self.params and device is not None and not _same_device(device, self.device):
            raise RuntimeError(f"scene parameters live on {self.device} but the query is on "
                               f"{device}; move the scene with .to(device)")
        track = torch.is_grad_enabled() and any(p.requires_grad for p in self.params)
        if not track:
            key = (tuple((p._version, p.data_ptr()) for p in self.params), str(device or self.device))
            if self._par_cache is not None and self._par_cache[0] == key:
                return self._par_cache[1]
            with torch.no_grad():
                par = self.lower_theta(self.theta(device)).contiguous()
            self._par_cache = (key, par)
            return par
        return self.lower_theta(self.theta(device)).contiguous()

    def describe(self) -> str:
        return (f"Tape(instr={self.n_instr}, par={self.n_par} floats, leaves={self.n_leaves}, "
                f"unions={self.n_nodes}, depth={self.max_depth}, theta={self.n_theta}, "
                f"flops/eval={self.flops_per_eval})")


def _same_device(a, b) -> bool:
    """``cuda`` (no index) names the current device."""
    a, b = torch.device(a), torch.device(b)
    if a.type != b.type:
        return False
    if a.type != "cuda":
        return True
    cur = torch.cuda.current_device() if (a.index is None or b.index is None) else 0
    return (cur if a.index is None else a.index) == (cur if b.index is None else b.index)


def tape_of(module: nn.Module) -> Tape:
    """Tape of ``module``'s sub-tree, cached on the module (structure is assumed immutable)."""
    tape = module.__dict__.get("_rm_tape")
    if tape is None:
        tape = Tape(module)
        module.__dict__["_rm_tape"] = tape
    return tape
