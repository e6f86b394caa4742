"""Tape-specialised render kernels: the scene STRUCTURE compiled into the kernel at scene-build time.

The general path interprets the tape (csrc/rm_eval.cuh).  For small tapes the interpreter's control
-- loop, dispatch chain, tail selection -- is ~40 % of the issued instructions (profiles/r01/
variants.md).  This module emits, for one tape, the straight-line sequence of exactly the statements
the interpreter would execute (the same hand-written primitive bodies of csrc/rm_prims.cuh, the same
fold order, constant record offsets), and compiles the unchanged forward translation unit around it
with nvcc for sm_100a into ``build/spec/librm_spec_<hash>.so``, which exports ``rm_render`` and the
backward entry points (``rm_render_bwd``, ``rm_march_bwd``, ``rm_sdf_vjp``) with the ABI of
include/rm_b200.h.  The backward kernels are specialised by the compiler: the header carries the tape
words as a switch, their tape loops are fully unrolled and every decode constant-folds (rm_eval.cuh).  Scene PARAMETERS stay run-time data (staged from ``par`` as before): an
optimiser step never recompiles.  Structure hash = FNV-1a of the instruction words, checked by the
library at every launch.

Opt-in: ``RenderLoop.specialise()`` / :func:`specialised_lib`.  Without nvcc, or for tapes above
``MAX_SPEC_INSTR``, callers keep the interpreter.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Optional

import numpy as np

from . import _lib, tape as T

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
SPEC_DIR = os.path.abspath(os.environ.get("RM_B200_SPEC_DIR", os.path.join(ROOT, "build", "spec")))
MAX_SPEC_INSTR = 40          # straight-line code beyond this stops fitting the instruction cache
MAX_SPEC_INSTR_BWD = int(os.environ.get("RM_B200_SPEC_BWD_LIMIT", "40"))   # backward-only libraries up to here
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

_PRIM_CALL = {
    T.OP_SPHERE: "prim_sphere<FAST>(P[{p}], x, y, z)",
    T.OP_BOX: "prim_box<FAST>(P[{p}], x, y, z)",
    T.OP_TORUS: "prim_torus<FAST>(P[{p}], x, y, z)",
    T.OP_DISK: "prim_disk<FAST>(P[{p}], x, y, z)",
    T.OP_LINE: "prim_line<FAST>(P[{p}], P[{p1}], P[{p2}], x, y, z)",
    T.OP_PLANE: "x",
}
_PRIM_F4 = {T.OP_SPHERE: 1, T.OP_BOX: 1, T.OP_PLANE: 0, T.OP_LINE: 3, T.OP_DISK: 1, T.OP_TORUS: 1}


def code_hash(code: np.ndarray) -> int:
    """FNV-1a over the instruction words (csrc/rm_host.h: tape_code_hash)."""
    h = 2166136261
    for b in np.ascontiguousarray(code, dtype=np.int32).tobytes():
        h = ((h ^ b) * 16777619) & 0xFFFFFFFF
    return h


def generate_source(tape) -> str:
    """The generated header: ``eval_spec<K, FAST>`` for this tape + its structure hash."""
    out: List[str] = []
    emit = out.append
    stack: List[int] = []                       # ids of the open unions, innermost last
    n_unions = 0

    folded = {}                                 # union id -> children folded so far

    def fold_into_parent(fold: int):
        if fold == T.FOLD_NONE:
            return
        u = stack[-1]
        first = folded.get(u, 0) == 0
        folded[u] = folded.get(u, 0) + 1
        if first:
            # first child: min(+inf, d) = d, resp. the online log-sum-exp started at (x, 1) -- what
            # the general statements below give from the initial (-inf, 0) for every finite x
            if fold == T.FOLD_MIN:
                emit(f"    RM_FORK a0_{u}[k] = d[k];")
            else:
                emit(f"    RM_FORK {{ a0_{u}[k] = c_{u} * d[k]; a1_{u}[k] = 1.f; }}")
            return
        if fold == T.FOLD_MIN:
            emit(f"    RM_FORK a0_{u}[k] = min_nan(a0_{u}[k], d[k]);")
        else:
            emit(f"    RM_FORK {{ const float x = c_{u} * d[k]; const float e = M::exp(-fabsf(x - a0_{u}[k]));")
            emit(f"              a1_{u}[k] = (x > a0_{u}[k]) ? fmaf(a1_{u}[k], e, 1.f) : a1_{u}[k] + e;")
            emit(f"              a0_{u}[k] = fmaxf(a0_{u}[k], x); }}")

    def posts_of(kinds_, rec):
        if not kinds_:
            return
        emit(f"    {{ const float4 pp = P[{rec}];")
        for i, kind in enumerate(kinds_):
            comp = "xyz"[i]
            if kind == T.POST_ONION:
                emit(f"      RM_FORK d[k] = fabsf(d[k]) - pp.{comp};")
            else:
                emit(f"      RM_FORK d[k] = d[k] - pp.{comp};")
        emit("    }")

    for w0, w1 in np.asarray(tape.code).tolist():
        op, frame, fold, root, posts, ordinal, off, node, cull = T.decode(w0, w1)
        if op in (T.OP_OPEN_MIN, T.OP_OPEN_SMOOTH):
            u = n_unions
            n_unions += 1
            stack.append(u)
            if op == T.OP_OPEN_MIN:
                emit(f"    float a0_{u}[K];")
            else:
                emit(f"    const float c_{u} = M::exp_scale(P[{off}].x);")
                emit(f"    float a0_{u}[K], a1_{u}[K];")
            continue
        if op in (T.OP_CLOSE_MIN, T.OP_CLOSE_SMOOTH):
            u = stack.pop()
            if not folded.get(u):
                raise ValueError("union without children cannot be specialised")
            if op == T.OP_CLOSE_MIN:
                emit(f"    RM_FORK d[k] = a0_{u}[k];")
            else:
                emit(f"    RM_FORK d[k] = M::div(a0_{u}[k] + M::log(a1_{u}[k]), c_{u});")
            posts_of(posts, off)
            fold_into_parent(fold)
            continue
        # leaf
        prim = off + (3 if frame else 0)
        call = _PRIM_CALL[op].format(p=prim, p1=prim + 1, p2=prim + 2)
        if frame:
            emit(f"    {{ const float4 r0 = P[{off}], r1 = P[{off + 1}], r2 = P[{off + 2}];")
            emit("      RM_FORK { const float x = fmaf(r0.x, px[k], fmaf(r0.y, py[k], fmaf(r0.z, pz[k], r0.w)));")
            emit("                const float y = fmaf(r1.x, px[k], fmaf(r1.y, py[k], fmaf(r1.z, pz[k], r1.w)));")
            emit("                const float z = fmaf(r2.x, px[k], fmaf(r2.y, py[k], fmaf(r2.z, pz[k], r2.w)));")
            emit(f"                (void)y; (void)z; d[k] = {call}; }} }}")
        else:
            emit("    RM_FORK { const float x = px[k], y = py[k], z = pz[k]; (void)y; (void)z;")
            emit(f"              d[k] = {call}; }}")
        posts_of(posts, prim + _PRIM_F4[op])
        fold_into_parent(fold)
    assert not stack, "unbalanced tape"
    body = "\n".join(out)
    words = np.asarray(tape.code, dtype=np.int64).reshape(-1, 2)
    cases = "\n".join(f"        case {2 * i + j}: return (int)0x{int(w) & 0xFFFFFFFF:08x}u;"
                      for i, row in enumerate(words.tolist()) for j, w in enumerate(row))
    return f"""// GENERATED by ray_marching_b200/specialise.py for {tape.describe()}
// Straight-line form of the tape: the statements csrc/rm_eval.cuh executes for it, in order.
#pragma once
#define RM_SPEC_HASH 0x{code_hash(tape.code):08x}u
#define RM_SPEC_N {tape.n_instr}
namespace rm {{
// the tape words (rm_tape.h encoding): constant-folded by the unrolled tape loops of the recording
// interpreter and the reverse sweep
__host__ __device__ __forceinline__ int spec_word(int pc, int j) {{
    switch (pc * 2 + j) {{
{cases}
        default: return 0;
    }}
}}
}}  // namespace rm
#define RM_FORK _Pragma("unroll") for (int k = 0; k < K; ++k)
namespace rm {{
template <int K, bool FAST>
__device__ __forceinline__ void eval_spec(const float4* __restrict__ P, const float (&px)[K],
                                          const float (&py)[K], const float (&pz)[K], float (&out)[K]) {{
    using M = Math<FAST>;
    float d[K];
    RM_FORK d[k] = 0.f;
{body}
    RM_FORK out[k] = d[k];
}}
}}  // namespace rm
#undef RM_FORK
"""


def backward_only(tape) -> bool:
    """Tapes between the two limits get a library with the backward entry points only (the forward
    keeps the interpreter, which already amortises its decode over 4 rays per thread)."""
    return MAX_SPEC_INSTR < tape.n_instr <= MAX_SPEC_INSTR_BWD


def spec_paths(tape):
    h = code_hash(tape.code)
    stem = "librm_specb" if backward_only(tape) else "librm_spec"
    return os.path.join(SPEC_DIR, f"rm_spec_{h:08x}.inc"), os.path.join(SPEC_DIR, f"{stem}_{h:08x}.so")


def build(tape, force: bool = False) -> str:
    """Generates and compiles the specialised library of ``tape`` (cached by structure hash)."""
    if tape.n_instr > MAX_SPEC_INSTR_BWD:
        raise ValueError(f"tape has {tape.n_instr} instructions; specialisation is for <= {MAX_SPEC_INSTR} "
                         f"(forward + backward) or <= {MAX_SPEC_INSTR_BWD} (backward only)")
    inc, lib = spec_paths(tape)
    src = generate_source(tape)
    # freshness by CONTENT (file times do not survive a copy to another machine): digest of the
    # generated header, every kernel source and the extra flags, kept beside the library
    digest = _source_digest(src)
    stamp = lib + ".digest"
    fresh = (os.path.exists(lib) and os.path.exists(inc) and os.path.exists(stamp)
             and open(stamp).read().strip() == digest and open(inc).read() == src)
    if fresh and not force:
        return lib
    if not os.path.exists(NVCC):
        raise RuntimeError(f"nvcc not found at {NVCC}: cannot specialise (the interpreter path is unaffected)")
    os.makedirs(SPEC_DIR, exist_ok=True)
    with open(inc, "w") as f:
        f.write(src)
    tmp = lib + f".tmp{os.getpid()}"
    cmd = [NVCC, "-shared", "-o", tmp, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo",
           "-std=c++17", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", f'-DRM_SPEC_INC="{inc}"',
           *os.environ.get("RM_B200_SPEC_FLAGS", "").split(),      # launch-shape experiments
           os.path.join(CSRC, "rm_host.cu"),
           *([] if backward_only(tape) else [os.path.join(CSRC, "rm_forward.cu")]),
           os.path.join(CSRC, "rm_backward.cu")]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed for the specialised library:\n" + res.stdout + res.stderr)
    os.replace(tmp, lib)
    with open(stamp, "w") as f:
        f.write(digest + "\n")
    return lib


def _source_digest(generated: str) -> str:
    import hashlib
    h = hashlib.sha1()
    h.update(generated.encode())
    h.update(os.environ.get("RM_B200_SPEC_FLAGS", "").encode())
    deps = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h")))
    deps.append(os.path.join(ROOT, "include", "rm_b200.h"))
    for path in deps:
        with open(path, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


_loaded = {}
# entry points a specialised library exports besides rm_abi_version / rm_last_error
SPEC_EXPORTS = ("rm_render", "rm_render_bwd", "rm_render_loss_bwd", "rm_march_bwd", "rm_sdf_vjp",
                "rm_bwd_scratch_floats", "rm_loss_scratch_floats")

# Libraries selected by RenderLoop.specialise(), per TAPE OBJECT: the autograd ops (ops.py:
# render_gbuffer / render_bwd) receive the tape's host code tensor, not the Tape, so the key is that
# tensor's address (owned by the tape for its whole life; the entry dies with the tape).  Two
# RenderLoops over scenes of the same structure hold different tapes and choose independently.
_selected = {}


def select(tape, handle: Optional[C.CDLL]) -> None:
    import weakref
    key = tape.code_tensor().data_ptr()
    if handle is None:
        _selected.pop(key, None)
        return
    if key not in _selected:
        weakref.finalize(tape, _selected.pop, key, None)
    _selected[key] = handle


def selected_for(code) -> Optional[C.CDLL]:
    """The specialised library selected for the tape this code block (a host int32 tensor) belongs to."""
    if not _selected:
        return None
    return _selected.get(code.data_ptr())


def spec_check(handle: C.CDLL, status: int, what: str) -> None:
    if status != 0:
        raise _lib.RmError(f"{what} (specialised) failed with status {status}: "
                           f"{handle.rm_last_error().decode(errors='replace')}")


def specialised_lib(tape, build_if_missing: bool = True) -> Optional[C.CDLL]:
    """ctypes handle of the library specialised for ``tape`` (its ``rm_render`` has the signature
    of the general one), or None when it does not exist and may not be built."""
    inc, path = spec_paths(tape)
    if path in _loaded:
        return _loaded[path]
    if build_if_missing:
        path = build(tape)
    elif not (os.path.exists(path) and os.path.exists(inc) and open(inc).read() == generate_source(tape)):
        return None
    handle = C.CDLL(path)
    for name in SPEC_EXPORTS:
        if name == "rm_render" and backward_only(tape):
            continue
        fn = getattr(handle, name)
        fn.restype, fn.argtypes = _lib._SIGNATURES[name]
    handle.has_forward = not backward_only(tape)
    handle.rm_last_error.restype = C.c_char_p
    if handle.rm_abi_version() != _lib.ABI_VERSION:
        raise RuntimeError(f"{path}: ABI version mismatch; rebuild")
    _loaded[path] = handle
    return handle
