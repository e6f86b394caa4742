"""CPU-only: the C-ABI shared library builds, loads, and exports every symbol include/rm_b200.h declares."""
import ctypes
import os
import re

from ray_marching_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "rm_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rm_[a-z_0-9]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound():
    names = _declared()
    assert "rm_render" in names and "rm_sdf_eval" in names
    handle = _lib.lib()
    for n in names:
        assert hasattr(handle, n), f"{n} declared in rm_b200.h but not exported"
    assert sorted(_lib.exported_symbols()) == names, "ctypes signature table out of sync with header"
    assert handle.rm_abi_version() == _lib.ABI_VERSION


def test_struct_layout_matches_header():
    assert ctypes.sizeof(_lib.RmTape) == 32
    assert ctypes.sizeof(_lib.RmCamera) == 24
    assert ctypes.sizeof(_lib.RmRenderOut) == 16 * 8
    assert ctypes.sizeof(_lib.RmRenderArgs) == 88
    assert ctypes.sizeof(_lib.RmLossArgs) == 40
    assert ctypes.sizeof(_lib.RmRenderGrads) == 48


def test_argument_errors_do_not_need_a_gpu():
    handle = _lib.lib()
    tape = _lib.RmTape(0, 0, 0, 0, 0, 0)
    assert handle.rm_sdf_eval(ctypes.byref(tape), 0, 4, 0, 0) == -1
    assert b"tape" in handle.rm_last_error()


def test_shard_arguments_are_validated_on_the_host():
    """rm_render / rm_render_loss_bwd check the (possibly interleaved) ray range before any CUDA call."""
    import numpy as np
    handle = _lib.lib()
    code = np.array([[1, 0]], dtype=np.int32)                        # one world-space sphere
    tape = _lib.RmTape(code.ctypes.data, 8, 1, 4, 0, 0)              # fake device pointer: never dereferenced
    cam = _lib.RmCamera(1, 16, 8, 1e-3, 2e-3, 1e-3)
    out = _lib.RmRenderOut()

    def render(begin, n, strip, stride):
        a = _lib.RmRenderArgs(begin, n, 4, 0, 5e-2, 1e-3, 0.1, 0, -1, 1, 0, 1, 0, 0.0, strip, stride)
        st = handle.rm_render(ctypes.byref(tape), ctypes.byref(cam), 8, ctypes.byref(a), ctypes.byref(out), 0)
        return st, handle.rm_last_error()
    assert render(0, 129, 0, 0)[0] == -1                             # contiguous range past the batch (128 rays)
    st, msg = render(0, 64, 6, 12)
    assert st == -1 and b"multiple of 4" in msg
    st, msg = render(0, 64, 8, 4)
    assert st == -1 and b"strip_stride" in msg
    st, msg = render(8, 72, 8, 16)                                   # local ray 71 = global 8 + 8*16 + 7 = 143 >= 128
    assert st == -1 and b"interleaved" in msg
    assert render(-1, 4, 0, 0)[0] == -1
    assert render(0, 0, 8, 16)[0] == 0                               # an empty shard is legal (no launch)
    loss = _lib.RmLossArgs(8, 8, 1.0, 1.0, 0)
    a = _lib.RmRenderArgs(0, 4, 4, 0, 5e-2, 0.0, 0.0, 0, -1, 1, 0, 0, 0, 0.0, 0, 0)
    assert handle.rm_render_loss_bwd(ctypes.byref(tape), ctypes.byref(cam), 8, ctypes.byref(a),
                                     ctypes.byref(loss), 8, 8, 8, 0) == -1
    assert b"loss_out" in handle.rm_last_error()


def test_cpu_tensors_are_rejected_loudly():
    import pytest
    import torch
    from ray_marching_b200 import scene as S
    with pytest.raises(_lib.RmError):
        S.SDFSphere(0.5)(torch.zeros(4, 3))


def test_loss_scratch_size_follows_the_route(monkeypatch):
    """rm_loss_scratch_floats (host arithmetic only): short tapes take the recording route (trajectory, slot
    indexed); long tapes the split route (per-ray planes of trajectory, node records and (conv, u, p_S));
    ranges whose planes would exceed 8 GB fall back to the slot-indexed node-record route; the test override
    RM_B200_LOSS_ROUTE moves the size with the route."""
    handle = _lib.lib()
    steps, block = 32, 128
    size = lambda tape, n: int(handle.rm_loss_scratch_floats(ctypes.byref(tape), n, steps))

    def geometry(n, rays_per_block):
        want = -(-n // rays_per_block)
        iters = max(1, -(-want // (148 * 8)))
        return -(-want // iters), iters                       # blocks, block-uniform iterations

    short = _lib.RmTape(0, 0, 6, 24, 1, 1)                    # scene2-like: 6 instructions
    long_ = _lib.RmTape(0, 0, 69, 1420, 2, 2)                 # c4b-like: 69 instructions, 2 union nodes
    monkeypatch.delenv("RM_B200_LOSS_ROUTE", raising=False)
    n = 1 << 20
    grid, iters = geometry(n, block)
    slots = grid * block
    assert size(short, n) == steps * slots * 3
    ncols = slots * iters
    nr = 1 + 2 * 2
    assert size(long_, n) == steps * ncols * 3 + (steps + 4) * nr * ncols + 7 * ncols + 16
    assert ncols >= n
    # 2^26 rays: the planes would take ~19 G floats -> the slot-indexed route (independent of the ray count)
    big = 1 << 26
    gridb, _ = geometry(big, block)
    assert size(long_, big) == steps * gridb * block * 3 + (steps + 4) * nr * gridb * block
    assert size(long_, big) < 1 << 31
    for route, expect in (("record", steps * slots * 3),
                          ("onfly", steps * slots * 3 + (steps + 4) * nr * slots)):
        monkeypatch.setenv("RM_B200_LOSS_ROUTE", route)
        assert size(long_, n) == expect, route
    monkeypatch.setenv("RM_B200_LOSS_ROUTE", "onfly4")
    g4, _ = geometry(n, block * 4)
    assert size(long_, n) == steps * 4 * g4 * block * 3 + (steps + 4) * nr * 4 * g4 * block
    assert size(long_, 0) == 0 and size(long_, -5) == 0
