"""rm_march_bwd_static: the allocation-free variant of rm_march_bwd (its trajectory scratch is a file-scope
device array inside the library; no counterpart in the reference).

Called through the C ABI (ctypes) and compared with the existing entry point rm_march_bwd -- same launch
geometry, so the two differ only in the order of the float atomics that sum d loss / d par -- within the
gradient tolerance of the other tests (1e-3 of the largest component of each tensor), on TWO CALLS IN A ROW
with different inputs (the shared scratch keeps the previous call's contents), plus the size limit.
This file sorts last on purpose."""
import ctypes as C

import pytest
import torch

from ray_marching_b200 import _lib, scene as S
from ray_marching_b200.tape import tape_of

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _rel(a, b):
    return (a - b).abs().max().item() / (b.abs().max().item() + 1e-30)


def _case(seed, n, z0):
    g = torch.Generator().manual_seed(seed)
    o = torch.tensor([0.0, 0.0, z0]) + 0.02 * torch.randn(n, 3, generator=g)
    v = torch.nn.functional.normalize(torch.randn(n, 3, generator=g) * 0.3 + torch.tensor([0.0, 0.0, 1.0]), dim=-1)
    w = torch.randn(n, 3, generator=g)
    return o.to(DEV).contiguous(), v.to(DEV).contiguous(), w.to(DEV).contiguous()


def _both(lib, tape_struct, n_par, o, v, w, steps, tol=1e-7):
    n = o.shape[0]
    st = torch.cuda.current_stream().cuda_stream
    p = lambda t: C.c_void_p(t.data_ptr())
    # existing entry point: caller-provided scratch
    traj = torch.empty(max(int(lib.rm_bwd_scratch_floats(n, steps)), 1), device=DEV)
    do_a, dv_a, dp_a = torch.empty_like(o), torch.empty_like(v), torch.empty(n_par, device=DEV)
    _lib.check(lib.rm_march_bwd(C.byref(tape_struct), p(o), p(v), p(w), n, steps, tol, p(traj), p(do_a), p(dv_a),
                                p(dp_a), st), "rm_march_bwd")
    # new entry point: outputs pre-allocated here, NOTHING allocated inside the call
    do_b, dv_b = torch.full_like(o, float("nan")), torch.full_like(v, float("nan"))
    dp_b = torch.full((n_par,), float("nan"), device=DEV)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    _lib.check(lib.rm_march_bwd_static(C.byref(tape_struct), p(o), p(v), p(w), n, steps, tol, p(do_b), p(dv_b),
                                       p(dp_b), st), "rm_march_bwd_static")
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    return (do_a, dv_a, dp_a), (do_b, dv_b, dp_b), free_before - free_after


def test_march_bwd_static_matches_march_bwd_on_two_calls_in_a_row():
    lib = _lib.lib()
    scene = S.make_test_scene2().to(DEV)
    tape = tape_of(scene)
    with torch.no_grad():
        par = tape.lowered_params(torch.device(DEV)).clone()
    code = tape.code_tensor()
    ts = _lib.RmTape(code.data_ptr(), par.data_ptr(), tape.n_instr, tape.n_par, tape.max_depth, tape.n_nodes)
    # set-up call (fetches the array's device address once per device): n_rays = 0 is legal and launches nothing
    st = torch.cuda.current_stream().cuda_stream
    dp0 = torch.ones(tape.n_par, device=DEV)
    assert lib.rm_march_bwd_static(C.byref(ts), None, None, None, 0, 32, 1e-7, None, None,
                                   C.c_void_p(dp0.data_ptr()), st) == 0
    torch.cuda.synchronize()
    assert not dp0.any()                                   # outputs are cleared even for an empty call
    drops = []
    # two calls in a row, different ray counts / step counts / rays: the second must not see the first's scratch
    for seed, n, steps in ((3, 5000, 32), (4, 3001, 17)):
        o, v, w = _case(seed, n, 1.0)
        (do_a, dv_a, dp_a), (do_b, dv_b, dp_b), drop = _both(lib, ts, tape.n_par, o, v, w, steps)
        drops.append(drop)
        for a, b, what in ((do_a, do_b, "d origin"), (dv_a, dv_b, "d dirs"), (dp_a, dp_b, "d par")):
            assert torch.isfinite(b).all(), what
            assert _rel(b, a) < 1e-3, (what, seed, _rel(b, a))
        # per-ray outputs do not depend on the order of atomics: identical
        assert torch.equal(do_a, do_b) and torch.equal(dv_a, dv_b)
        assert dp_a.abs().max().item() > 0
    # no device memory was taken by the calls (a cudaMalloc of the scratch would be 32 MiB)
    assert all(d < (1 << 20) for d in drops), drops


def test_march_bwd_static_rejects_sizes_beyond_its_scratch():
    lib = _lib.lib()
    scene = S.make_test_scene2().to(DEV)
    tape = tape_of(scene)
    with torch.no_grad():
        par = tape.lowered_params(torch.device(DEV)).clone()
    code = tape.code_tensor()
    ts = _lib.RmTape(code.data_ptr(), par.data_ptr(), tape.n_instr, tape.n_par, tape.max_depth, tape.n_nodes)
    n, steps = 200_000, 32
    assert int(lib.rm_bwd_scratch_floats(n, steps)) > _lib.MARCH_BWD_STATIC_MAX_FLOATS
    o, v, w = _case(5, 8, 1.0)                              # never dereferenced: the size check comes first
    dp = torch.empty(tape.n_par, device=DEV)
    p = lambda t: C.c_void_p(t.data_ptr())
    rc = lib.rm_march_bwd_static(C.byref(ts), p(o), p(v), p(w), n, steps, 1e-7, p(o), p(v), p(dp),
                                 torch.cuda.current_stream().cuda_stream)
    msg = lib.rm_last_error().decode()
    assert rc == -3 and "RM_MARCH_BWD_STATIC_MAX_FLOATS" in msg and "rm_march_bwd" in msg, (rc, msg)
    # the largest size that fits at 32 steps is accepted by the size check (682 blocks x 128 rays)
    assert int(lib.rm_bwd_scratch_floats(682 * 128, 32)) <= _lib.MARCH_BWD_STATIC_MAX_FLOATS
    torch.cuda.synchronize()
