"""Custom-op conformance (SURVEY.md §4 tier 4): ``torch.library.opcheck`` on every ``rm_b200::*`` op --
schema (no undeclared mutation / aliasing), fake-tensor kernels (``register_fake`` shapes, dtypes and
devices equal the real kernel's), autograd registration, and AOT dispatch on the differentiable ones.

Inputs are small enough for one thread block, so the float atomics of the backward kernels add in a
fixed order and repeated calls are bit-identical (opcheck compares runs).
"""
import pytest
import torch

from ray_marching_b200 import ops, scene as S
from ray_marching_b200.tape import tape_of

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
UTILS = ("test_schema", "test_autograd_registration", "test_faketensor", "test_aot_dispatch_dynamic")


@pytest.fixture(scope="module")
def tape_args():
    scene = S.make_test_scene().to(DEV)                  # smooth union of 4 affine children
    tape = tape_of(scene)
    with torch.no_grad():
        par = tape.lowered_params(DEV).clone()
    return tape, tape.code_tensor(), par


def _pts(n, seed=0, grad=False):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(n, 3, generator=g) * 0.5).to(DEV).requires_grad_(grad)


def test_opcheck_sdf_eval(tape_args):
    tape, code, par = tape_args
    for pts in (_pts(65), _pts(8).reshape(2, 4, 3)):
        torch.library.opcheck(torch.ops.rm_b200.sdf_eval.default, (code, par, pts, tape.n_nodes, tape.max_depth),
                              test_utils=UTILS)
    torch.library.opcheck(torch.ops.rm_b200.sdf_eval.default,
                          (code, par.clone().requires_grad_(True), _pts(33, grad=True), tape.n_nodes, tape.max_depth),
                          test_utils=UTILS)


def test_opcheck_sdf_vjp(tape_args):
    tape, code, par = tape_args
    pts = _pts(40)
    torch.library.opcheck(torch.ops.rm_b200.sdf_vjp.default,
                          (code, par, pts, torch.randn(40, 1, device=DEV), tape.n_nodes, tape.max_depth),
                          test_utils=UTILS[:3])


def test_opcheck_march_and_march_bwd(tape_args):
    tape, code, par = tape_args
    o = _pts(48, 1) + torch.tensor([0.0, 0.0, -1.5], device=DEV)
    v = torch.nn.functional.normalize(_pts(48, 2) * 0.2 + torch.tensor([0.0, 0.0, 1.0], device=DEV), dim=-1)
    torch.library.opcheck(torch.ops.rm_b200.march.default, (code, par, o, v, 8, False, tape.n_nodes, tape.max_depth),
                          test_utils=UTILS)
    torch.library.opcheck(torch.ops.rm_b200.march.default,
                          (code, par.clone().requires_grad_(True), o.clone().requires_grad_(True),
                           v.clone().requires_grad_(True), 8, False, tape.n_nodes, tape.max_depth),
                          test_utils=UTILS)
    torch.library.opcheck(torch.ops.rm_b200.march_bwd.default,
                          (code, par, o, v, torch.randn(48, 3, device=DEV), 8, tape.n_nodes, tape.max_depth, 1e-7),
                          test_utils=UTILS[:3])


def test_opcheck_normals(tape_args):
    tape, code, par = tape_args
    torch.library.opcheck(torch.ops.rm_b200.normals.default, (code, par, _pts(50), 5e-2, tape.n_nodes, tape.max_depth),
                          test_utils=UTILS[:3])


def test_opcheck_render_gbuffer_and_render_bwd(tape_args):
    tape, code, par = tape_args
    q = torch.nn.functional.normalize(torch.tensor([[0.99, 0.03, -0.08, 0.05]], device=DEV), dim=-1)
    t = torch.tensor([[0.0, 0.1, -1.5]], device=DEV)
    poses = ops.pose_records(q, t)
    W, H, px = 10, 6, 3.45e-6
    args = (code, par, poses, 1, W, H, px * H, px * W, px * H, 8, 5e-2, tape.n_nodes, tape.max_depth)
    torch.library.opcheck(torch.ops.rm_b200.render_gbuffer.default, args, test_utils=UTILS)
    args_g = (code, par.clone().requires_grad_(True), poses.clone().requires_grad_(True)) + args[3:]
    torch.library.opcheck(torch.ops.rm_b200.render_gbuffer.default, args_g, test_utils=UTILS)
    n = W * H
    g3 = lambda: torch.randn(n, 3, device=DEV)
    torch.library.opcheck(torch.ops.rm_b200.render_bwd.default,
                          (code, par, poses, 1, W, H, px * H, px * W, px * H, 8, 5e-2,
                           g3(), g3(), g3(), torch.randn(n, device=DEV), g3(), None,
                           tape.n_nodes, tape.max_depth, 1e-7),
                          test_utils=UTILS[:3])


def test_fake_kernels_give_the_real_shapes(tape_args):
    """register_fake outputs == real outputs in shape / dtype / device, for odd leading dimensions."""
    from torch._subclasses.fake_tensor import FakeTensorMode
    tape, code, par = tape_args
    pts = _pts(24).reshape(2, 3, 4, 3)
    real = torch.ops.rm_b200.sdf_eval(code, par, pts, tape.n_nodes, tape.max_depth)
    rn, rl = torch.ops.rm_b200.normals(code, par, pts, 5e-2, tape.n_nodes, tape.max_depth)
    with FakeTensorMode() as mode:
        fp, fpar = mode.from_tensor(pts), mode.from_tensor(par)
        fcode = mode.from_tensor(code)
        fake = torch.ops.rm_b200.sdf_eval(fcode, fpar, fp, tape.n_nodes, tape.max_depth)
        fn, fl = torch.ops.rm_b200.normals(fcode, fpar, fp, 5e-2, tape.n_nodes, tape.max_depth)
    for a, b in ((real, fake), (rn, fn), (rl, fl)):
        assert a.shape == b.shape and a.dtype == b.dtype and a.device == b.device


def test_opcheck_render_loss(tape_args):
    """rm_b200::render_loss (the fused loss + backward kernel): schema, fake kernel, autograd registration and
    AOT dispatch; one block of rays, so the float atomics add in a fixed order."""
    tape, code, par = tape_args
    q = torch.nn.functional.normalize(torch.tensor([[0.99, 0.03, -0.08, 0.05]], device=DEV), dim=-1)
    t = torch.tensor([[0.0, 0.1, -1.5]], device=DEV)
    poses = ops.pose_records(q, t)
    W, H, px = 10, 6, 3.45e-6
    n = W * H
    gen = torch.Generator().manual_seed(9)
    lam_t = torch.rand(n, generator=gen).to(DEV)
    dep_t = (1.0 + 0.2 * torch.rand(n, generator=gen)).to(DEV)
    common = (1, W, H, px * H, px * W, px * H, 8, 5e-2)
    tail = (1.0 / n, 1.0 / n, 0, n, tape.n_nodes, tape.max_depth, 1e-7)
    # no gradient requested: loss only
    torch.library.opcheck(torch.ops.rm_b200.render_loss.default,
                          (code, par, poses) + common + (lam_t, dep_t) + tail + (False, 0, 0), test_utils=UTILS[:3])
    # differentiable inputs, one target missing, an interleaved shard description
    torch.library.opcheck(torch.ops.rm_b200.render_loss.default,
                          (code, par.clone().requires_grad_(True), poses.clone().requires_grad_(True)) + common +
                          (lam_t, None) + tail + (True, 0, 0), test_utils=UTILS)
    real = torch.ops.rm_b200.render_loss(code, par, poses, *common, lam_t, dep_t, *tail, True, 0, 0)
    assert real[0].shape == () and real[1].shape == (2,) and real[2].shape == par.shape and real[3].shape == poses.shape
