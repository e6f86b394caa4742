"""Tape-specialised render kernels (specialise.py) against the interpreter and the oracle."""
import numpy as np
import pytest
import torch

from oracle import sdf_oracle as O
from ray_marching_b200 import RenderLoop, _lib
from ray_marching_b200.scene.scene_registry import make_test_scene, make_test_scene2
from test_specialise_cpu import nested_scene

pytestmark = pytest.mark.gpu
PX = 3.45e-6


def make_loop(scene, w=192, h=108):
    return RenderLoop(scene, num_cameras=1, px_width=w, px_height=h, focal_length=PX * h,
                      sensor_width=PX * w, sensor_height=PX * h, normals_eps=5e-2).cuda()


@pytest.mark.parametrize("build,t", [(make_test_scene2, (0.0, 0.0, 1.0)), (make_test_scene, (0.0, 0.0, -1.5)),
                                     (nested_scene, (0.0, 0.0, -1.5))])
@pytest.mark.parametrize("fast", [True, False])
def test_specialised_render_equals_interpreter(build, t, fast):
    loop = make_loop(build())
    loop.fast_math = fast
    q = torch.nn.functional.normalize(torch.tensor([[0.98, 0.05, -0.1, 0.02]], device="cuda"), dim=-1)
    tt = torch.tensor([t], device="cuda")
    ref = loop.render_planes(q, tt, gbuffer=True, modes=(0, 1, 3, 4, 5))
    ref_planes = {k: getattr(ref, k).clone() for k in ("pos", "sdf", "normals", "laplacian")}
    ref_imgs = [None if x is None else x.clone() for x in ref.images]
    loop.specialise()
    got = loop.render_planes(q, tt, gbuffer=True, modes=(0, 1, 3, 4, 5))
    # same statements in the same order: agreement to rounding of re-associated FMAs at most
    settled = ref_planes["sdf"].abs() < 1e-4
    assert settled.float().mean() > 0.1
    for k, r in ref_planes.items():
        g = getattr(got, k)
        m = settled if r.dim() == 1 else settled[:, None].expand_as(r)
        scale = 1.0 + r[m].abs()
        tol = 2e-3 if k == "laplacian" else 1e-5          # the Laplacian amplifies rounding by 6/eps^2
        assert ((g[m] - r[m]).abs() / scale).max().item() < tol, k
        assert torch.equal(torch.isnan(g), torch.isnan(r)), k
    for m_, r in enumerate(ref_imgs):
        if r is not None:
            ok = settled if r.shape[1] == 1 else settled
            d = (got.images[m_] - r).abs()[ok]
            assert torch.nan_to_num(d, nan=0.0).max().item() < 1e-3, m_
    loop.specialise(False)
    back = loop.render_planes(q, tt, gbuffer=True, modes=(0,))
    assert torch.equal(torch.nan_to_num(back.pos), torch.nan_to_num(ref_planes["pos"]))


def test_specialised_render_matches_oracle_depth():
    from oracle import cases
    W, H = 160, 120
    intr = cases.config1_camera(W, H)
    loop = RenderLoop(make_test_scene2(), normals_eps=5e-2, **intr).cuda().specialise()
    q = torch.tensor([[1.0, 0.0, 0.0, 0.0]], device="cuda")
    t = torch.tensor([[0.0, 0.0, 1.0]], device="cuda")
    res = loop.render_planes(q, t, gbuffer=True, modes=(0,))
    orc = O.RenderLoopOracle(make_test_scene2(), normals_eps=5e-2, **intr)
    with torch.no_grad():
        gb = orc.gbuffer(q.cpu(), t.cpu(), 32)
        lam = O.shade_lambertian(gb["dirs"], gb["normals"]).reshape(-1)
    depth = (res.pos - res.origin).norm(dim=-1).cpu()
    depth_ref = (gb["pos"] - gb["px"]).norm(dim=-1).reshape(-1)
    assert ((depth - depth_ref).abs() / depth_ref).max().item() < 1e-4
    assert (res.images[0].reshape(-1).cpu() - lam).abs().max().item() < 1e-3


def test_specialised_library_refuses_another_tape():
    loop_a = make_loop(make_test_scene2()).specialise()
    loop_b = make_loop(make_test_scene())
    from ray_marching_b200.tape import tape_of
    tape_of(loop_b.scene).spec_lib = tape_of(loop_a.scene).spec_lib      # wrong library on purpose
    q = torch.tensor([[1.0, 0.0, 0.0, 0.0]], device="cuda")
    t = torch.tensor([[0.0, 0.0, -1.5]], device="cuda")
    with pytest.raises(_lib.RmError, match="specialised for another tape"):
        loop_b.render_planes(q, t, modes=(0,))


# ---- backward: the compiler-specialised kernels (unrolled tape loops) -------------------------
def _every_plane_loss(gb, weights):
    return sum((gb[k] * w).sum() for k, w in weights.items())


def _grads_through(loop, scene, q0, t0, weights):
    for p in scene.parameters():
        p.grad = None
    q, t = q0.clone().requires_grad_(True), t0.clone().requires_grad_(True)
    _every_plane_loss(loop.gbuffer(q, t, 32), weights).backward()
    flat = torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1)
                      for p in scene.parameters()])
    return q.grad.clone(), t.grad.clone(), flat.clone()


def _rel(a, b):
    return ((a - b).abs().max() / (b.abs().max() + 1e-30)).item()


def _in_shell(build):
    """The scene2 pattern (scene_registry.py:55-60): an enclosing onion box, so that every ray hits
    and no NaN normals enter the gradients."""
    from ray_marching_b200 import scene as S
    return lambda: S.SDFUnion([S.SDFOnion(S.SDFBox((5.0, 5.0, 5.0)), 0.1), build()])


@pytest.mark.parametrize("build,t", [(make_test_scene2, (0.05, 0.1, 1.0)),
                                     (_in_shell(make_test_scene), (0.1, 0.0, -1.5)),
                                     (_in_shell(nested_scene), (0.0, 0.1, -1.5))])
def test_specialised_backward_equals_interpreter(build, t):
    """Random upstream gradients on every G-buffer plane, two views, odd image size: the specialised
    forward (IEEE) + backward kernels against the interpreter's -- same statements, same order."""
    scene = build()
    loop = RenderLoop(scene, num_cameras=2, px_width=37, px_height=21, focal_length=PX * 21,
                      sensor_width=PX * 37, sensor_height=PX * 21, normals_eps=5e-2).cuda()
    q0 = torch.nn.functional.normalize(torch.tensor([[0.99, 0.03, -0.08, 0.05], [0.97, -0.1, 0.12, 0.02]],
                                                    device="cuda"), dim=-1)
    t0 = torch.tensor([list(t), [t[0] - 0.1, t[1], t[2] * 0.95]], device="cuda")
    gen = torch.Generator().manual_seed(5)
    with torch.no_grad():
        probe = loop.gbuffer(q0, t0, 32)
    assert all(torch.isfinite(v).all() for v in probe.values())
    weights = {k: torch.randn(v.shape, generator=gen).cuda() * (1e-3 if k == "laplacian" else 1.0)
               for k, v in probe.items()}
    ref = _grads_through(loop, scene, q0, t0, weights)
    loop.specialise()
    got = _grads_through(loop, scene, q0, t0, weights)
    loop.specialise(False)
    again = _grads_through(loop, scene, q0, t0, weights)
    for g, r, name in zip(got, ref, ("dq", "dt", "dparams")):
        assert torch.isfinite(r).all(), name
        assert _rel(g, r) < 2e-4, name
    for g, r in zip(again, ref):                   # de-selecting returns to the interpreter
        assert _rel(g, r) < 1e-5


def test_specialised_backward_matches_reference_fixture(golden):
    """Config-4 loss through the specialised fused path against the reference-autograd fixture."""
    from oracle import cases
    from rm_testutil import t32
    g = golden("grads")
    tag = "scene2"
    scene = make_test_scene2()
    loop = RenderLoop(scene, normals_eps=5e-2, **cases.config1_camera(16, 16)).cuda().specialise()
    q = t32(g[f"grads/{tag}/q"]).cuda().requires_grad_(True)
    t = t32(g[f"grads/{tag}/t"]).cuda().requires_grad_(True)
    gb = loop.gbuffer(q, t, 32)
    lam = (-(gb["dirs"] * gb["normals"]).sum(-1, keepdim=True)).clamp(0, 1)
    depth = (gb["pos"] - gb["origin"]).norm(dim=-1, keepdim=True)
    loss = ((lam - t32(g[f"grads/{tag}/lam_target"]).cuda()) ** 2).mean() + \
           ((depth - t32(g[f"grads/{tag}/depth_target"]).cuda()) ** 2).mean()
    assert abs(loss.item() - float(g[f"grads/{tag}/loss"])) < 1e-4 * max(1.0, float(g[f"grads/{tag}/loss"]))
    loss.backward()
    flat = torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1)
                      for p in scene.parameters()]).cpu()
    rel = lambda a, b: (np.abs(a - b).max() / (np.abs(b).max() + 1e-30))
    assert rel(q.grad.cpu().numpy(), g[f"grads/{tag}/dq"]) < 1e-3
    assert rel(t.grad.cpu().numpy(), g[f"grads/{tag}/dt"]) < 1e-3
    assert rel(flat.numpy(), g[f"grads/{tag}/dparams_flat"]) < 1e-3


def test_specialised_march_and_sdf_vjp_follow_the_selection():
    """SDFMarcher / scene-node autograd (rm_march_bwd, rm_sdf_vjp) through the specialised library."""
    from ray_marching_b200 import SDFMarcher
    scene = make_test_scene2()
    loop = make_loop(scene, 24, 16)
    gen = torch.Generator().manual_seed(3)
    pts = (torch.rand(500, 3, generator=gen) * 2 - 1).cuda()
    w = torch.randn(500, 1, generator=gen).cuda()
    o = torch.tensor([[0.0, 0.0, 1.0]]).repeat(200, 1).cuda() + 0.01 * torch.randn(200, 3, generator=gen).cuda()
    v = torch.nn.functional.normalize(torch.randn(200, 3, generator=gen).cuda() * 0.2 + torch.tensor([0.0, 0.0, 1.0]).cuda(), dim=-1)
    marcher = SDFMarcher(scene).cuda()

    def run():
        for p in scene.parameters():
            p.grad = None
        x = pts.clone().requires_grad_(True)
        (scene(x) * w).sum().backward()
        g_sdf = torch.cat([p.grad.reshape(-1) for p in scene.parameters()]).clone()
        for p in scene.parameters():
            p.grad = None
        oo, vv = o.clone().requires_grad_(True), v.clone().requires_grad_(True)
        (marcher(oo, vv, 32) ** 2).sum().backward()
        g_m = torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in scene.parameters()]).clone()
        return x.grad.clone(), g_sdf, oo.grad.clone(), vv.grad.clone(), g_m

    ref = run()
    loop.specialise()
    got = run()
    for g, r in zip(got, ref):
        assert _rel(g, r) < 2e-4


def test_specialised_backward_refuses_another_tape():
    """The backward entry points of a specialised library check the structure hash too."""
    import ctypes as C
    from ray_marching_b200 import ops, specialise as SP
    from ray_marching_b200.tape import tape_of
    loop_a = make_loop(make_test_scene2(), 16, 16).specialise()
    handle = tape_of(loop_a.scene).spec_lib
    other = tape_of(make_loop(make_test_scene(), 16, 16).scene)
    with torch.no_grad():
        par = other.lowered_params()
    t = ops._tape_struct(other.code_tensor(), par, other.n_nodes, other.max_depth)
    pts = torch.zeros(8, 3, device="cuda")
    g = torch.ones(8, device="cuda")
    dpts, dpar = torch.empty_like(pts), torch.empty_like(par)
    st = handle.rm_sdf_vjp(C.byref(t), pts.data_ptr(), g.data_ptr(), 8, dpts.data_ptr(), dpar.data_ptr(),
                           torch.cuda.current_stream().cuda_stream)
    assert st != 0 and b"specialised for another tape" in handle.rm_last_error()
    with pytest.raises(_lib.RmError, match="specialised"):
        SP.spec_check(handle, st, "rm_sdf_vjp")


def test_specialised_early_exit_is_bit_identical():
    """The specialised kernel has two march loops (with and without the fixed-point test): same planes."""
    loop = make_loop(make_test_scene2(), 128, 72).specialise()
    q = torch.nn.functional.normalize(torch.tensor([[0.98, 0.05, -0.1, 0.02]], device="cuda"), dim=-1)
    t = torch.tensor([[0.0, 0.0, 1.0]], device="cuda")
    full = loop.render_planes(q, t, gbuffer=True, modes=(0, 4, 5))
    keep = [x.clone() for x in (full.pos, full.sdf, full.normals, full.laplacian, full.images[0], full.images[5])]
    loop.early_exit = True
    fast = loop.render_planes(q, t, gbuffer=True, modes=(0, 4, 5))
    for a, b in zip(keep, (fast.pos, fast.sdf, fast.normals, fast.laplacian, fast.images[0], fast.images[5])):
        assert torch.equal(torch.nan_to_num(a, nan=-7.0), torch.nan_to_num(b, nan=-7.0))
