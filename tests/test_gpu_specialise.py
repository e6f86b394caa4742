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
