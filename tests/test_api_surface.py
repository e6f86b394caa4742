"""The drop-in boundary, pinned: constructor / ``forward`` parameter names and
defaults of every reference class on the path, the free functions' signatures, and the
``state_dict`` keys / shapes / dtypes of a ``RenderLoop`` -- against ``tests/golden/api_surface.json``,
generated from the UNMODIFIED reference by ``oracle/make_golden.py`` (``gen_api_surface``).

Also the host-side drop-in functions (``quaternion.*``, ``control.make_reflection_directions``;
device-agnostic torch glue) against ``tests/golden/quaternion.npz``.  CPU only: no kernel is called.
"""
import inspect
import json
import os

import numpy as np
import pytest
import torch

import ray_marching_b200 as R
from ray_marching_b200 import quaternion as Q
from ray_marching_b200 import scene as S
from rm_testutil import t32

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def api():
    with open(os.path.join(GOLDEN, "api_surface.json")) as f:
        return json.load(f)


def _mine(name):
    if hasattr(S, name):
        return getattr(S, name)
    if name == "QuaternionToSO3":
        return Q.QuaternionToSO3
    return getattr(R, name)


def _defaults(fn):
    return {k: v.default for k, v in inspect.signature(fn).parameters.items()
            if v.default is not inspect.Parameter.empty and isinstance(v.default, (int, float, str, bool))}


def test_every_reference_class_keeps_its_signatures(api):
    assert len(api["classes"]) >= 17
    for name, want in api["classes"].items():
        cls = _mine(name)
        init = list(inspect.signature(cls.__init__).parameters)
        fwd = list(inspect.signature(cls.forward).parameters)
        assert init == want["init"], f"{name}.__init__: {init} != reference {want['init']}"
        # extra arguments are allowed only after the reference's, and only keyword-only
        assert fwd[:len(want["forward"])] == want["forward"], \
            f"{name}.forward: {fwd} != reference {want['forward']}"
        for extra in fwd[len(want["forward"]):]:
            kind = inspect.signature(cls.forward).parameters[extra].kind
            assert kind is inspect.Parameter.KEYWORD_ONLY, f"{name}.forward: extra positional '{extra}'"
        assert _defaults(cls.__init__) == want["init_defaults"], name
        mine_fd = _defaults(cls.forward)
        assert {k: mine_fd.get(k) for k in want["forward_defaults"]} == want["forward_defaults"], name


def test_free_functions_keep_their_signatures(api):
    table = {"quaternion.cross_product": Q.cross_product, "quaternion.multiply": Q.multiply,
             "quaternion.conjugate": Q.conjugate, "quaternion.rotation": Q.rotation,
             "quaternion.to_versor": Q.to_versor,
             "control.make_reflection_directions": R.make_reflection_directions,
             "scene_registry.make_test_scene": S.make_test_scene,
             "scene_registry.make_test_scene2": S.make_test_scene2}
    assert set(table) == set(api["functions"])
    for name, fn in table.items():
        got = [k for k, v in inspect.signature(fn).parameters.items()
               if v.default is inspect.Parameter.empty]          # ours may add defaulted extras
        assert got == api["functions"][name], f"{name}: {got} != reference {api['functions'][name]}"


@pytest.mark.parametrize("tag,make,n", [("scene2_n1_8x6", S.make_test_scene2, 1),
                                        ("scene1_n2_8x6", S.make_test_scene, 2)])
def test_state_dict_matches_the_reference(api, tag, make, n):
    loop = R.RenderLoop(make(), n, 8, 6, 1e-3, 2e-3, 1.5e-3, 5e-2)
    got = [[k, list(v.shape), str(v.dtype)] for k, v in loop.state_dict().items()]
    assert got == api["state_dict"][tag]
    # a checkpoint written by either side loads strictly into this module
    sd = loop.state_dict()
    fresh = R.RenderLoop(make(), n, 8, 6, 1e-3, 2e-3, 1.5e-3, 5e-2)
    with torch.no_grad():
        for p in fresh.parameters():
            p.add_(1.0)
    res = fresh.load_state_dict(sd, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    for (ka, a), (kb, b) in zip(loop.state_dict().items(), fresh.state_dict().items()):
        assert ka == kb and torch.equal(a, b), ka
    bad = dict(sd)
    bad.pop("camera.ray_directions")
    with pytest.raises(RuntimeError, match="ray_directions"):
        fresh.load_state_dict(bad, strict=True)
    bad = dict(sd)
    bad["camera.ray_positions"] = torch.zeros(1, 2, 2, 3)
    with pytest.raises(RuntimeError, match="ray_positions"):
        fresh.load_state_dict(bad, strict=True)


def test_camera_state_dict_rays_equal_the_reference_buffers(golden):
    g = golden("camera")
    cam = R.PinholeCamera(2, 8, 6, 17e-3, 24e-3, 18e-3)
    sd = cam.state_dict()
    np.testing.assert_allclose(sd["ray_positions"].numpy(), g["camera/base_pos"], rtol=0, atol=1e-9)
    np.testing.assert_allclose(sd["ray_directions"].numpy(), g["camera/base_dirs"], rtol=0, atol=2e-7)


def test_combinators_accept_the_reference_keyword():
    """transformations.py:67,90,117,131 name the argument ``query_coords``; primitives and the affine
    transform ``query_positions``.  Binding by keyword must not raise TypeError before the (CUDA-only)
    evaluation rejects the CPU tensor."""
    from ray_marching_b200._lib import RmError
    pts = torch.zeros(2, 3)
    box = S.SDFBox((0.1, 0.2, 0.3))
    for node, kw in ((S.SDFUnion([box]), "query_coords"), (S.SDFSmoothUnion([box], 4.0), "query_coords"),
                     (S.SDFRounding(box, 0.1), "query_coords"), (S.SDFOnion(box, 0.1), "query_coords"),
                     (box, "query_positions"),
                     (S.SDFAffineTransformation(box, (1.0, 0.0, 0.0, 0.0), (0.0, 0.0, 0.0)), "query_positions")):
        with pytest.raises(RmError):
            node(**{kw: pts})


# ---- host-side drop-in functions vs the reference fixture -------------------------------------
def test_quaternion_functions_match_reference_fixture(golden):
    g = golden("quaternion")
    p, q, v, small = (t32(g[f"quat/{k}"]) for k in ("p", "q", "v", "small"))
    tol = dict(rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(Q.cross_product(v, small).numpy(), g["quat/cross"], **tol)
    np.testing.assert_allclose(Q.multiply(p, q).numpy(), g["quat/multiply"], **tol)
    np.testing.assert_allclose(Q.conjugate(q).numpy(), g["quat/conjugate"], rtol=0, atol=0)
    np.testing.assert_allclose(Q.rotation(v, q).numpy(), g["quat/rotation"], rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(Q.to_versor(small).numpy(), g["quat/to_versor"], **tol)
    np.testing.assert_allclose(Q.QuaternionToSO3()(q[:, None, None, :]).numpy(), g["quat/so3"], **tol)
    n = torch.nn.functional.normalize(small, dim=-1)
    np.testing.assert_allclose(R.make_reflection_directions(n, v).numpy(), g["quat/reflect"], **tol)


def test_quaternion_edge_cases():
    # to_versor: NaN when |V| > 1, as the reference (quaternion.py:75-91)
    assert torch.isnan(Q.to_versor(torch.tensor([[0.9, 0.9, 0.0]]))[0, 0])
    # broadcasting of rotation (camera: (N,H,W,3) against (N,1,1,4)) and of cross_product
    V = torch.randn(2, 3, 4, 3)
    q = torch.randn(2, 1, 1, 4)
    out = Q.rotation(V, q)
    assert out.shape == V.shape
    ref = V + 2 * q[..., :1] * torch.linalg.cross(q[..., 1:].expand_as(V), V) + \
        2 * torch.linalg.cross(q[..., 1:].expand_as(V), torch.linalg.cross(q[..., 1:].expand_as(V), V))
    np.testing.assert_allclose(out.numpy(), ref.numpy(), rtol=1e-5, atol=1e-5)
    assert Q.cross_product(V, torch.randn(3)).shape == V.shape
    # gradients of rotation w.r.t. a NON-unit quaternion equal those of the reference polynomial
    qq = torch.tensor([[0.9014, 0.25, 0.25, 0.25]], requires_grad=True)
    vv = torch.tensor([[0.3, -0.2, 0.7]])
    Q.rotation(vv, qq).sum().backward()
    q2 = qq.detach().clone().requires_grad_(True)
    w, u = q2[..., :1], q2[..., 1:]
    tt = 2 * torch.linalg.cross(u, vv)
    (vv + w * tt + torch.linalg.cross(u, tt)).sum().backward()
    np.testing.assert_allclose(qq.grad.numpy(), q2.grad.numpy(), rtol=1e-5, atol=1e-6)


def test_fast_math_levels():
    """RenderLoop.fast_math / .cull -> RmRenderArgs.fast_math (include/rm_b200.h): culling only exists on the
    approximate-math kernels; "exact" is the bit-identical subset."""
    from ray_marching_b200 import ops
    assert ops.fast_math_level(False, False) == 0 and ops.fast_math_level(False, True) == 0
    assert ops.fast_math_level(False, "exact") == 0
    assert ops.fast_math_level(True, False) == 1 and ops.fast_math_level(True, True) == 2
    assert ops.fast_math_level(True, "exact") == 3
    with pytest.raises(ValueError):
        ops.fast_math_level(True, "fuzzy")
