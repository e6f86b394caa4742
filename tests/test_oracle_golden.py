"""Pins the CPU oracle (oracle/sdf_oracle.py) to outputs of the unmodified reference.

The fixtures in tests/golden/ were produced by oracle/make_golden.py from /root/reference.
Values are compared bit-for-bit where the oracle uses the same torch operators in the same
order (everything except explicitly noted), NaN masks must coincide, gradients to 1e-6.
"""
import numpy as np
import pytest
import torch

from oracle import cases, sdf_oracle as O
from ray_marching_b200 import scene as S
from ray_marching_b200.scene.scene_registry import _default_classes
from rm_testutil import t32

C = _default_classes()


def same(a, b, tol=0.0):
    a = np.asarray(a.detach() if isinstance(a, torch.Tensor) else a)
    b = np.asarray(b)
    assert a.shape == b.shape, (a.shape, b.shape)
    assert np.array_equal(np.isnan(a), np.isnan(b))
    if tol == 0.0:
        assert np.array_equal(np.nan_to_num(a), np.nan_to_num(b))
    else:
        np.testing.assert_allclose(np.nan_to_num(a), np.nan_to_num(b), rtol=tol, atol=tol)


def test_quaternion(golden):
    g = golden("quaternion")
    p, q, v, small = (t32(g[f"quat/{k}"]) for k in ("p", "q", "v", "small"))
    same(O.q_cross(v, small), g["quat/cross"])
    same(O.q_multiply(p, q), g["quat/multiply"])
    same(O.q_conjugate(q), g["quat/conjugate"])
    same(O.q_rotation(v, q), g["quat/rotation"])
    same(O.q_to_versor(small), g["quat/to_versor"])
    same(O.q_to_so3(q[:, None, None, :]), g["quat/so3"])
    same(O.reflect(torch.nn.functional.normalize(small, dim=-1), v), g["quat/reflect"])


def test_known_answers():
    """Hand-checkable vectors of SURVEY.md §8(c)."""
    pts = torch.tensor([[0.3, -0.2, 0.7], [0.0, 0.0, 0.0], [-1.5, 2.0, 0.25]])
    want = {
        "sphere": [0.28740078, -0.5, 2.01246881],
        "box": [0.68007350, -0.05, 2.28910446],
        "plane": [0.3, 0.0, -1.5],
        "line": [1.10000002, 1.12474489, 1.77916205],
        "disk": [0.37681428, 0.0, 2.13235450],
        "torus": [0.22927600, 0.40000001, 2.14539742],
        "rounding_box": [0.63007349, -0.1, 2.23910451],
        "onion_box": [0.58007348, -0.05, 2.18910456],
        "affine_torus": [0.29802340, 0.51803386, 1.80668497],
    }
    nodes = cases.node_cases(C)
    for name, vals in want.items():
        got = O.sdf(nodes[name], pts)[:, 0]
        np.testing.assert_allclose(got.detach().numpy(), vals, rtol=2e-6, atol=2e-7)
    np.testing.assert_allclose(O.sdf(S.make_test_scene(), pts)[:, 0].detach().numpy(),
                               [-0.03099138, 0.04288458, 1.75918460], rtol=2e-6, atol=2e-7)
    np.testing.assert_allclose(O.sdf(S.make_test_scene2(), pts)[:, 0].detach().numpy(),
                               [0.06119990, -0.5, 1.81666851], rtol=2e-6, atol=2e-7)


@pytest.mark.parametrize("name", list(cases.node_cases(C).keys()))
def test_nodes_value_and_grad(golden, name):
    g = golden("nodes")
    node = cases.node_cases(C)[name]
    x = t32(g["nodes/points"]).requires_grad_(True)
    w = t32(g["nodes/weights"])
    d = O.sdf(node, x)
    same(d, g[f"nodes/{name}/value"])
    params = list(node.named_parameters())
    grads = torch.autograd.grad((d * w).sum(), [x] + [p for _, p in params], allow_unused=True)
    same(grads[0], g[f"nodes/{name}/dpoints"], 1e-6)
    for (pn, p), gp in zip(params, grads[1:]):
        same(gp if gp is not None else torch.zeros_like(p), g[f"nodes/{name}/dparam/{pn}"], 1e-6)


@pytest.mark.parametrize("name,build", [
    ("test_scene", S.make_test_scene), ("test_scene2", S.make_test_scene2),
    ("synthetic8_shell", lambda: S.make_synthetic_scene(8, 0, True)),
    ("synthetic64", lambda: S.make_synthetic_scene(64, 0, False)),
])
def test_scenes(golden, name, build):
    g = golden("nodes")
    scene = build()
    flat = torch.cat([p.reshape(-1) for p in scene.parameters()])
    same(flat, g[f"scenes/{name}/params_flat"])           # same parameters, same order
    x = t32(g["nodes/points"]).requires_grad_(True)
    w = t32(g["nodes/weights"])
    d = O.sdf(scene, x)
    same(d, g[f"scenes/{name}/value"])
    params = list(scene.parameters())
    grads = torch.autograd.grad((d * w).sum(), [x] + params, allow_unused=True)
    same(grads[0], g[f"scenes/{name}/dpoints"], 1e-6)
    gflat = torch.cat([(gp if gp is not None else torch.zeros_like(p)).reshape(-1)
                       for p, gp in zip(params, grads[1:])])
    same(gflat, g[f"scenes/{name}/dparams_flat"], 1e-5)


def test_camera(golden):
    g = golden("camera")
    base = O.camera_base_rays(2, 8, 6, 17e-3, 24e-3, 18e-3)
    same(base[0], g["camera/base_pos"])
    same(base[1], g["camera/base_dirs"])
    pos, frames, pos2, dirs = O.camera_rays(*base, t32(g["camera/q"]), t32(g["camera/t"]))
    assert pos is pos2
    same(pos, g["camera/pos"])
    same(dirs, g["camera/dirs"])
    same(frames, g["camera/frames"])


def test_config1_probe_pixels(golden):
    """640x480 config-1 camera; three probe pixels of SURVEY.md §8(c) through march + stencil."""
    g = golden("camera")
    base = O.camera_base_rays(**cases.config1_camera(640, 480))
    pos0, _, _, dirs = O.camera_rays(*base, torch.tensor([[1.0, 0, 0, 0]]),
                                     torch.tensor([[0.0, 0, 1.0]]))
    sel = ([240, 0, 100], [320, 0, 500])
    p0, v = pos0[0][sel], dirs[0][sel]
    same(p0, g["camera/c1_pos_sel"])
    same(v, g["camera/c1_dirs_sel"])
    scene = S.make_test_scene2()
    with torch.no_grad():
        hit = O.march(scene, p0, v, 32)
        n, lap = O.normals_laplacian(scene, hit, 5e-2)
        same(hit, g["camera/c1_hit_sel"])
        same(O.sdf(scene, hit), g["camera/c1_sdf_sel"])
        same(n, g["camera/c1_normals_sel"])
        same(lap, g["camera/c1_lap_sel"])
    np.testing.assert_allclose((hit - p0).norm(dim=-1).numpy(),
                               [0.25000021, 0.30783942, 0.27159759], rtol=1e-6)


@pytest.mark.parametrize("tag,build", [
    ("scene2", S.make_test_scene2), ("scene1", S.make_test_scene),
    ("syn8", lambda: S.make_synthetic_scene(8, 0, True)),
])
def test_render_loop(golden, tag, build):
    g = golden("render")
    from ray_marching_b200.rendering.shader import load_cmap
    loop = O.RenderLoopOracle(build(), normals_eps=5e-2, cmap=load_cmap(),
                              **cases.config1_camera(32, 24))
    q, t = t32(g[f"render/{tag}/q"]), t32(g[f"render/{tag}/t"])
    with torch.no_grad():
        gb = loop.gbuffer(q, t, 32)
        same(gb["px"], g[f"render/{tag}/pos0"])
        same(gb["dirs"], g[f"render/{tag}/dirs"])
        same(gb["pos"], g[f"render/{tag}/pos"])
        same(gb["dists"], g[f"render/{tag}/sdf"])
        same(gb["normals"], g[f"render/{tag}/normals"])
        same(gb["lap"], g[f"render/{tag}/lap"])
        for mode in range(8):
            img = loop(q, t, mode, 3 if mode >= 6 else 1, 32)
            same(img.contiguous(), g[f"render/{tag}/mode{mode}"])
        same(loop(q, t, 0, 1, 7).contiguous(), g[f"render/{tag}/steps7_mode0"])


@pytest.mark.parametrize("tag,build", [
    ("scene2", S.make_test_scene2), ("syn8", lambda: S.make_synthetic_scene(8, 0, True)),
])
def test_config4_gradients(golden, tag, build):
    g = golden("grads")
    scene = build()
    loop = O.RenderLoopOracle(scene, normals_eps=5e-2, **cases.config1_camera(16, 16))
    q = t32(g[f"grads/{tag}/q"]).requires_grad_(True)
    t = t32(g[f"grads/{tag}/t"]).requires_grad_(True)
    gb = loop.gbuffer(q, t, 32)
    target = dict(lam=t32(g[f"grads/{tag}/lam_target"]), depth=t32(g[f"grads/{tag}/depth_target"]))
    loss = O.config4_loss(gb, target)
    same(loss, g[f"grads/{tag}/loss"], 1e-6)
    params = list(scene.parameters())
    grads = torch.autograd.grad(loss, [q, t] + params, allow_unused=True, retain_graph=True)
    same(grads[0], g[f"grads/{tag}/dq"], 2e-5)
    same(grads[1], g[f"grads/{tag}/dt"], 2e-5)
    flat = torch.cat([(gp if gp is not None else torch.zeros_like(p)).reshape(-1)
                      for p, gp in zip(params, grads[2:])])
    same(flat, g[f"grads/{tag}/dparams_flat"], 2e-5)


def test_camera_state_update_path(golden):
    """control.py:148-165 restated (oracle.camera_state_update) against 24 frames x 3 cameras produced
    by the reference's own get_state source lines."""
    g = golden("camera_path")
    pos, ori = t32(g["path/position0"]), t32(g["path/orientation0"])
    ts, rs = [float(x) for x in g["path/sens"]]
    script = t32(g["path/script"])
    for f in range(script.shape[0]):
        pos, ori = O.camera_state_update(pos, ori, script[f, :, :3], script[f, :, 3:], ts, rs)
        np.testing.assert_allclose(pos.numpy(), g["path/positions"][f], rtol=0, atol=1e-6)
        np.testing.assert_allclose(ori.numpy(), g["path/orientations"][f], rtol=0, atol=1e-6)
    assert np.allclose(np.linalg.norm(g["path/orientations"], axis=-1), 1.0, atol=1e-6)
