"""GPU parity of the forward path: CUDA kernels (through the drop-in modules -> torch ops ->
ctypes -> C ABI) against the CPU oracle and the committed reference fixtures.

Tolerances (BASELINE.json north_star): hit depth and SDF value 1e-4 relative (an absolute floor
of 1e-5 scene units covers converged hits whose |f| is rounding noise), shaded pixels 1e-3
absolute.  NaN masks must coincide wherever the reference is NaN on purpose.
"""
import numpy as np
import pytest
import torch

from oracle import cases, sdf_oracle as O
from ray_marching_b200 import scene as S
from ray_marching_b200 import RenderLoop, PinholeCamera, SDFMarcher, SDFNormals, Shader
from ray_marching_b200.rendering.shader import load_cmap
from ray_marching_b200.scene.scene_registry import _default_classes
from rm_testutil import t32

pytestmark = pytest.mark.gpu
C = _default_classes()
DEV = "cuda"


def lap_close(lap_gpu, lap32, scene_fn, pos32, what=""):
    """Laplacian = (f(p) - mean_4 f(p + off)) * 6/eps^2: differences of O(1e-2) float32 values scaled by
    2400, i.e. an ABSOLUTE rounding noise of a few 1e-3 in the reference itself.  Yardstick: the float64
    oracle at the reference's hit points is the exact value; the CUDA result must be as close to it as the
    reference's own float32 evaluation is, within a factor 2 (floor 1e-3)."""
    import copy
    s64 = copy.deepcopy(scene_fn()).double()
    with torch.no_grad():
        _, lap64 = O.normals_laplacian(s64, torch.as_tensor(pos32).double(), 5e-2)
    lap64 = lap64.reshape(-1).numpy()
    lap_gpu = lap_gpu.detach().cpu().double().numpy().reshape(-1) if isinstance(lap_gpu, torch.Tensor) else np.asarray(lap_gpu, dtype=np.float64).reshape(-1)
    e_ref = np.abs(np.asarray(lap32, dtype=np.float64).reshape(-1) - lap64).max()
    e_gpu = np.abs(lap_gpu - lap64).max()
    # Floor: ONE float32 rounding of a coordinate of size |p| moves f by up to ulp(|p|) and the Laplacian by
    # 6/eps^2 ulp(|p|) (= 1.1e-3 at |p| = 5, 1.4e-4 at |p| = 0.75); the frame transform and the five samples
    # contribute a handful of such roundings, in a different order than the reference's quaternion form
    # (its own error `e_ref` is a max over mostly flat hits and can be lucky): 4 of them.  This replaces the
    # flat 2e-2 of round 1.  (The GPU evaluates at ITS hit points; |grad L| x 1e-6 is below this floor.)
    floor = 4.0 * (6.0 / 5e-2 ** 2) * float(np.spacing(np.float32(np.abs(np.asarray(pos32)).max())))
    assert e_gpu <= max(2.0 * e_ref, floor), (what, e_gpu, e_ref, floor)


def close(a, b, rtol, atol, what=""):
    a = a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
    b = b.detach().cpu().numpy() if isinstance(b, torch.Tensor) else np.asarray(b)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    assert np.array_equal(np.isnan(a), np.isnan(b)), f"{what}: NaN masks differ"
    np.testing.assert_allclose(np.nan_to_num(a), np.nan_to_num(b), rtol=rtol, atol=atol, err_msg=what)


SCENES = {
    "test_scene": S.make_test_scene, "test_scene2": S.make_test_scene2,
    "syn8": lambda: S.make_synthetic_scene(8, 0, True),
    "syn64": lambda: S.make_synthetic_scene(64, 0, False),
}


@pytest.mark.parametrize("name", list(cases.node_cases(C).keys()) + list(SCENES.keys()))
def test_sdf_eval_matches_oracle_and_golden(golden, name):
    node = cases.node_cases(C)[name] if name not in SCENES else SCENES[name]()
    pts = cases.eval_points()
    want = O.sdf(node, pts).detach()
    got = node.to(DEV)(pts.to(DEV))
    assert got.shape == (pts.shape[0], 1) and got.is_cuda
    close(got, want, 1e-4, 1e-5, name)
    g = golden("nodes")
    key = {"test_scene": "scenes/test_scene/value", "test_scene2": "scenes/test_scene2/value",
           "syn8": "scenes/synthetic8_shell/value", "syn64": "scenes/synthetic64/value"}.get(
               name, f"nodes/{name}/value")
    close(got, g[key], 1e-4, 1e-5, name + " vs reference fixture")


def test_sdf_eval_shapes_and_edge_cases():
    scene = S.make_test_scene2().to(DEV)
    assert scene(torch.zeros(0, 3, device=DEV)).shape == (0, 1)
    for n in (1, 2, 3, 5, 511, 513):                     # ragged tails of the 4-per-thread mapping
        p = torch.randn(n, 3, device=DEV)
        close(scene(p), O.sdf(S.make_test_scene2(), p.cpu()).detach(), 1e-4, 1e-5, f"n={n}")
    p = torch.randn(2, 5, 4, 3, device=DEV)              # arbitrary leading dims, stencil-style
    assert scene(p).shape == (2, 5, 4, 1)
    close(scene(p), O.sdf(S.make_test_scene2(), p.cpu()).detach(), 1e-4, 1e-5)
    nc = p.transpose(0, 1)                               # non-contiguous input
    close(scene(nc), O.sdf(S.make_test_scene2(), nc.cpu()).detach(), 1e-4, 1e-5)
    with pytest.raises(ValueError):
        scene(torch.zeros(4, 2, device=DEV))


def test_camera_matches_reference_fixture(golden):
    g = golden("camera")
    cam = PinholeCamera(2, 8, 6, 17e-3, 24e-3, 18e-3)
    pos, frames, pos2, dirs = cam(t32(g["camera/q"]).to(DEV), t32(g["camera/t"]).to(DEV))
    assert pos is pos2 and pos.shape == (2, 6, 8, 3) and frames.shape == (2, 3, 3)
    close(pos, g["camera/pos"], 1e-5, 1e-7, "pos")
    close(dirs, g["camera/dirs"], 1e-5, 1e-6, "dirs")
    close(frames, g["camera/frames"], 1e-5, 1e-6, "frames")
    close(cam.ray_positions, g["camera/base_pos"], 1e-5, 1e-9, "base pos")
    close(cam.ray_directions, g["camera/base_dirs"], 1e-5, 1e-6, "base dirs")


def _loop(scene, w, h, n=1):
    return RenderLoop(scene.to(DEV), normals_eps=5e-2,
                      **dict(cases.config1_camera(w, h), num_cameras=n)).to(DEV)


@pytest.mark.parametrize("tag,build,fast", [
    ("scene2", S.make_test_scene2, True), ("scene2", S.make_test_scene2, False),
    ("scene1", S.make_test_scene, False), ("syn8", lambda: S.make_synthetic_scene(8, 0, True), True),
])
def test_gbuffer_matches_reference_fixture(golden, tag, build, fast):
    """camera + march + f(p) + stencil on the 32x24 fixtures (config-1 camera, 32 steps)."""
    g = golden("render")
    loop = _loop(build(), 32, 24)
    loop.fast_math = fast
    q, t = t32(g[f"render/{tag}/q"]).to(DEV), t32(g[f"render/{tag}/t"]).to(DEV)
    res = loop.render_planes(q, t, gbuffer=True)
    ref = {k: g[f"render/{tag}/{k}"].reshape(-1, g[f"render/{tag}/{k}"].shape[-1])
           for k in ("pos0", "dirs", "pos", "sdf", "normals", "lap")}
    close(res.origin, ref["pos0"], 1e-5, 1e-7, "origin")
    close(res.dirs, ref["dirs"], 1e-5, 1e-6, "dirs")
    depth_ref = np.linalg.norm(ref["pos"] - ref["pos0"], axis=-1)
    hit = np.abs(ref["sdf"][:, 0]) < 1e-3               # rays that converged in the reference
    assert hit.sum() > 50
    depth = (res.pos - res.origin).norm(dim=-1).cpu().numpy()
    np.testing.assert_allclose(depth[hit], depth_ref[hit], rtol=1e-4)
    np.testing.assert_allclose(res.pos.cpu().numpy()[hit], ref["pos"][hit], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(res.sdf.cpu().numpy()[hit], ref["sdf"][hit, 0], rtol=1e-4, atol=1e-5)
    n_got, n_ref = res.normals.cpu().numpy()[hit], ref["normals"][hit]
    assert np.isfinite(n_got).all()
    np.testing.assert_allclose(n_got, n_ref, atol=1e-3)
    lap_close(res.laplacian[torch.from_numpy(hit).to(DEV)], ref["lap"][hit, 0], build, ref["pos"][hit], tag)
    if tag == "scene2":                                  # closed scene: every ray hits
        assert hit.all()


@pytest.mark.parametrize("tag,build", [("scene2", S.make_test_scene2),
                                       ("syn8", lambda: S.make_synthetic_scene(8, 0, True))])
@pytest.mark.parametrize("mode", range(8))
def test_shaders_match_reference_fixture(golden, tag, build, mode):
    """RenderLoop.forward for every shader mode; pixels within 1e-3 absolute."""
    g = golden("render")
    loop = _loop(build(), 32, 24)
    q, t = t32(g[f"render/{tag}/q"]).to(DEV), t32(g[f"render/{tag}/t"]).to(DEV)
    degree = 3 if mode >= 6 else 1
    with torch.no_grad():
        img = loop(q, t, mode, degree, 32)
    want = g[f"render/{tag}/mode{mode}"]
    assert img.shape == (1, 24, 32, 3) and img.shape == want.shape
    got = img.contiguous().cpu().numpy()
    assert np.array_equal(np.isnan(got), np.isnan(want)), "NaN masks differ"
    err = np.abs(np.nan_to_num(got) - np.nan_to_num(want))
    if mode >= 6:
        # colour-wheel lookup: a 1-ulp angle difference may move a pixel to the neighbouring LUT
        # entry; allow 1% of pixels to be one entry off (the LUT is smooth), none further.
        assert (err > 1e-3).mean() < 0.01 and err.max() < 2e-2
    elif mode in (1, 2, 5):
        # x^(1/2.33) has unbounded slope at 0: at the arg-min pixel the reference is exactly 0 and
        # a 1-ulp difference in the global minimum becomes 1e-6^(0.43) = 2.6e-3.  Near zero the
        # comparison is therefore made before the gamma (x = y^2.33), elsewhere on the pixel.
        low = np.nan_to_num(want) < 0.05
        assert err[~low].max(initial=0.0) <= 1e-3, (mode, err[~low].max())
        pre = np.abs(np.nan_to_num(got) ** 2.33 - np.nan_to_num(want) ** 2.33)
        assert pre[low].max(initial=0.0) <= 1e-4, (mode, pre[low].max())
    else:
        assert err.max() <= 1e-3, (mode, err.max())
    # headless display buffer (H,W,4), alpha 1, same pixels
    rgba = loop.display_frame(q, t, mode, degree, 32)
    assert rgba.shape == (24, 32, 4) and rgba.is_contiguous()
    assert torch.equal(rgba[..., 3], torch.ones_like(rgba[..., 3]))
    e2 = np.abs(np.nan_to_num(rgba[..., :3].cpu().numpy()) - np.nan_to_num(want[0]))
    e2 = e2[np.nan_to_num(want[0]) >= 0.05] if mode in (1, 2, 5) else e2
    assert (e2 > 1e-3).sum() <= (0.01 if mode >= 6 else 0.0) * e2.size
    assert np.array_equal(np.isnan(rgba[..., :3].cpu().numpy()), np.isnan(want[0]))


def test_proximity_is_all_nan_on_converged_scene(golden):
    """On scene2 every ray converges, log-range is 0 and the reference emits 0/0 = NaN."""
    g = golden("render")
    assert np.isnan(g["render/scene2/mode2"]).all()
    loop = _loop(S.make_test_scene2(), 32, 24)
    with torch.no_grad():
        img = loop(torch.tensor([[1.0, 0, 0, 0]], device=DEV), torch.tensor([[0.0, 0, 1.0]], device=DEV), 2)
    assert torch.isnan(img).all()


def test_marcher_and_normals_modules(golden):
    g = golden("render")
    scene = S.make_test_scene2().to(DEV)
    pos0, dirs = t32(g["render/scene2/pos0"]).to(DEV), t32(g["render/scene2/dirs"]).to(DEV)
    with torch.no_grad():
        pos = SDFMarcher(scene)(pos0, dirs, 32)
        n, lap = SDFNormals(scene, 5e-2).to(DEV)(pos)
        pos7 = SDFMarcher(scene)(pos0, dirs, 7)
    assert pos.shape == pos0.shape and n.shape == pos0.shape and lap.shape == (1, 24, 32, 1)
    close(pos, g["render/scene2/pos"], 1e-4, 1e-5, "march")
    close(n, g["render/scene2/normals"], 0, 1e-3, "normals")
    lap_close(lap, g["render/scene2/lap"], S.make_test_scene2, g["render/scene2/pos"].reshape(-1, 3), "laplacian")
    ref7 = O.march(S.make_test_scene2(), pos0.cpu(), dirs.cpu(), 7)
    close(pos7, ref7.detach(), 1e-4, 1e-5, "march 7 steps")
    assert SDFMarcher(scene)(pos0, dirs, 0).equal(pos0)


def test_early_exit_is_bit_identical():
    loop = _loop(S.make_test_scene2(), 160, 120)
    q = torch.tensor([[1.0, 0, 0, 0]], device=DEV)
    t = torch.tensor([[0.0, 0, 1.0]], device=DEV)
    a = loop.render_planes(q, t, gbuffer=True, modes=(0, 4))
    loop.early_exit = True
    b = loop.render_planes(q, t, gbuffer=True, modes=(0, 4))
    for x, y in ((a.pos, b.pos), (a.sdf, b.sdf), (a.normals, b.normals), (a.images[0], b.images[0])):
        assert torch.equal(x, y)


def test_multi_view_and_ray_ranges():
    """N=3 views: per-view vignette (the reference only broadcasts for N=1), and a render split
    into ray ranges is bit-identical to the one-shot render (rays are independent)."""
    scene = S.make_synthetic_scene(8, 0, True)
    loop = _loop(scene, 40, 30, n=3)
    q = torch.nn.functional.normalize(torch.tensor(
        [[1.0, 0, 0, 0], [0.9, 0.1, -0.3, 0.05], [0.7, -0.2, 0.6, 0.1]], device=DEV), dim=-1)
    t = torch.tensor([[0.0, 0, -4.0], [0.5, 0.2, -3.5], [-2.0, 0.3, -3.0]], device=DEV)
    full = loop.render_planes(q, t, gbuffer=True, modes=(0, 3, 4))
    orc = O.RenderLoopOracle(S.make_synthetic_scene(8, 0, True), normals_eps=5e-2,
                             **dict(cases.config1_camera(40, 30), num_cameras=3))
    with torch.no_grad():
        gb = orc.gbuffer(q.cpu(), t.cpu(), 32)
        lam = O.shade_lambertian(gb["dirs"], gb["normals"])
        vig = torch.stack([O.shade_vignette(gb["dirs"][i:i + 1], gb["frames"][i:i + 1]) for i in range(3)])
    conv = (gb["dists"].abs() < 1e-4)[..., 0]             # grazing rays still moving after 32 steps are
    assert conv.float().mean() > 0.99                      # chaotic in float32; compare converged ones
    close(full.pos.reshape(3, 30, 40, 3)[conv], gb["pos"][conv], 1e-4, 1e-5, "pos")
    close(full.images[0].reshape(3, 30, 40, 1), lam, 0, 1e-3, "lambertian")
    close(full.images[3].reshape(3, 30, 40, 1), vig.reshape(3, 30, 40, 1), 0, 1e-3, "vignette")
    total = 3 * 30 * 40
    cuts = [0, 1001, 2402, total]                         # deliberately not multiples of 4
    for a, b in zip(cuts[:-1], cuts[1:]):
        part = loop.render_planes(q, t, gbuffer=True, modes=(0, 4), ray_begin=a, n_rays=b - a)
        assert torch.equal(part.pos, full.pos[a:b])
        assert torch.equal(part.images[4], full.images[4][a:b])
        assert torch.equal(part.images[0], full.images[0][a:b])


def test_reflected_march_matches_oracle_composition():
    """Config-3 semantics: B=2 bounces, tau=1e-3, delta=2*eps."""
    scene = S.make_synthetic_scene(8, 0, True)
    loop = _loop(scene, 48, 36)
    loop.fast_math = False
    q = torch.tensor([[1.0, 0, 0, 0]], device=DEV)
    t = torch.tensor([[0.0, 0, -4.0]], device=DEV)
    res = loop.render_planes(q, t, gbuffer=True, modes=(0,), bounces=2)
    orc = O.RenderLoopOracle(S.make_synthetic_scene(8, 0, True), normals_eps=5e-2,
                             **cases.config1_camera(48, 36))
    with torch.no_grad():
        px, frames, pos0, dirs = O.camera_rays(*orc.base, q.cpu(), t.cpu())
        pos, dirs2, hit, normals, lap = O.reflected_march(orc.scene, pos0, dirs, 32, 5e-2, 2)
    # a reflection decision flips if |f| straddles tau within rounding: compare where positions agree
    got = res.pos.reshape(pos.shape).cpu()
    same = (got - pos).norm(dim=-1) < 1e-3
    assert same.float().mean() > 0.97
    close(res.dirs.reshape(pos.shape)[same], dirs2[same], 0, 2e-3, "reflected dirs")
    close(res.normals.reshape(pos.shape)[same], normals[same], 0, 2e-3, "normals")


def test_config1_full_size_properties():
    """640x480 config 1: closed scene => every ray converges (|f| < 1e-5), depth of the centre
    pixel is the SURVEY.md probe value, normals are unit, distance image spans [0,1]."""
    loop = _loop(S.make_test_scene2(), 640, 480)
    q = torch.tensor([[1.0, 0, 0, 0]], device=DEV)
    t = torch.tensor([[0.0, 0, 1.0]], device=DEV)
    res = loop.render_planes(q, t, gbuffer=True, modes=(1, 2))
    assert res.sdf.abs().max().item() < 1e-5
    depth = (res.pos - res.origin).norm(dim=-1).reshape(480, 640)
    assert abs(depth[240, 320].item() - 0.25000021) < 1e-4 * 0.25
    assert abs(depth[0, 0].item() - 0.30783942) < 1e-4 * 0.31
    assert abs(depth[100, 500].item() - 0.27159759) < 1e-4 * 0.27
    assert (res.normals.norm(dim=-1) - 1).abs().max().item() < 1e-5
    d = res.images[1]
    assert d.min().item() == 0.0 and abs(d.max().item() - 1.0) < 1e-6
    assert torch.isnan(res.images[2]).all()


@pytest.mark.parametrize("n_prims", [64, 512])
def test_culling_matches_full_evaluation(n_prims):
    """Bounding-sphere culling of smooth-union children (march steps only): hit
    positions equal the fully evaluated march to 1e-5 relative depth and the oracle to 1e-4."""
    scene = S.make_synthetic_scene(n_prims, 0, True)
    intr = cases.config1_camera(96, 54)
    loop = RenderLoop(scene, normals_eps=5e-2, **intr).to(DEV)
    q = torch.tensor([[1.0, 0.0, 0.0, 0.0]], device=DEV)
    t = torch.tensor([[0.0, 0.0, -4.0]], device=DEV)
    full = loop.render_planes(q, t, gbuffer=True, modes=(0,))
    loop.cull = True
    cul = loop.render_planes(q, t, gbuffer=True, modes=(0,))
    depth = lambda r: (r.pos - r.origin).norm(dim=-1)
    rel = ((depth(cul) - depth(full)).abs() / depth(full)).max().item()
    assert rel < 1e-5, rel
    assert (cul.images[0] - full.images[0]).abs().max().item() < 1e-3
    if n_prims == 64:
        ref_scene = S.make_synthetic_scene(n_prims, 0, True)
        orc = O.RenderLoopOracle(ref_scene, normals_eps=5e-2, **intr)
        with torch.no_grad():
            gb = orc.gbuffer(q.cpu(), t.cpu(), 32)
        dref = (gb["pos"] - gb["px"]).norm(dim=-1).reshape(-1)
        settled = gb["dists"].reshape(-1).abs() < 1e-4       # grazing rays are still moving at step 32
        assert settled.float().mean().item() > 0.9
        assert ((depth(cul).cpu() - dref).abs() / dref)[settled].max().item() < 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize("n_prims,W,H,bounces", [(64, 128, 72, 2), (64, 96, 54, 2), (512, 64, 36, 0), (8, 64, 32, 1)])
def test_exact_culling_is_bit_identical(n_prims, W, H, bounces):
    """RenderLoop.cull = "exact" (fast_math = 3): the only thing skipped is a whole smooth union under a hard
    union where its bounding spheres prove, for every ray of the warp, that it loses the min -- min(parent, f_U)
    = parent -- so every plane equals the fully evaluated frame BIT FOR BIT (march positions, stencil, shaders,
    reflected segments; patch tiles and linear tiles)."""
    scene = S.make_synthetic_scene(n_prims, 0, True)
    loop = _loop(scene, W, H)
    q = torch.nn.functional.normalize(torch.tensor([[0.98, 0.05, -0.15, 0.02]], device=DEV), dim=-1)
    t = torch.tensor([[0.2, -0.1, -4.0]], device=DEV)
    kw = dict(marching_steps=32, bounces=bounces, hit_tau=1e-3, lift_delta=0.1)
    full = loop.render_planes(q, t, modes=(0, 3, 4), **kw)
    full_g = loop.render_planes(q, t, gbuffer=True, modes=(0,), **kw)
    loop.cull = "exact"
    ex = loop.render_planes(q, t, modes=(0, 3, 4), **kw)
    ex_g = loop.render_planes(q, t, gbuffer=True, modes=(0,), **kw)
    same = lambda a, b: torch.equal(a, b) or torch.equal(torch.nan_to_num(a, nan=-7.0), torch.nan_to_num(b, nan=-7.0))
    for m in (0, 3, 4):
        assert same(ex.images[m], full.images[m]), m
    for name in ("pos", "sdf", "normals", "laplacian", "dirs", "origin"):
        assert same(getattr(ex_g, name), getattr(full_g, name)), name
    with pytest.raises(ValueError):
        loop.cull = "roughly"
        loop.render_planes(q, t, modes=(0,))


@pytest.mark.gpu
@pytest.mark.parametrize("W,H,n", [(128, 40, 1), (64, 24, 2), (192, 16, 1)])
def test_culled_patch_tiles_put_every_pixel_in_place(W, H, n):
    """The culled kernel walks 16 x 8-pixel warp patches when the width is a multiple of 64 (render_kernel):
    every pixel is still written, in its own place -- whole frames (a ragged last patch row, two views) and
    row-interleaved strip shards, against the fully evaluated frame with linear tiles."""
    from ray_marching_b200 import parallel
    scene = S.make_synthetic_scene(64, 0, True)
    loop = _loop(scene, W, H, n=n)
    q = torch.nn.functional.normalize(torch.tensor([[1.0, 0, 0, 0], [0.9, 0.1, -0.3, 0.05]], device=DEV), dim=-1)[:n]
    t = torch.tensor([[0.0, 0, -4.0], [0.5, 0.2, -3.5]], device=DEV)[:n]
    full = loop.render_planes(q, t, gbuffer=True, modes=(0, 3))
    depth = lambda r: (r.pos - r.origin).norm(dim=-1)
    loop.cull = True
    cul = loop.render_planes(q, t, gbuffer=True, modes=(0, 3))
    assert torch.equal(cul.origin, full.origin) and torch.equal(cul.dirs, full.dirs)      # exact: placement
    assert torch.equal(cul.images[3], full.images[3])                                   # vignette: ray only
    assert ((depth(cul) - depth(full)).abs() / depth(full)).max().item() < 1e-5
    total = n * W * H
    for G, strip in ((2, 8 * W), (3, 4 * W)):
        got = torch.full_like(full.pos, float("nan"))
        for g_ in range(G):
            sh = parallel.shard_strips(total, g_, G, strip)
            part = loop.render_planes(q, t, gbuffer=True, modes=(0,), **sh)
            idx = parallel.strip_ray_index(sh, device=DEV)
            assert torch.equal(part.origin, full.origin[idx]) and torch.equal(part.dirs, full.dirs[idx])
            got[idx] = part.pos
        d = (got - full.origin).norm(dim=-1)
        assert ((d - depth(full)).abs() / depth(full)).max().item() < 1e-5


@pytest.mark.gpu
def test_inline_rounded_math_is_bit_identical_to_ieee():
    """csrc/rm_math.cuh: the branch-free sqrt / div / reciprocal the IEEE-policy kernels use are the
    in-range instruction sequences of sqrt.rn / div.rn; they must agree bit for bit with the
    compiler's versions (specials included) on this path's operand range."""
    import ctypes as C
    from ray_marching_b200 import _lib
    g = torch.Generator().manual_seed(7)
    n = 1 << 20
    mag = torch.exp2(torch.empty(n).uniform_(-58.0, 58.0, generator=g))      # quotients stay inside 2^+-98
    a = mag * torch.where(torch.rand(n, generator=g) < 0.5, -1.0, 1.0)
    b = torch.exp2(torch.empty(n).uniform_(-40.0, 40.0, generator=g)) * torch.where(torch.rand(n, generator=g) < 0.5, -1.0, 1.0)
    a[:16] = torch.tensor([0.0, -0.0, float("inf"), float("-inf"), float("nan"), 1.0, 4.0, 2.0, 1e-38, 1e-40, 1e-44,
                           3e38, 0.0, 5.0, float("inf"), 1.0])
    b[:16] = torch.tensor([1.0, 1.0, 2.0, 2.0, 1.0, 0.0, -0.0, float("inf"), 1e-10, 1.0, 1.0,
                           1.0, 0.0, float("nan"), float("inf"), 3.0])
    a_d, b_d = a.cuda(), b.cuda()
    out = torch.empty(6, n, device="cuda")
    st = _lib.lib().rm_probe_math(C.c_void_p(a_d.data_ptr()), C.c_void_p(b_d.data_ptr()), n,
                                  C.c_void_p(out.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    _lib.check(st, "rm_probe_math")
    bits = out.cpu().view(torch.int32)
    nan = torch.isnan(out.cpu())
    # sqrt below sqrt.rn's in-range window (0 < x < 2^-101, lengths under 6e-16) is the hardware's
    # approximate root: <= 2 ulp, denormal arguments flush to 0
    tiny = (a > 0) & (a < 3.9443045e-31)
    o = out.cpu()
    assert tiny.any() and ((o[0] - o[1]).abs()[tiny] <= 2.4e-7 * o[1][tiny] + 1.1e-19).all()
    for mine, ref, what in ((0, 1, "sqrt"), (2, 3, "div"), (4, 5, "rcp")):
        same = (bits[mine] == bits[ref]) | (nan[mine] & nan[ref])
        if what == "sqrt":
            same = same | tiny
        bad = (~same).nonzero().flatten()
        assert bad.numel() == 0, (what, a[bad[:5]], b[bad[:5]], out[mine].cpu()[bad[:5]], out[ref].cpu()[bad[:5]])


@pytest.mark.gpu
def test_half_precision_image_planes_for_transport():
    """rm_planes_to_f16: binary16 copies of the finalised image planes equal
    torch's round-to-nearest conversion bit for bit, stay within the 1e-3 pixel tolerance, keep NaNs,
    handle ray counts that are not multiples of 4, and leave unrequested modes alone."""
    from ray_marching_b200 import RenderLoop, ops
    from ray_marching_b200.scene.scene_registry import make_test_scene2
    PXS = 3.45e-6
    for w, h in ((61, 33), (64, 48)):
        loop = RenderLoop(make_test_scene2(), num_cameras=1, px_width=w, px_height=h, focal_length=PXS * h,
                          sensor_width=PXS * w, sensor_height=PXS * h, normals_eps=5e-2).cuda()
        q = torch.tensor([[1.0, 0.0, 0.0, 0.0]], device="cuda")
        t = torch.tensor([[0.0, 0.0, 1.0]], device="cuda")
        res = loop.render_planes(q, t, modes=(0, 1, 2, 3, 4, 5))
        mask = 0b111101                                    # every rendered mode but distance (1)
        half = ops.planes_to_half(res, mask)
        assert half[1] is None and half[6] is None
        for m in (0, 2, 3, 4, 5):
            ref = res.images[m].half()
            assert half[m].dtype == torch.float16 and half[m].shape == res.images[m].shape
            same = (half[m].view(torch.int16) == ref.view(torch.int16)) | (torch.isnan(half[m]) & torch.isnan(ref))
            assert same.all(), m
            err = (half[m].float() - res.images[m]).abs()
            assert torch.nan_to_num(err, nan=0.0).max().item() < 1e-3, m
        assert torch.isnan(half[2]).all()                  # proximity is all-NaN on this scene, as in the reference
        again = ops.planes_to_half(res, mask, out=half)    # buffers are reused
        assert all(a is b for a, b in zip(again, half))


@pytest.mark.gpu
def test_planes_to_f16_argument_errors():
    import ctypes as C
    from ray_marching_b200 import _lib
    lib = _lib.lib()
    stream = torch.cuda.current_stream().cuda_stream
    src = (C.c_void_p * 8)()
    dst = (C.c_void_p * 8)()
    out = torch.empty(4, dtype=torch.float16, device="cuda")
    dst[0] = out.data_ptr()
    assert lib.rm_planes_to_f16(1, 4, src, dst, stream) != 0            # mode 0 asked for, source plane missing
    assert b"image[0] is null" in lib.rm_last_error()
    assert lib.rm_planes_to_f16(1, -1, src, dst, stream) != 0
    assert lib.rm_planes_to_f16(0, 4, src, dst, stream) == 0            # nothing requested: no launch
    assert lib.rm_planes_to_f16(1, 0, src, dst, stream) != 0            # still validates the table


@pytest.mark.gpu
def test_supersampled_frame_is_the_box_filter_of_the_fine_camera():
    """RenderLoop.supersampled: factor^2 rays per pixel = the oracle's PinholeCamera at
    factor x the resolution on the same sensor, shaded, then averaged per pixel."""
    from oracle import sdf_oracle as O
    from ray_marching_b200 import RenderLoop
    from ray_marching_b200.scene.scene_registry import make_test_scene2
    PXS, W, H, F = 3.45e-6, 40, 24, 3
    intr = dict(num_cameras=1, focal_length=PXS * H, sensor_width=PXS * W, sensor_height=PXS * H)
    loop = RenderLoop(make_test_scene2(), px_width=W, px_height=H, normals_eps=5e-2, **intr).cuda()
    q = torch.nn.functional.normalize(torch.tensor([[0.99, 0.03, -0.08, 0.05]]), dim=-1)
    t = torch.tensor([[0.05, 0.1, 1.0]])
    orc = O.RenderLoopOracle(make_test_scene2(), px_width=W * F, px_height=H * F, normals_eps=5e-2, **intr)
    for mode in (0, 4):
        got = loop.supersampled(q.cuda(), t.cuda(), mode=mode, factor=F)
        assert got.shape == (1, H, W, 3)
        with torch.no_grad():
            fine = orc(q, t, mode, 1, 32)                                   # (1, H F, W F, 3)
        want = fine.reshape(1, H, F, W, F, 3).mean(dim=(2, 4))
        assert (got.cpu() - want).abs().max().item() < 1e-3, mode
    with torch.no_grad():                              # the fused path (autograd on takes the G-buffer route)
        one = loop.supersampled(q.cuda(), t.cuda(), mode=0, factor=1)
        assert torch.equal(one, loop(q.cuda(), t.cuda(), 0))


@pytest.mark.gpu
def test_interleaved_strip_shards_equal_the_full_frame():
    """parallel.shard_strips (SURVEY 8(e) "row-interleaved tiles"): G interleaved shards of one frame,
    scattered back by strip_ray_index, are bit-identical to the one-shot render (rays are independent);
    two views, a short last strip, a normalising shader whose reduce scratch the shards share."""
    from ray_marching_b200 import parallel
    scene = S.make_synthetic_scene(8, 0, True)
    loop = _loop(scene, 40, 30, n=2)
    q = torch.nn.functional.normalize(torch.tensor([[1.0, 0, 0, 0], [0.9, 0.1, -0.3, 0.05]], device=DEV), dim=-1)
    t = torch.tensor([[0.0, 0, -4.0], [0.5, 0.2, -3.5]], device=DEV)
    full = loop.render_planes(q, t, gbuffer=True, modes=(0, 1, 4))
    total = 2 * 30 * 40
    for G, strip in ((3, 160), (4, 4 * 40), (2, 1000)):
        pos = torch.empty_like(full.pos)
        lam = torch.empty_like(full.images[0])
        nrm = torch.empty_like(full.images[4])
        dist = torch.empty_like(full.images[1])
        reduce = torch.empty(8, device=DEV)
        from ray_marching_b200 import ops
        ops.check(ops.lib().rm_reduce_init(reduce.data_ptr(), torch.cuda.current_stream().cuda_stream), "init")
        parts = []
        for g_ in range(G):
            sh = parallel.shard_strips(total, g_, G, strip)
            part = loop.render_planes(q, t, gbuffer=True, modes=(0, 1, 4), finalize=False, **sh)
            assert part.pos.shape[0] == sh["n_rays"]
            # shared scratch: combine like the MAX all-reduce would
            reduce = torch.maximum(reduce, part.reduce)
            parts.append((sh, part))
        for sh, part in parts:
            part.reduce = reduce
            ops.shade_finalize(part, 1 << 1)
            idx = parallel.strip_ray_index(sh, device=DEV)
            pos[idx], lam[idx], nrm[idx], dist[idx] = part.pos, part.images[0], part.images[4], part.images[1]
        assert torch.equal(pos, full.pos) and torch.equal(lam, full.images[0]) and torch.equal(nrm, full.images[4])
        assert torch.equal(dist, full.images[1])


@pytest.mark.gpu
@pytest.mark.parametrize("tag,build,t0,bounces,cull", [
    ("scene2", S.make_test_scene2, (0.0, 0.0, 1.0), 0, False),
    ("syn8", lambda: S.make_synthetic_scene(8, 0, True), (0.3, -0.2, -4.0), 0, False),
    ("syn64", lambda: S.make_synthetic_scene(64, 0, True), (0.0, 0.0, -4.0), 0, True),
    ("syn64", lambda: S.make_synthetic_scene(64, 0, True), (0.0, 0.0, -4.0), 2, False),
])
def test_fast_stencil_path_matches_the_rounded_one(tag, build, t0, bounces, cull):
    """rm_render picks march_stencil_fast (one approximate-math interpreter instance for the march AND the
    stencil samples) when the frame outputs neither a Laplacian, nor f(p), nor a G-buffer plane.  Its pixels
    must equal those of the correctly rounded stencil (requested here through gbuffer=True) far inside the
    1e-3 pixel tolerance: the approximate sqrt / exp move f by 2^-22 relative, the normal by ~1e-5."""
    loop = _loop(build(), 96, 64)
    loop.cull = cull
    q = torch.nn.functional.normalize(torch.tensor([[0.99, 0.03, -0.08, 0.05]], device=DEV), dim=-1)
    t = torch.tensor([list(t0)], device=DEV)
    modes = (0, 3, 4, 6, 7)
    ieee = loop.render_planes(q, t, gbuffer=True, modes=modes, bounces=bounces, degree=2)
    fast = loop.render_planes(q, t, modes=modes, bounces=bounces, degree=2)
    assert fast.pos is None and fast.normals is None
    if bounces == 0:
        ok = torch.ones(96 * 64, dtype=torch.bool, device=DEV)
    else:
        # multi-bounce rays are chaotic (tests/test_gpu_fullsize.py): compare where the math policy of the
        # MARCH does not move the end point (IEEE march vs approximate march, both with the rounded stencil)
        loop.fast_math = False
        exact = loop.render_planes(q, t, gbuffer=True, modes=modes, bounces=bounces, degree=2)
        loop.fast_math = True
        ok = (exact.pos - ieee.pos).norm(dim=-1) < 1e-5
        assert ok.float().mean().item() > 0.5
    for m in (0, 3, 4):
        d = (fast.images[m] - ieee.images[m]).abs()[ok]
        assert torch.equal(torch.isnan(fast.images[m]), torch.isnan(ieee.images[m]))
        frac = (d <= 1e-4).float().mean().item()
        assert frac >= (1.0 if bounces == 0 else 0.97) and (bounces or d.max().item() <= 1e-4), (m, frac, d.max().item())
    for m in (6, 7):                                     # colour wheel: a LUT entry off on < 1 % of pixels
        d = (fast.images[m] - ieee.images[m]).abs()[ok]
        assert (d > 1e-3).float().mean().item() < 0.01
    # the display buffer takes the same path
    rgba = loop.display_frame(q, t, mode=0)
    assert (rgba[..., 0].reshape(-1) - fast.images[0].reshape(-1)).abs().max().item() == 0.0 or bounces
