"""Full-size, oracle-backed parity of the BASELINE configs (VERDICT r1 "parity holes").

Every test renders through the public modules (-> torch ops -> ctypes -> C ABI) and compares with
``oracle.sdf_oracle`` on the SAME inputs at the size bench.py runs (c1, c2) or at the largest size the
CPU oracle finishes in seconds (c3, c5, c4b), for the interpreter AND the tape-specialised kernels.

Tolerances are BASELINE.json's: hit depth / SDF 1e-4 relative, pixels 1e-3 absolute, gradients 1e-3
relative (max-norm per tensor), NaN masks equal.  Where a quantity is ill-conditioned in float32 --
the Laplacian (differences of O(1e-2) values scaled by 6/eps^2 = 2400) and chaotic multi-bounce rays --
the yardstick is stated next to the assertion: the float64 oracle is the exact value, and the CUDA
result must be as close to it as the reference's own float32 evaluation is (factor 2), or the
comparison is restricted to the rays on which the reference's float32 and float64 runs agree.
"""
import copy

import numpy as np
import pytest
import torch

from oracle import cases, sdf_oracle as O
from ray_marching_b200 import RenderLoop, scene as S
from ray_marching_b200.rendering.shader import load_cmap
from rm_testutil import t32

pytestmark = pytest.mark.gpu
DEV = "cuda"
PX = 3.45e-6


def _loop(scene, w, h, n=1):
    return RenderLoop(scene.to(DEV), normals_eps=5e-2, **dict(cases.config1_camera(w, h), num_cameras=n)).to(DEV)


def _oracle(scene, w, h, n=1, dtype=torch.float32):
    orc = O.RenderLoopOracle(scene if dtype == torch.float32 else copy.deepcopy(scene).double(),
                             normals_eps=5e-2, cmap=load_cmap(), **dict(cases.config1_camera(w, h), num_cameras=n))
    if dtype != torch.float32:
        orc.base = tuple(b.to(dtype) for b in orc.base)
    return orc


def _cmp_pixels(got, want, what):
    got, want = got.detach().cpu().numpy(), want.detach().cpu().numpy()
    assert got.shape == want.shape, (what, got.shape, want.shape)
    assert np.array_equal(np.isnan(got), np.isnan(want)), f"{what}: NaN masks differ"
    err = np.abs(np.nan_to_num(got) - np.nan_to_num(want)).max()
    assert err <= 1e-3, (what, err)


def _lap_check(lap_gpu, lap32, lap64, what, pmax=1.0):
    """Laplacian: |gpu - exact| <= 2 x |reference float32 - exact| (max over the frame).  Floor: 4 roundings
    of a coordinate of size pmax, each worth 6/eps^2 ulp(pmax) (see tests/test_gpu_forward.py: lap_close)."""
    e_ref = (lap32.double() - lap64).abs().max().item()
    e_gpu = (lap_gpu.cpu().double() - lap64).abs().max().item()
    floor = 4.0 * 2400.0 * float(np.spacing(np.float32(pmax)))
    assert e_gpu <= max(2.0 * e_ref, floor), (what, e_gpu, e_ref, floor)
    return e_gpu, e_ref


@pytest.mark.parametrize("specialise", [False, True])
def test_config1_full_frame_matches_oracle(specialise):
    """c1: make_test_scene2, 640x480, 32 steps, distance + proximity shaders, full frame."""
    W, H = 640, 480
    loop = _loop(S.make_test_scene2(), W, H)
    if specialise:
        loop.specialise(True)
    q = torch.tensor([[1.0, 0.0, 0.0, 0.0]])
    t = torch.tensor([[0.0, 0.0, 1.0]])
    res = loop.render_planes(q.to(DEV), t.to(DEV), gbuffer=True, modes=(1, 2))
    orc = _oracle(S.make_test_scene2(), W, H)
    with torch.no_grad():
        gb = orc.gbuffer(q, t, 32)
        dist = O.shade_distance(gb["px"], gb["pos"]).reshape(-1, 1)
        prox = O.shade_proximity(gb["dists"]).reshape(-1, 1)
    depth_ref = (gb["pos"] - gb["px"]).norm(dim=-1).reshape(-1)
    depth = (res.pos - res.origin).norm(dim=-1).cpu()
    assert ((depth - depth_ref).abs() / depth_ref).max().item() < 1e-4
    assert (res.sdf.cpu() - gb["dists"].reshape(-1)).abs().max().item() < 1e-5      # converged: |f| ~ 1e-8
    # distance: x^(1/2.33) has unbounded slope at 0 (the arg-min pixel): compare before the gamma there
    got, want = res.images[1].cpu(), dist
    low = want < 0.05
    assert (got - want).abs()[~low].max().item() <= 1e-3
    assert (got.pow(2.33) - want.pow(2.33)).abs()[low].max().item() <= 1e-4
    # proximity: every ray converges, the log range is 0 and the reference emits 0/0 = NaN everywhere
    assert torch.isnan(prox).all() and torch.isnan(res.images[2]).all()


@pytest.mark.parametrize("specialise", [False, True])
def test_config2_full_frame_matches_oracle(specialise):
    """c2: make_test_scene2, 1920x1080, normal + laplacian + lambertian + vignette from ONE march --
    exactly the frame bench.py times (specialised) and its interpreter twin."""
    W, H = 1920, 1080
    loop = _loop(S.make_test_scene2(), W, H)
    if specialise:
        loop.specialise(True)
    q = torch.tensor([[1.0, 0.0, 0.0, 0.0]])
    t = torch.tensor([[0.0, 0.0, 1.0]])
    res = loop.render_planes(q.to(DEV), t.to(DEV), gbuffer=True, modes=(4, 5, 0, 3))
    orc = _oracle(S.make_test_scene2(), W, H)
    o64 = _oracle(S.make_test_scene2(), W, H, dtype=torch.float64)
    with torch.no_grad():
        gb = orc.gbuffer(q, t, 32)
        g64 = o64.gbuffer(q.double(), t.double(), 32)
    depth_ref = (gb["pos"] - gb["px"]).norm(dim=-1).reshape(-1)
    depth = (res.pos - res.origin).norm(dim=-1).cpu()
    assert ((depth - depth_ref).abs() / depth_ref).max().item() < 1e-4
    assert (res.normals.cpu() - gb["normals"].reshape(-1, 3)).abs().max().item() < 1e-3
    _lap_check(res.laplacian, gb["lap"].reshape(-1), g64["lap"].reshape(-1), "raw laplacian")
    with torch.no_grad():
        _cmp_pixels(res.images[4], O.shade_normal(gb["normals"]).reshape(-1, 3), "normal")
        _cmp_pixels(res.images[0], O.shade_lambertian(gb["dirs"], gb["normals"]).reshape(-1, 1), "lambertian")
        _cmp_pixels(res.images[3], O.shade_vignette(gb["dirs"], gb["frames"]).reshape(-1, 1), "vignette")
        _cmp_pixels(res.images[5], O.shade_laplacian(gb["lap"]).reshape(-1, 1), "laplacian")


@pytest.mark.parametrize("mode", range(8))
def test_open_scene_shaders_match_reference_fixture(golden, mode):
    """make_test_scene from (0,0,-1.5): 15 % of the rays hit, the rest run off -- 68 % of all rays to
    |p| >= 1e7, where the stencil offsets (0.04) vanish below ulp(|p|), all five samples coincide, the
    normal is 0/0 = NaN and so are the lambertian / normal / tangent / spin pixels.  The reference's own
    images (render/scene1/mode*) through RenderLoop.forward, in three classes taken from its G-buffer:
      hit   (|f| < 1e-3)        pixel within 1e-3, never NaN;
      far   (|p| >= 1e7)        NaN exactly where the reference is NaN (stencil shaders), value otherwise;
      other (missed, |p| < 1e7) not compared for the stencil shaders: there ulp(|p|) is comparable to the
            offsets, the reference's own normal is rounding noise of 1e5-sized numbers (55 of the 86 rays
            with 1e5 <= |p| < 1e7 happen to be NaN in the fixture, 31 are not), and which way it falls
            depends on the last bit of p."""
    g = golden("render")
    loop = _loop(S.make_test_scene(), 32, 24)
    q, t = t32(g["render/scene1/q"]).to(DEV), t32(g["render/scene1/t"]).to(DEV)
    want = g[f"render/scene1/mode{mode}"].reshape(-1, 3)
    with torch.no_grad():
        img = loop(q, t, mode, 3 if mode >= 6 else 1, 32)
    assert img.shape == (1, 24, 32, 3)
    got = img.contiguous().cpu().numpy().reshape(-1, 3)
    sdf = g["render/scene1/sdf"].reshape(-1)
    radius = np.linalg.norm(g["render/scene1/pos"].reshape(-1, 3), axis=1)
    hit, far = np.abs(sdf) < 1e-3, radius >= 1e7
    assert hit.sum() > 100 and far.sum() > 500
    stencil_mode = mode in (0, 4, 5, 6, 7)
    if mode == 5:
        # max|L| -- the normaliser of EVERY pixel -- is taken over the missed rays too, where L is
        # 2400 x rounding noise of 1e5..1e9-sized numbers: the reference's image is noise-scaled.  Compare
        # the un-normalised plane on the hit rays instead.
        res = loop.render_planes(q, t, gbuffer=True)
        s64 = copy.deepcopy(S.make_test_scene()).double()
        with torch.no_grad():
            _, lap64 = O.normals_laplacian(s64, t32(g["render/scene1/pos"].reshape(-1, 3)[hit]).double(), 5e-2)
        _lap_check(res.laplacian[torch.from_numpy(hit).to(DEV)], t32(g["render/scene1/lap"].reshape(-1)[hit]),
                   lap64.reshape(-1), "open-scene laplacian (hit rays)", pmax=2.0)
        return
    assert not np.isnan(got[hit]).any() and not np.isnan(want[hit]).any()
    if stencil_mode:
        assert np.isnan(want[far]).all() and np.isnan(got[far]).all(), f"mode {mode}: far rays must be NaN"
        sel = hit
    else:
        assert np.array_equal(np.isnan(got), np.isnan(want)), f"mode {mode}: NaN masks differ"
        sel = np.ones_like(hit)
    err = np.abs(np.nan_to_num(got[sel]) - np.nan_to_num(want[sel]))
    if mode >= 6:
        assert (err > 1e-3).mean() < 0.01 and err.max() < 2e-2   # one LUT entry off on <1 % of pixels
    elif mode in (1, 2):
        low = np.nan_to_num(want[sel]) < 0.05
        assert err[~low].max(initial=0.0) <= 1e-3
        pre = np.abs(np.nan_to_num(got[sel]) ** 2.33 - np.nan_to_num(want[sel]) ** 2.33)
        assert pre[low].max(initial=0.0) <= 1e-4
    else:
        assert err.max() <= 1e-3, (mode, err.max())


_C3_CACHE = {}


def _c3_oracle(W, H, q, t):
    """float32 and float64 oracle runs of the c3 composition (shared by the parametrisations)."""
    if (W, H) not in _C3_CACHE:
        ref_scene = S.make_synthetic_scene(64, 0, True)
        orc = _oracle(ref_scene, W, H)
        o64 = _oracle(ref_scene, W, H, dtype=torch.float64)
        with torch.no_grad():
            px, frames, pos0, dirs = O.camera_rays(*orc.base, q, t)
            pos, dirs2, hit, normals, lap = O.reflected_march(orc.scene, pos0, dirs, 32, 5e-2, 2)
            lam = O.shade_lambertian(dirs2, normals)
            px6, _, pos06, dirs6 = O.camera_rays(*o64.base, q.double(), t.double())
            pos64, dirs64, hit64, _, _ = O.reflected_march(o64.scene, pos06, dirs6, 32, 5e-2, 2)
        _C3_CACHE[(W, H)] = (pos, dirs2, normals, lam, pos64, dirs64)
    return _C3_CACHE[(W, H)]


@pytest.mark.parametrize("cull", [False, True])
def test_config3_reflections_on_synthetic64_with_bench_math(cull):
    """c3 semantics at 192x108 on the bench's scene (synthetic-64 in the shell) with the bench's math
    policy (approximate sqrt/exp inside march steps) and 2 bounces.

    Multi-bounce rays are chaotic: a reflection off a curved surface amplifies a last-bit difference, and a
    hit test |f| < tau flips when |f| straddles tau.  The comparison set is therefore the rays the
    REFERENCE itself reproduces: those whose float32 and float64 oracle runs end within 5e-4 of each
    other.  On that set >= 99.5 % of the CUDA rays must end within 1e-3 of the float32 oracle (the rest
    are rays whose conditioning sits at the edge of the set), with lambertian within 1e-3 on them."""
    W, H = 192, 108
    scene = S.make_synthetic_scene(64, 0, True)
    loop = _loop(scene, W, H)
    loop.cull = cull
    assert loop.fast_math
    q = torch.tensor([[1.0, 0.0, 0.0, 0.0]])
    t = torch.tensor([[0.0, 0.0, -4.0]])
    res = loop.render_planes(q.to(DEV), t.to(DEV), gbuffer=True, modes=(0,), bounces=2)
    pos, dirs2, normals, lam, pos64, dirs64 = _c3_oracle(W, H, q, t)
    stable = ((pos.double() - pos64).norm(dim=-1) < 5e-4) & ((dirs2.double() - dirs64).norm(dim=-1) < 5e-4)
    frac_stable = stable.float().mean().item()
    assert frac_stable > 0.80, frac_stable                   # the reference reproduces itself on most rays
    got_pos = res.pos.reshape(pos.shape).cpu()
    dist = (got_pos - pos).norm(dim=-1)
    same = dist < 1e-3
    agree = same[stable].float().mean().item()
    tight = dist < 1e-5
    print(f"c3 parity (cull={cull}): {frac_stable:.4f} of rays reproducible by the reference; of those "
          f"{agree:.5f} end within 1e-3 and {tight[stable].float().mean().item():.5f} within 1e-5 of the "
          f"float32 oracle; |dp| quantiles 50/90/99 % = "
          f"{[float(f'{x:.2e}') for x in torch.quantile(dist[stable], torch.tensor([0.5, 0.9, 0.99])).tolist()]}")
    assert agree >= 0.995, (agree, frac_stable)
    both = stable & same
    assert (res.dirs.reshape(pos.shape).cpu() - dirs2)[both].abs().max().item() < 2e-3
    # the pixel: d lambertian / d p is the surface curvature (up to ~10 / unit on these primitives), so the
    # 1e-3 pixel tolerance is asserted where the hit point agrees to 1e-5 (position-limited elsewhere)
    pin = stable & tight
    assert pin.float().mean().item() > 0.5
    e_lam = (res.images[0].reshape(lam.shape).cpu() - lam)[pin].abs()
    assert e_lam.max().item() <= 1e-3, e_lam.max().item()
    e_all = (res.images[0].reshape(lam.shape).cpu() - lam)[both].abs()
    assert (e_all <= 1e-3).float().mean().item() >= 0.99 and e_all.max().item() < 5e-2


def test_config5_synthetic512_matches_oracle():
    """c5's scene (512 primitives in the shell) at 64x36 for two ring views, full evaluation and culled:
    depth vs the oracle on the rays that have settled (grazing rays still move after 32 steps)."""
    W, H = 64, 36
    import math
    scene = S.make_synthetic_scene(512, 0, True)
    loop = _loop(scene, W, H, n=2)
    qs, ts = [], []
    for j in (0, 5):
        th = 2 * math.pi * j / 64
        qj = torch.tensor([[math.cos(th / 2), 0.0, math.sin(th / 2), 0.0]])
        qs.append(qj)
        ts.append(O.q_rotation(torch.tensor([[0.0, 0.0, -4.0]]), qj))
    q, t = torch.cat(qs), torch.cat(ts)
    orc = _oracle(S.make_synthetic_scene(512, 0, True), W, H, n=2)
    with torch.no_grad():
        gb = orc.gbuffer(q, t, 32)
        lam = O.shade_lambertian(gb["dirs"], gb["normals"]).reshape(-1, 1)
    dref = (gb["pos"] - gb["px"]).norm(dim=-1).reshape(-1)
    settled = gb["dists"].reshape(-1).abs() < 1e-4          # the rest graze the blend and still move
    assert settled.float().mean().item() > 0.7
    for cull in (False, True):
        loop.cull = cull
        res = loop.render_planes(q.to(DEV), t.to(DEV), gbuffer=True, modes=(0,))
        depth = (res.pos - res.origin).norm(dim=-1).cpu()
        rel = ((depth - dref).abs() / dref)[settled]
        assert rel.max().item() < 1e-4, (cull, rel.max().item())
        e = (res.images[0].cpu() - lam).abs()[settled]
        assert (e <= 1e-3).float().mean().item() > 0.999, (cull, e.max().item())


@pytest.mark.parametrize("route", ["gbuffer", "loss_fused"])
def test_config4b_gradients_match_reference_fixture(golden, route):
    """c4b: the config-4 loss on synthetic-64 in the shell (650 parameters) at 32x32 -- gradients of every
    scene parameter and of the camera pose vs the REFERENCE's autograd (tests/golden/grads_c4b.npz)."""
    from ray_marching_b200 import ops
    g = golden("grads_c4b")
    scene = S.make_synthetic_scene(64, 0, True)
    loop = _loop(scene, 32, 32)
    q = t32(g["grads/syn64/q"]).to(DEV).requires_grad_(True)
    t = t32(g["grads/syn64/t"]).to(DEV).requires_grad_(True)
    lam_t, depth_t = t32(g["grads/syn64/lam_target"]).to(DEV), t32(g["grads/syn64/depth_target"]).to(DEV)
    if route == "gbuffer":
        gb = loop.gbuffer(q, t, 32)
        lam = (-(gb["dirs"] * gb["normals"]).sum(-1, keepdim=True)).clamp(0, 1)
        depth = (gb["pos"] - gb["origin"]).norm(dim=-1, keepdim=True)
        # (grazing rays are still moving after 32 steps: the depth PLANE is compared on settled rays in
        #  test_config5 / test_culling; here the loss pins the forward and the fixture the gradients)
        d_err = (depth.detach().cpu().numpy() - g["grads/syn64/depth"]).__abs__() / g["grads/syn64/depth"]
        assert np.median(d_err) < 1e-6 and (d_err < 1e-4).mean() > 0.9
        loss = ((lam - lam_t) ** 2).mean() + ((depth - depth_t) ** 2).mean()
    else:
        if not hasattr(loop, "inverse_loss"):
            pytest.skip("loss-fused entry point not built")
        loss = loop.inverse_loss(q, t, lam_t, depth_t, marching_steps=32)
    ref_loss = float(g["grads/syn64/loss"])
    assert abs(loss.item() - ref_loss) < 1e-4 * max(1.0, ref_loss)
    loss.backward()
    rel = lambda a, b: np.abs(a - b).max() / (np.abs(b).max() + 1e-30)
    flat = torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1)
                      for p in scene.parameters()]).cpu().numpy()
    assert flat.shape == g["grads/syn64/dparams_flat"].shape and flat.size >= 640
    # Tolerance: north_star's 1e-3 PLUS the reference's own float32 rounding error, measured as the distance
    # between its float32 autograd and the same modules run in float64 (both committed in the fixture:
    # 3.9e-4 on the parameters, 3.3e-4 on dq).  Against the exact (float64) gradient the budget is 1e-3 + once
    # that noise, against the float32 reference -- two noisy values -- twice.
    for got, k in ((q.grad.cpu().numpy(), "dq"), (t.grad.cpu().numpy(), "dt"), (flat, "dparams_flat")):
        ref32, ref64 = g[f"grads/syn64/{k}"], g[f"grads/syn64/{k}64"]
        noise = rel(ref32, ref64)
        e32, e64 = rel(got, ref32), rel(got, ref64)
        print(f"c4b {route} {k}: vs reference float32 {e32:.2e}, vs float64 {e64:.2e}; reference float32 vs float64 {noise:.2e}")
        assert noise < 1e-3, "the fixture's own float32 gradient is unusable"
        assert e64 < 1e-3 + noise and e32 < 1e-3 + 2 * noise, (k, e32, e64, noise)


def test_config4b_loss_routes_agree_over_several_block_iterations(monkeypatch):
    """c4b's scene at 512 x 512 -- 262 144 rays, more than one grid-stride iteration of the loss kernels, so the
    slot-indexed trajectory / node-record scratch is reused -- through the four routes of rm_render_loss_bwd
    (recording walk; node records; node records with the four-ray replay; replay in a kernel of its own): loss equal to 2e-5, every gradient
    to 3e-4 of its size (the routes differ only in the order of float32 sums and, for the liveness test, in
    the math policy of a value that decides about weights below 1e-9)."""
    scene = S.make_synthetic_scene(64, 0, True)
    loop = _loop(scene, 512, 512)
    q0 = torch.nn.functional.normalize(torch.tensor([[0.98, 0.05, -0.15, 0.02]], device=DEV), dim=-1)
    t0 = torch.tensor([[0.2, -0.1, -4.0]], device=DEV)
    gen = torch.Generator().manual_seed(11)
    lam_t = torch.rand(1, 512, 512, 1, generator=gen).to(DEV)
    depth_t = (3.0 + 0.5 * torch.rand(1, 512, 512, 1, generator=gen)).to(DEV)
    out = {}
    for route in ("record", "onfly", "onfly4", "split"):
        monkeypatch.setenv("RM_B200_LOSS_ROUTE", route)
        for p in scene.parameters():
            p.grad = None
        q, t = q0.clone().requires_grad_(True), t0.clone().requires_grad_(True)
        loss = loop.inverse_loss(q, t, lam_t, depth_t, marching_steps=32)
        loss.backward()
        flat = torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in scene.parameters()])
        out[route] = (loss.item(), q.grad.clone(), t.grad.clone(), flat.clone())
    rel = lambda a, b: ((a - b).abs().max() / (b.abs().max() + 1e-30)).item()
    for route in ("onfly", "onfly4", "split"):
        # (the loss is a float32 sum of 262 144 terms added by atomics: the order differs per launch)
        assert abs(out[route][0] - out["record"][0]) < 2e-5 * max(1.0, abs(out["record"][0])), route
        for i in (1, 2, 3):
            assert rel(out[route][i], out["record"][i]) < 3e-4, (route, i, rel(out[route][i], out["record"][i]))
