"""Headless frame loop (main.py:52-88 without the window), device-side camera paths
(control.py:148-165) and the frame writers, on the GPU."""
import numpy as np
import pytest
import torch

from oracle import sdf_oracle as O
from ray_marching_b200 import RenderLoop, frames
from ray_marching_b200.scene.scene_registry import make_test_scene, make_test_scene2
from rm_testutil import t32

pytestmark = pytest.mark.gpu
PX = 3.45e-6


def make_loop(scene, w=160, h=120):
    return RenderLoop(scene, num_cameras=1, px_width=w, px_height=h, focal_length=PX * h,
                      sensor_width=PX * w, sensor_height=PX * h, normals_eps=5e-2).cuda()


def test_scripted_camera_path_matches_reference_fixture(golden):
    g = golden("camera_path")
    ts, rs = [float(x) for x in g["path/sens"]]
    path = frames.CameraPath(g["path/position0"], g["path/orientation0"], translation_sensitivity=ts,
                             rotation_sensitivity=rs, script=t32(g["path/script"]))
    n_frames = g["path/script"].shape[0]
    for f in range(n_frames):
        path.advance()
        np.testing.assert_allclose(path.position.cpu().numpy(), g["path/positions"][f], rtol=0, atol=2e-6)
        np.testing.assert_allclose(path.orientation.cpu().numpy(), g["path/orientations"][f], rtol=0, atol=2e-6)
    assert int(path.counter.item()) == n_frames
    # the script wraps around: frame n_frames applies row 0 again
    pos, ori = path.position.cpu().clone(), path.orientation.cpu().clone()
    path.advance()
    s0 = t32(g["path/script"])[0]
    want_p, want_o = O.camera_state_update(pos, ori, s0[:, :3], s0[:, 3:], ts, rs)
    np.testing.assert_allclose(path.position.cpu().numpy(), want_p.numpy(), rtol=0, atol=2e-6)
    np.testing.assert_allclose(path.orientation.cpu().numpy(), want_o.numpy(), rtol=0, atol=2e-6)


def test_explicit_camera_steps_and_versor_nan():
    path = frames.CameraPath([(0.0, 0.0, 1.0), (1.0, 2.0, 3.0)], [(1.0, 0.0, 0.0, 0.0), (0.5, 0.5, 0.5, 0.5)],
                             translation_sensitivity=0.1, rotation_sensitivity=0.05)
    pos, ori = path.position.cpu().clone(), path.orientation.cpu().clone()
    ti, oi = torch.tensor([[1.0, -2.0, 0.5]]), torch.tensor([[0.3, 0.1, -0.2]])
    path.step(ti[0], oi[0])
    want_p, want_o = O.camera_state_update(pos, ori, ti, oi, 0.1, 0.05)
    np.testing.assert_allclose(path.position.cpu().numpy(), want_p.numpy(), rtol=0, atol=2e-6)
    np.testing.assert_allclose(path.orientation.cpu().numpy(), want_o.numpy(), rtol=0, atol=2e-6)
    # |orientation_input * sensitivity| > 1: Q.to_versor takes sqrt of a negative number -> NaN state,
    # exactly what the reference does (quaternion.py:75-91)
    path.step((0.0, 0.0, 0.0), (30.0, 0.0, 0.0))
    assert torch.isnan(path.orientation).all()


@pytest.mark.parametrize("mode", [0, 1, 5, 6])
def test_graph_replay_is_bit_identical_to_eager(mode):
    g = torch.Generator().manual_seed(mode)
    script = torch.cat([torch.randn(12, 1, 3, generator=g) * 0.2, torch.randn(12, 1, 3, generator=g) * 0.5], -1)
    loop = make_loop(make_test_scene2())
    outs = []
    for use_graph in (False, True):
        path = frames.CameraPath([(0.0, 0.0, 1.0)], [(1.0, 0.0, 0.0, 0.0)], script=script)
        fl = frames.FrameLoop(loop, path, mode=mode, graph=use_graph)
        outs.append([fl.frame().clone() for _ in range(14)])       # 14 > 12: wraps around
        assert fl.frames == 14 and int(path.counter.item()) == 14
    for a, b in zip(*outs):
        assert torch.equal(torch.nan_to_num(a, nan=-7.0), torch.nan_to_num(b, nan=-7.0))
    # frames differ from one another (the camera really moved)
    assert not torch.equal(torch.nan_to_num(outs[1][0]), torch.nan_to_num(outs[1][5]))


def test_frame_equals_display_frame_and_recaptures_on_parameter_change(tmp_path):
    scene = make_test_scene().cuda()
    loop = make_loop(scene)
    path = frames.CameraPath([(0.0, 0.0, -1.5)], [(1.0, 0.0, 0.0, 0.0)])
    fl = frames.FrameLoop(loop, path, mode=4)
    a = fl.frame((0.0, 0.0, 1.0), (0.0, 0.2, 0.0)).clone()
    want = loop.display_frame(path.orientation.clone(), path.position.clone(), mode=4)
    assert torch.equal(torch.nan_to_num(a), torch.nan_to_num(want))
    graph_before = fl._graph
    b = fl.frame().clone()                                  # static camera, same graph
    assert fl._graph is graph_before and torch.equal(torch.nan_to_num(a), torch.nan_to_num(b))
    with torch.no_grad():
        next(iter(scene.parameters())).add_(0.05)
    c = fl.frame().clone()
    assert fl._graph is not graph_before                    # new lowered parameters -> new capture
    want = loop.display_frame(path.orientation.clone(), path.position.clone(), mode=4)
    assert torch.equal(torch.nan_to_num(c), torch.nan_to_num(want))
    assert not torch.equal(torch.nan_to_num(b), torch.nan_to_num(c))
    frames.write_png(tmp_path / "f.png", c)
    assert (frames.read_png(tmp_path / "f.png") == frames.to_rgba8(c)).all()
    assert c[..., 3].min().item() == 1.0


@pytest.mark.parametrize("specialise,two_graphs", [(False, False), (True, False), (True, True)])
def test_graphed_optimisation_step_follows_the_eager_trajectory(specialise, two_graphs):
    """frames.GraphedStep: config-4 style inverse rendering (lambertian + depth MSE, Adam on the scene
    parameters and the camera) -- five replays of the captured step against five eager steps from the
    same start (the warm-up steps of the capture must not count)."""
    def setup():
        torch.manual_seed(0)
        scene = make_test_scene2()
        loop = make_loop(scene, 64, 48)
        if specialise:
            loop.specialise()
        q = torch.nn.functional.normalize(torch.tensor([[0.99, 0.03, -0.08, 0.05]], device="cuda"), dim=-1)
        t = torch.tensor([[0.05, 0.1, 1.0]], device="cuda")
        with torch.no_grad():
            gb = loop.gbuffer(q, t, 32)
            lam_t = (-(gb["dirs"] * gb["normals"]).sum(-1, keepdim=True)).clamp(0, 1) * 0.9
            depth_t = (gb["pos"] - gb["origin"]).norm(dim=-1, keepdim=True) * 1.05
        q, t = q.clone().requires_grad_(True), t.clone().requires_grad_(True)
        params = list(scene.parameters()) + [q, t]
        opt = torch.optim.Adam(params, lr=1e-2, foreach=True, capturable=True)

        def body():
            gb = loop.gbuffer(q, t, 32)
            lam = (-(gb["dirs"] * gb["normals"]).sum(-1, keepdim=True)).clamp(0, 1)
            depth = (gb["pos"] - gb["origin"]).norm(dim=-1, keepdim=True)
            loss = ((lam - lam_t) ** 2).mean() + ((depth - depth_t) ** 2).mean()
            loss.backward()
            return loss
        return params, opt, body

    params, opt, body = setup()
    eager_losses = []
    for _ in range(5):
        opt.zero_grad(set_to_none=True)
        eager_losses.append(body().item())
        opt.step()
    eager = torch.cat([p.detach().reshape(-1) for p in params]).clone()

    params, opt, body = setup()
    calls = []
    between = (lambda: calls.append(1)) if two_graphs else None     # stands in for the gradient all-reduce
    step = frames.GraphedStep(body, opt, between=between)
    calls.clear()
    graph_losses = [step().item() for _ in range(5)]
    graphed = torch.cat([p.detach().reshape(-1) for p in params])
    assert eager_losses[-1] < eager_losses[0]                       # it optimises
    assert len(calls) == (5 if two_graphs else 0)
    np.testing.assert_allclose(graph_losses, eager_losses, rtol=1e-4)
    # atomics order differs run to run; Adam's normalised update amplifies that slightly
    assert (graphed - eager).abs().max().item() < 2e-4


def test_pipelined_frames_equal_the_plain_sequence():
    """parallel.PipelinedFrames: the all-reduce + finalize of frame i on a side stream under the render
    of frame i+1 leaves exactly the planes of the plain render -> finalize sequence (single process:
    the collective is a no-op, stream ordering and buffer rotation are what is tested)."""
    from ray_marching_b200 import parallel
    loop = make_loop(make_test_scene2(), 96, 64)
    modes = (0, 1, 4, 5)
    pipe = parallel.PipelinedFrames(loop, modes)
    poses = []
    for i in range(5):
        q = torch.nn.functional.normalize(torch.tensor([[1.0, 0.02 * i, -0.03 * i, 0.01]], device="cuda"), dim=-1)
        t = torch.tensor([[0.02 * i, 0.0, 1.0 - 0.03 * i]], device="cuda")
        poses.append((q, t))
    got = []
    for q, t in poses:
        res = pipe.submit(q, t)
        pipe.wait(res)
        got.append([res.images[m].clone() for m in modes])
    # back-to-back without waiting in between: only the last two frames are still in their buffers
    inflight = [pipe.submit(q, t) for q, t in poses]
    pipe.wait()
    torch.cuda.synchronize()
    for (q, t), planes in zip(poses, got):
        ref = loop.render_planes(q, t, modes=modes)
        for m, p in zip(modes, planes):
            assert torch.equal(torch.nan_to_num(p, nan=-7.0), torch.nan_to_num(ref.images[m], nan=-7.0)), m
    for (q, t), res in list(zip(poses, inflight))[-2:]:
        ref = loop.render_planes(q, t, modes=modes)
        for m in modes:
            assert torch.equal(torch.nan_to_num(res.images[m], nan=-7.0), torch.nan_to_num(ref.images[m], nan=-7.0)), m
