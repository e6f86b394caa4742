"""Generate tests/golden/*.npz from the UNMODIFIED reference.  TEST INFRASTRUCTURE ONLY.

Run in the build container only (needs /root/reference):

    python -m oracle.make_golden

Everything is float32 CPU, seeded.  The fixtures pin the oracle (tests/test_oracle_golden.py)
and, through it, the CUDA kernels.  Re-running must reproduce the committed files bit for bit
on the same torch build.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import cases, ref_import  # noqa: E402
from ray_marching_b200.scene.scene_registry import make_test_scene, make_test_scene2  # noqa: E402
from ray_marching_b200.scene.synthetic import make_synthetic_scene  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def np32(t):
    return t.detach().cpu().numpy()


def param_list(module):
    return [(n, p) for n, p in module.named_parameters()]


def gen_quaternion(ref, out):
    g = torch.Generator().manual_seed(11)
    p = torch.randn(48, 4, generator=g)
    q = torch.randn(48, 4, generator=g)
    v = torch.randn(48, 3, generator=g)
    small = torch.rand(48, 3, generator=g) * 0.5
    so3 = ref.Q.QuaternionToSO3()
    out.update({
        "quat/p": np32(p), "quat/q": np32(q), "quat/v": np32(v), "quat/small": np32(small),
        "quat/cross": np32(ref.Q.cross_product(v, small)),
        "quat/multiply": np32(ref.Q.multiply(p, q)),
        "quat/conjugate": np32(ref.Q.conjugate(q)),
        "quat/rotation": np32(ref.Q.rotation(v, q)),
        "quat/to_versor": np32(ref.Q.to_versor(small)),
        "quat/so3": np32(so3(q[:, None, None, :])),
        "quat/reflect": np32(ref.make_reflection_directions(
            torch.nn.functional.normalize(small, dim=-1), v)),
    })


def gen_nodes(ref, out):
    pts = cases.eval_points()
    g = torch.Generator().manual_seed(5)
    w = torch.randn(pts.shape[0], 1, generator=g)
    # keep the far points out of the weighted sum used for gradients (1e9 swamps fp32)
    w[14:16] = 0.0
    out["nodes/points"] = np32(pts)
    out["nodes/weights"] = np32(w)
    for name, node in cases.node_cases(ref).items():
        x = pts.clone().requires_grad_(True)
        d = node(x)
        params = param_list(node)
        grads = torch.autograd.grad((d * w).sum(), [x] + [p for _, p in params],
                                    allow_unused=True)
        out[f"nodes/{name}/value"] = np32(d)
        out[f"nodes/{name}/dpoints"] = np32(grads[0])
        for (pn, p), gp in zip(params, grads[1:]):
            out[f"nodes/{name}/dparam/{pn}"] = np32(gp if gp is not None else torch.zeros_like(p))
    for name, scene in (("test_scene", ref.SR.make_test_scene()),
                        ("test_scene2", ref.SR.make_test_scene2()),
                        ("synthetic8_shell", make_synthetic_scene(8, 0, True, classes=ref)),
                        ("synthetic64", make_synthetic_scene(64, 0, False, classes=ref))):
        x = pts.clone().requires_grad_(True)
        d = scene(x)
        params = param_list(scene)
        grads = torch.autograd.grad((d * w).sum(), [x] + [p for _, p in params], allow_unused=True)
        out[f"scenes/{name}/value"] = np32(d)
        out[f"scenes/{name}/dpoints"] = np32(grads[0])
        flat = torch.cat([(gp if gp is not None else torch.zeros_like(p)).reshape(-1)
                          for (_, p), gp in zip(params, grads[1:])])
        out[f"scenes/{name}/dparams_flat"] = np32(flat)
        out[f"scenes/{name}/params_flat"] = np32(torch.cat([p.reshape(-1) for _, p in params]))


def gen_camera(ref, out):
    g = torch.Generator().manual_seed(3)
    q = torch.nn.functional.normalize(torch.randn(2, 4, generator=g), dim=-1)
    t = torch.randn(2, 3, generator=g)
    cam = ref.RM.PinholeCamera(2, 8, 6, 17e-3, 24e-3, 18e-3)
    pos, frames, _, dirs = cam(q, t)
    out.update({"camera/q": np32(q), "camera/t": np32(t), "camera/pos": np32(pos),
                "camera/dirs": np32(dirs), "camera/frames": np32(frames),
                "camera/base_pos": np32(cam.ray_positions), "camera/base_dirs": np32(cam.ray_directions)})
    # config-1 geometry at full 640x480: three probe pixels
    c1 = cases.config1_camera(640, 480)
    cam1 = ref.RM.PinholeCamera(**c1)
    pos, frames, _, dirs = cam1(torch.tensor([[1.0, 0, 0, 0]]), torch.tensor([[0.0, 0, 1.0]]))
    sel = ([240, 0, 100], [320, 0, 500])
    out["camera/c1_pos_sel"] = np32(pos[0][sel])
    out["camera/c1_dirs_sel"] = np32(dirs[0][sel])
    scene = ref.SR.make_test_scene2()
    marched = ref.RM.SDFMarcher(scene)(pos, dirs, 32)
    normals, lap = ref.RM.SDFNormals(scene, 5e-2)(marched)
    out["camera/c1_hit_sel"] = np32(marched[0][sel])
    out["camera/c1_sdf_sel"] = np32(scene(marched)[0][sel])
    out["camera/c1_normals_sel"] = np32(normals[0][sel])
    out["camera/c1_lap_sel"] = np32(lap[0][sel])


def gen_render(ref, out):
    W, H = 32, 24
    intr = cases.config1_camera(W, H)
    for tag, scene, q, t in (
        ("scene2", ref.SR.make_test_scene2(), [[1.0, 0, 0, 0]], [[0.0, 0, 1.0]]),
        ("scene1", ref.SR.make_test_scene(), [[1.0, 0, 0, 0]], [[0.0, 0, -1.5]]),
        ("syn8", make_synthetic_scene(8, 0, True, classes=ref), [[0.96, 0.1, -0.2, 0.15]], [[0.3, -0.2, -4.0]]),
    ):
        q = torch.tensor(q)
        q = q / q.norm(dim=-1, keepdim=True)
        t = torch.tensor(t)
        loop = ref.RenderLoop(scene, normals_eps=5e-2, **intr)
        with torch.no_grad():
            px, frames, pos0, dirs = loop.camera(q, t)
            pos = loop.marcher(pos0, dirs, 32)
            sd = scene(pos)
            normals, lap = loop.normals(pos)
            out[f"render/{tag}/q"] = np32(q)
            out[f"render/{tag}/t"] = np32(t)
            out[f"render/{tag}/pos0"] = np32(pos0)
            out[f"render/{tag}/dirs"] = np32(dirs)
            out[f"render/{tag}/pos"] = np32(pos)
            out[f"render/{tag}/sdf"] = np32(sd)
            out[f"render/{tag}/normals"] = np32(normals)
            out[f"render/{tag}/lap"] = np32(lap)
            for mode in range(8):
                img = loop(q, t, mode, 3 if mode >= 6 else 1, 32)
                out[f"render/{tag}/mode{mode}"] = np32(img.contiguous())
            out[f"render/{tag}/steps7_mode0"] = np32(loop(q, t, 0, 1, 7).contiguous())


def gen_grads(ref, out):
    """Config-4 style loss through the reference's unrolled autograd graph."""
    W = H = 16
    intr = cases.config1_camera(W, H)
    for tag, scene_fn, t0 in (("scene2", lambda: ref.SR.make_test_scene2(), [0.0, 0.1, 1.0]),
                              ("syn8", lambda: make_synthetic_scene(8, 0, True, classes=ref),
                               [0.3, -0.2, -4.0])):
        scene = scene_fn()
        loop = ref.RenderLoop(scene, normals_eps=5e-2, **intr)
        q = torch.nn.functional.normalize(torch.tensor([[0.98, 0.05, -0.12, 0.08]]), dim=-1)
        t = torch.tensor([t0])

        def gbuf(q, t):
            px, frames, pos0, dirs = loop.camera(q, t)
            pos = loop.marcher(pos0, dirs, 32)
            normals, lap = loop.normals(pos)
            lam = (dirs * normals).sum(-1, keepdim=True).neg().clamp(0, 1)
            depth = (pos - px).norm(dim=-1, keepdim=True)
            return lam, depth, lap

        # target: seeded perturbation of the parameters (p + 0.05 randn, seed 1)
        params = param_list(scene)
        saved = [p.detach().clone() for _, p in params]
        g = torch.Generator().manual_seed(1)
        with torch.no_grad():
            for _, p in params:
                p.add_(0.05 * torch.randn(p.shape, generator=g))
            lam_t, depth_t, _ = gbuf(q, t)
            for (_, p), s in zip(params, saved):
                p.copy_(s)
        q = q.clone().requires_grad_(True)
        t = t.clone().requires_grad_(True)
        lam, depth, lap = gbuf(q, t)
        loss = ((lam - lam_t) ** 2).mean() + ((depth - depth_t) ** 2).mean()
        grads = torch.autograd.grad(loss, [q, t] + [p for _, p in params], allow_unused=True,
                                    retain_graph=True)
        out[f"grads/{tag}/q"] = np32(q)
        out[f"grads/{tag}/t"] = np32(t)
        out[f"grads/{tag}/lam_target"] = np32(lam_t)
        out[f"grads/{tag}/depth_target"] = np32(depth_t)
        out[f"grads/{tag}/loss"] = np32(loss)
        out[f"grads/{tag}/dq"] = np32(grads[0])
        out[f"grads/{tag}/dt"] = np32(grads[1])
        out[f"grads/{tag}/dparams_flat"] = np32(torch.cat(
            [(gp if gp is not None else torch.zeros_like(p)).reshape(-1)
             for (_, p), gp in zip(params, grads[2:])]))
        # second loss exercising the Laplacian and sdf outputs
        loss2 = (lap * 1e-3).pow(2).mean()
        grads2 = torch.autograd.grad(loss2, [q, t] + [p for _, p in params], allow_unused=True)
        out[f"grads/{tag}/lap_dq"] = np32(grads2[0])
        out[f"grads/{tag}/lap_dt"] = np32(grads2[1])
        out[f"grads/{tag}/lap_dparams_flat"] = np32(torch.cat(
            [(gp if gp is not None else torch.zeros_like(p)).reshape(-1)
             for (_, p), gp in zip(params, grads2[2:])]))


def gen_grads_c4b(ref, out):
    """Config-4b: the same loss on synthetic-64 inside the shell (642 + 8 parameters), 32x32, through
    the reference's unrolled autograd graph -- pins the fused backward kernel on the large tape."""
    W = H = 32
    intr = cases.config1_camera(W, H)
    scene = make_synthetic_scene(64, 0, True, classes=ref)
    loop = ref.RenderLoop(scene, normals_eps=5e-2, **intr)
    q = torch.nn.functional.normalize(torch.tensor([[0.98, 0.05, -0.12, 0.08]]), dim=-1)
    t = torch.tensor([[0.0, 0.0, -4.0]])

    def gbuf(q, t):
        px, frames, pos0, dirs = loop.camera(q, t)
        pos = loop.marcher(pos0, dirs, 32)
        normals, lap = loop.normals(pos)
        lam = (dirs * normals).sum(-1, keepdim=True).neg().clamp(0, 1)
        depth = (pos - px).norm(dim=-1, keepdim=True)
        return lam, depth

    params = param_list(scene)
    saved = [p.detach().clone() for _, p in params]
    g = torch.Generator().manual_seed(1)
    with torch.no_grad():
        for _, p in params:
            p.add_(0.05 * torch.randn(p.shape, generator=g))
        lam_t, depth_t = gbuf(q, t)
        for (_, p), s in zip(params, saved):
            p.copy_(s)
    q = q.clone().requires_grad_(True)
    t = t.clone().requires_grad_(True)
    lam, depth = gbuf(q, t)
    loss = ((lam - lam_t) ** 2).mean() + ((depth - depth_t) ** 2).mean()
    grads = torch.autograd.grad(loss, [q, t] + [p for _, p in params], allow_unused=True)
    out["grads/syn64/q"] = np32(q)
    out["grads/syn64/t"] = np32(t)
    out["grads/syn64/lam_target"] = np32(lam_t)
    out["grads/syn64/depth_target"] = np32(depth_t)
    out["grads/syn64/lam"] = np32(lam)
    out["grads/syn64/depth"] = np32(depth)
    out["grads/syn64/loss"] = np32(loss)
    out["grads/syn64/dq"] = np32(grads[0])
    out["grads/syn64/dt"] = np32(grads[1])
    out["grads/syn64/dparams_flat"] = np32(torch.cat(
        [(gp if gp is not None else torch.zeros_like(p)).reshape(-1)
         for (_, p), gp in zip(params, grads[2:])]))
    # the same gradient through the SAME reference modules in float64 (the exact value up to 1e-16): the
    # yardstick for how much of a float32 discrepancy is the reference's own rounding noise
    loop.double()
    params64 = param_list(scene)
    q64 = q.detach().double().requires_grad_(True)
    t64 = t.detach().double().requires_grad_(True)
    lam, depth = gbuf(q64, t64)
    loss64 = ((lam - lam_t.double()) ** 2).mean() + ((depth - depth_t.double()) ** 2).mean()
    grads64 = torch.autograd.grad(loss64, [q64, t64] + [p for _, p in params64], allow_unused=True)
    out["grads/syn64/loss64"] = loss64.detach().numpy()
    out["grads/syn64/dq64"] = grads64[0].numpy()
    out["grads/syn64/dt64"] = grads64[1].numpy()
    out["grads/syn64/dparams_flat64"] = torch.cat(
        [(gp if gp is not None else torch.zeros_like(p)).reshape(-1)
         for (_, p), gp in zip(params64, grads64[2:])]).numpy()


def gen_api_surface(ref):
    """Parameter names of every drop-in class's ``__init__`` / ``forward`` and the ``state_dict``
    keys / shapes / dtypes of a reference ``RenderLoop``: the boundary, pinned."""
    import inspect
    import json
    classes = {}
    table = {name: getattr(ref, name) for name in dir(ref) if name.startswith("SDF")}
    table.update(PinholeCamera=ref.RM.PinholeCamera, SDFMarcher=ref.RM.SDFMarcher,
                 SDFNormals=ref.RM.SDFNormals, Shader=ref.SH.Shader,
                 QuaternionToSO3=ref.Q.QuaternionToSO3)
    loop1 = ref.RenderLoop(ref.SR.make_test_scene2(), 1, 8, 6, 1e-3, 2e-3, 1.5e-3, 5e-2)
    table["RenderLoop"] = type(loop1)
    for name, cls in sorted(table.items()):
        classes[name] = {"init": list(inspect.signature(cls.__init__).parameters),
                         "forward": list(inspect.signature(cls.forward).parameters),
                         "init_defaults": {k: v.default for k, v in inspect.signature(cls.__init__).parameters.items()
                                           if v.default is not inspect.Parameter.empty
                                           and isinstance(v.default, (int, float, str, bool))},
                         "forward_defaults": {k: v.default for k, v in inspect.signature(cls.forward).parameters.items()
                                              if v.default is not inspect.Parameter.empty
                                              and isinstance(v.default, (int, float, str, bool))}}
    functions = {}
    for name in ("cross_product", "multiply", "conjugate", "rotation", "to_versor"):
        functions[f"quaternion.{name}"] = list(inspect.signature(getattr(ref.Q, name)).parameters)
    functions["control.make_reflection_directions"] = list(
        inspect.signature(ref.make_reflection_directions).parameters)
    functions["scene_registry.make_test_scene"] = list(inspect.signature(ref.SR.make_test_scene).parameters)
    functions["scene_registry.make_test_scene2"] = list(inspect.signature(ref.SR.make_test_scene2).parameters)
    state = {}
    for tag, loop in (("scene2_n1_8x6", loop1),
                      ("scene1_n2_8x6", ref.RenderLoop(ref.SR.make_test_scene(), 2, 8, 6, 1e-3, 2e-3, 1.5e-3, 5e-2))):
        state[tag] = [[k, list(v.shape), str(v.dtype)] for k, v in loop.state_dict().items()]
    path = os.path.join(OUT, "api_surface.json")
    with open(path, "w") as f:
        json.dump({"classes": classes, "functions": functions, "state_dict": state}, f, indent=1, sort_keys=True)
    print(f"{path}: {len(classes)} classes, {len(functions)} functions")


def gen_camera_path(ref, out):
    """A scripted 24-frame path of 3 cameras through control.py:148-165 (the reference's own source)."""
    import types
    g = torch.Generator().manual_seed(23)
    n_frames, n_views = 24, 3
    script = torch.cat([torch.randn(n_frames, n_views, 3, generator=g),
                        torch.randn(n_frames, n_views, 3, generator=g) * 2.0], dim=-1)
    script[5, :, :3] = 0.0                      # pure rotation frame
    script[9, :, 3:] = 0.0                      # pure translation frame
    state = types.SimpleNamespace(
        position=torch.tensor([[0.0, 0.0, 1.0], [0.5, -0.25, -2.0], [-1.0, 2.0, 0.5]]),
        orientation=torch.nn.functional.normalize(torch.randn(n_views, 4, generator=g), dim=-1),
        translation_sensitivity=0.05, rotation_sensitivity=0.02)
    out["path/script"] = np32(script)
    out["path/position0"] = np32(state.position)
    out["path/orientation0"] = np32(state.orientation)
    out["path/sens"] = np.array([state.translation_sensitivity, state.rotation_sensitivity], np.float32)
    pos, ori = [], []
    for f in range(n_frames):
        ref.camera_state_update(state, script[f, :, :3], script[f, :, 3:])
        pos.append(np32(state.position).copy())
        ori.append(np32(state.orientation).copy())
    out["path/positions"] = np.stack(pos)
    out["path/orientations"] = np.stack(ori)


def main():
    assert ref_import.available(), "needs /root/reference (build container only)"
    torch.manual_seed(0)
    torch.set_num_threads(1)          # deterministic reduction order
    ref = ref_import.load()
    os.makedirs(OUT, exist_ok=True)
    only = set(sys.argv[1:])          # e.g. `python -m oracle.make_golden api grads_c4b`
    if not only or "api" in only:
        gen_api_surface(ref)
    for name, fn in (("quaternion", gen_quaternion), ("nodes", gen_nodes), ("camera", gen_camera),
                     ("render", gen_render), ("grads", gen_grads), ("grads_c4b", gen_grads_c4b),
                     ("camera_path", gen_camera_path)):
        if only and name not in only:
            continue
        out = {}
        fn(ref, out)
        path = os.path.join(OUT, f"{name}.npz")
        np.savez_compressed(path, **out)
        print(f"{path}: {len(out)} arrays, {os.path.getsize(path)/1024:.1f} KiB")
    if only:
        return
    cmap = ref.shader.cyclic_cmap.contiguous().numpy().astype(np.float32)
    # the colour LUT is an *input* of the tangent/spin shaders (data/cyclic_cmap.pt, (4096,3) f64,
    # loaded at shader.py:177); kept as a float32 data asset so the GPU box has it
    np.save(os.path.join(ROOT, "ray_marching_b200", "data", "cyclic_cmap_f32.npy"), cmap)
    print("cmap", cmap.shape, cmap.dtype)


if __name__ == "__main__":
    main()
