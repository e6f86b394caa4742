"""Host logic: scene tree -> tape lowering, checked on CPU against the oracle through a torch
executor that mirrors the CUDA interpreter (tests/tape_interp.py)."""
import numpy as np
import pytest
import torch

import tape_interp
from oracle import cases, sdf_oracle as O
from ray_marching_b200 import scene as S
from ray_marching_b200 import tape as T
from ray_marching_b200.scene.scene_registry import _default_classes

C = _default_classes()
ALL = dict(cases.node_cases(C))
ALL.update(test_scene=S.make_test_scene(), test_scene2=S.make_test_scene2(),
           syn8=S.make_synthetic_scene(8, 0, True), syn64=S.make_synthetic_scene(64, 0, False))


@pytest.mark.parametrize("name", list(ALL.keys()))
def test_lowered_tape_matches_oracle(name):
    """Values in float32 (the precision the kernels run in); gradients in float64, where the
    comparison is exact to rounding.  (In float32 the reference's own autograd loses e.g. d/dk of
    a smooth union to cancellation -- 0.0 instead of 0.0272 on `smooth_nested` -- so float32
    gradients are not a usable yardstick for the lowering logic.)"""
    import copy
    node = ALL[name]
    tape = T.Tape(node)
    pts = cases.eval_points()
    pts = torch.cat([pts[:14], pts[16:]])   # the 1e3 / 1e9 points are checked on the GPU, values only
    with torch.no_grad():
        d = tape_interp.run(tape, tape.lower_theta(tape.theta()), pts)
        np.testing.assert_allclose(d.numpy(), O.sdf(node, pts).numpy(), rtol=2e-5, atol=2e-6)
    node64 = copy.deepcopy(node).double()
    tape64 = T.Tape(node64)
    p64 = pts.double()
    theta = tape64.theta().detach().requires_grad_(True)
    x = p64.clone().requires_grad_(True)
    d = tape_interp.run(tape64, tape64.lower_theta(theta), x)
    xo = p64.clone().requires_grad_(True)
    do = O.sdf(node64, xo)
    np.testing.assert_allclose(d.detach().numpy(), do.detach().numpy(), rtol=1e-12, atol=1e-12)
    g = torch.Generator().manual_seed(7)
    w = torch.randn(pts.shape[0], 1, generator=g).double()
    params = list(node64.parameters())
    gi = torch.autograd.grad((d * w).sum(), [x, theta], allow_unused=True)
    go = torch.autograd.grad((do * w).sum(), [xo] + params, allow_unused=True)
    np.testing.assert_allclose(gi[0].numpy(), go[0].numpy(), rtol=1e-9, atol=1e-9)
    if params:
        flat = torch.cat([(gp if gp is not None else torch.zeros_like(p)).reshape(-1)
                          for p, gp in zip(params, go[1:])])
        np.testing.assert_allclose(gi[1].numpy(), flat.numpy(), rtol=1e-8, atol=1e-9)


def test_flops_per_eval_match_survey_table():
    """algorithmic flops per scene evaluation."""
    assert T.Tape(S.make_test_scene2()).flops_per_eval == 64
    assert T.Tape(S.make_test_scene()).flops_per_eval == 181
    assert T.Tape(S.make_synthetic_scene(64, 0, False)).flops_per_eval == 3382
    assert T.Tape(S.make_synthetic_scene(64, 0, True)).flops_per_eval == 3404
    assert T.Tape(S.make_synthetic_scene(512, 0, True)).flops_per_eval == 27070


def test_tape_structure_and_errors():
    t = T.Tape(S.make_test_scene2())
    assert t.n_nodes == 1 and t.max_depth == 1 and t.n_leaves == 4   # inner hard union is flattened
    assert t.code.dtype == np.int32 and t.code.shape[1] == 2 and t.n_par % 4 == 0
    dec = [T.decode(w0, w1) for w0, w1 in t.code.tolist()]
    assert [d[0] for d in dec] == [T.OP_OPEN_MIN, T.OP_BOX, T.OP_SPHERE, T.OP_TORUS, T.OP_LINE,
                                   T.OP_CLOSE_MIN]
    assert [d[5] for d in dec[1:5]] == [0, 1, 2, 3]        # child ordinals of the merged union
    assert dec[0][3] and dec[-1][3]                        # root flag on the outermost OPEN/CLOSE
    assert dec[1][4] == [T.POST_ONION] and not dec[1][1]   # onion post-op on the box, world frame
    a = T.Tape(S.make_test_scene())
    assert sum(T.decode(*w)[1] for w in a.code.tolist()) == 3     # three leaves carry a frame record
    # parameter order follows module.parameters(), so optimisers / state_dict carry over
    assert [n for n, _ in S.make_test_scene().named_parameters()][:3] == [
        "blend_k", "sdfs.0.translation", "sdfs.0.orientation"]
    with pytest.raises(NotImplementedError):
        T.Tape(torch.nn.Linear(3, 1))
    deep = S.SDFSphere(1.0)
    for _ in range(4):
        deep = S.SDFOnion(deep, 0.1)
    with pytest.raises(NotImplementedError):
        T.Tape(deep)
    nest = S.SDFSphere(1.0)
    for _ in range(9):
        nest = S.SDFSmoothUnion([nest, S.SDFPlane()], 4.0)
    with pytest.raises(NotImplementedError):
        T.Tape(nest)


def test_lowered_params_cache_tracks_parameter_updates():
    scene = S.make_test_scene()
    tape = T.tape_of(scene)
    with torch.no_grad():
        a = tape.lowered_params()
        assert tape.lowered_params() is a
        scene.blend_k.add_(1.0)
        b = tape.lowered_params()
    assert b is not a and not torch.equal(a, b)
    assert tape.lowered_params().requires_grad      # autograd on: fresh, connected to parameters


def test_lowered_params_cache_sees_parameter_writes():
    """the no-grad cache must not serve a stale block.  Version-bumping writes and a
    replaced ``.data`` are detected; in-place writes THROUGH ``.data`` need ``Tape.invalidate()``."""
    scene = S.make_test_scene()
    tape = T.tape_of(scene)
    k = scene.blend_k
    with torch.no_grad():
        a = tape.lowered_params()
        assert tape.lowered_params() is a                         # cached
        k.mul_(2.0)                                               # bumps the version counter
        b = tape.lowered_params()
        assert b is not a and not torch.equal(a, b)
        k.data = torch.tensor(5.0)                                # new storage, same version
        c = tape.lowered_params()
        assert c is not b and float(-c[tape._perm.tolist().index(1 + tape.n_theta + tape._slot(k))]) == 5.0
        k.data.fill_(7.0)                                         # invisible to autograd's counters
        tape.invalidate()
        d = tape.lowered_params()
        assert d is not c and float(-d[tape._perm.tolist().index(1 + tape.n_theta + tape._slot(k))]) == 7.0


def test_sphere_table_bounds_hold_and_need_no_lapack():
    """tape.py: the culled union's sphere table comes from closed-form 3x3 singular values / inverses (plain
    elementwise torch: the lowering runs inside captured CUDA graphs).  The bounds must bracket LAPACK's values,
    tightly for the near-isometric frames scenes use; every child's surface must lie inside its sphere; and the
    table must be there while autograd records (the loss kernel's reverse sweep reads it)."""
    g = torch.Generator().manual_seed(0)
    A = torch.randn(500, 3, 3, generator=g)
    A[:200] = torch.linalg.qr(torch.randn(200, 3, 3, generator=g))[0] * (0.5 + torch.rand(200, 1, 1, generator=g))
    hi, lo = T._singular_range_3x3(A)
    sv = torch.linalg.svdvals(A.double()).float()
    assert (hi >= sv[:, 0]).all() and (lo <= sv[:, 2]).all()
    assert (hi[:200] / sv[:200, 0]).max() < 1.002 and (lo[:200] / sv[:200, 2]).min() > 0.998
    assert ((T._inverse_3x3(A[:200]) @ A[:200] - torch.eye(3)).abs().max() < 1e-4)

    scene = S.make_synthetic_scene(16, 0, True)
    tape = T.tape_of(scene)
    assert tape._cull_union is not None
    par_grad = tape.lowered_params("cpu")                      # autograd recording: table present
    with torch.no_grad():
        par = tape.lowered_params("cpu")
    assert par_grad.requires_grad and torch.equal(par_grad.detach(), par)
    code = tape.code
    opens = [i for i, (w0, w1) in enumerate(code.tolist()) if T.decode(w0, w1)[0] == T.OP_OPEN_SMOOTH and T.decode(w0, w1)[8]]
    assert len(opens) == 1
    o = opens[0]
    table_f4 = T.decode(*code[o].tolist())[5]                  # ordinal field of the OPEN = table offset (float4)
    union_rec = par[T.decode(*code[o].tolist())[6] * 4: T.decode(*code[o].tolist())[6] * 4 + 4]
    assert union_rec[0] < 0 and 0 < union_rec[1] <= 1.0 and union_rec[2] >= 1.0 and union_rec[3] == 4.0   # -k, s_min, L_max, log2 16
    # every leaf child: sample its zero set through the oracle-free route -- points at which the child's own
    # value is ~0 must lie inside its sphere.  Use the children's evaluation on random points near the sphere.
    children = list(scene.sdfs[1].sdfs) if hasattr(scene.sdfs[1], "sdfs") else []
    assert len(children) == 16
    pts = (torch.rand(20000, 3, generator=g) - 0.5) * 8.0
    for j, child in enumerate(children):
        cx, cy, cz, r = par[(table_f4 + j) * 4: (table_f4 + j) * 4 + 4].tolist()
        with torch.no_grad():
            f = _cpu_value(child, pts)
        inside = f <= 0
        if inside.any():
            d = (pts[inside] - torch.tensor([cx, cy, cz])).norm(dim=-1)
            assert d.max().item() <= r * 1.0001, (j, d.max().item(), r)


def _cpu_value(node, pts):
    """Value of one scene node at CPU points through the oracle's restatement of the reference classes."""
    return O.sdf(node, pts).reshape(-1)
