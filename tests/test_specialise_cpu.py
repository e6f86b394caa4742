"""specialise.py on the CPU side: the generated straight-line source mirrors the tape, the structure
hash is stable, and (nvcc present) the specialised translation unit compiles for sm_100a."""
import os
import re

import numpy as np
import pytest

from ray_marching_b200 import scene as S, specialise as SP
from ray_marching_b200.scene.scene_registry import make_test_scene, make_test_scene2
from ray_marching_b200.scene.synthetic import make_synthetic_scene
from ray_marching_b200.tape import tape_of


def nested_scene():
    """Two sibling unions (smooth with a rounded onion-layered box, hard) under a hard root with a
    rounded close: every generator branch (frames, post-op chains, nested folds)."""
    a = S.SDFSmoothUnion([S.SDFAffineTransformation(S.SDFSphere(0.3), (0.9, 0.1, 0.2, 0.3), (0.2, 0.0, 0.1)),
                          S.SDFRounding(S.SDFOnion(S.SDFBox((0.2, 0.1, 0.3)), 0.05), 0.01),
                          S.SDFAffineTransformation(S.SDFPlane(), (1.0, 0.0, 0.0, 0.0), (-2.0, 0.0, 0.0))], blend_k=9.0)
    b = S.SDFOnion(S.SDFSmoothUnion([S.SDFTorus(0.4, 0.1), S.SDFDisk(0.3),
                                     S.SDFLine((-0.2, 0.1, 0.0), (0.3, 0.2, 0.1), 0.05)], blend_k=15.0), 0.02)
    return S.SDFRounding(S.SDFUnion([a, b, S.SDFSphere(0.1)]), 0.005)


def test_generated_source_mirrors_the_tape():
    t = tape_of(make_test_scene2())
    src = SP.generate_source(t)
    assert f"0x{SP.code_hash(t.code):08x}u" in src
    for prim in ("prim_box<FAST>(P[0]", "prim_sphere<FAST>(P[2]", "prim_torus<FAST>(P[3]",
                 "prim_line<FAST>(P[4], P[5], P[6]"):
        assert prim in src
    # 4 children, the first one assigned (min(+inf, d) = d), one onion
    assert src.count("min_nan") == 3 and "a0_0[k] = d[k];" in src and "fabsf(d[k]) - pp.x" in src
    t2 = tape_of(nested_scene())
    src2 = SP.generate_source(t2)
    assert len(re.findall(r"M::exp_scale", src2)) == 2
    assert src2.count("a1_1[k] = 1.f;") == 1 and src2.count("a1_2[k] = 1.f;") == 1   # log-sum-exp starts at (x, 1)
    assert "CUDART_INF_F" not in src2                                           # no accumulator is initialised
    assert SP.code_hash(t2.code) != SP.code_hash(t.code)
    assert SP.code_hash(tape_of(make_test_scene2()).code) == SP.code_hash(t.code)   # structure only


def test_large_tapes_are_refused():
    with pytest.raises(ValueError):
        SP.build(tape_of(make_synthetic_scene(64)))


@pytest.mark.skipif(not os.path.exists(SP.NVCC), reason="nvcc not available")
def test_specialised_translation_unit_compiles(tmp_path, monkeypatch):
    monkeypatch.setattr(SP, "SPEC_DIR", str(tmp_path))
    lib = SP.build(tape_of(nested_scene()))
    assert os.path.exists(lib)
    out = os.popen(f"nm -D --defined-only {lib}").read()
    assert " rm_render" in out and " rm_abi_version" in out and " rm_last_error" in out
    for name in SP.SPEC_EXPORTS:                                             # render + the backward entry points
        assert f" {name}\n" in out, name
    assert " rm_march\n" not in out and " rm_sdf_eval" not in out          # no stand-alone forward ops
    assert SP.build(tape_of(nested_scene())) == lib                          # cached


def test_selection_is_per_tape_not_per_structure():
    """two scenes of the same structure choose their kernels independently, and the
    entry dies with the tape."""
    import gc
    a, b = make_test_scene2(), make_test_scene2()
    ta, tb = tape_of(a), tape_of(b)
    handle = object()
    SP.select(ta, handle)
    assert SP.selected_for(ta.code_tensor()) is handle
    assert SP.selected_for(tb.code_tensor()) is None          # same code bytes, other tape
    SP.select(ta, None)
    assert SP.selected_for(ta.code_tensor()) is None
    SP.select(tb, handle)
    n = len(SP._selected)
    del b, tb
    gc.collect()
    assert len(SP._selected) == n - 1
