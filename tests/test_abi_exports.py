"""CPU-only: the C-ABI shared library builds, loads, and exports every symbol include/rm_b200.h declares."""
import ctypes
import os
import re

from ray_marching_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "rm_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rm_[a-z_0-9]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound():
    names = _declared()
    assert "rm_render" in names and "rm_sdf_eval" in names
    handle = _lib.lib()
    for n in names:
        assert hasattr(handle, n), f"{n} declared in rm_b200.h but not exported"
    assert sorted(_lib.exported_symbols()) == names, "ctypes signature table out of sync with header"
    assert handle.rm_abi_version() == _lib.ABI_VERSION


def test_struct_layout_matches_header():
    assert ctypes.sizeof(_lib.RmTape) == 32
    assert ctypes.sizeof(_lib.RmCamera) == 24
    assert ctypes.sizeof(_lib.RmRenderOut) == 16 * 8
    assert ctypes.sizeof(_lib.RmRenderArgs) == 72
    assert ctypes.sizeof(_lib.RmRenderGrads) == 48


def test_argument_errors_do_not_need_a_gpu():
    handle = _lib.lib()
    tape = _lib.RmTape(0, 0, 0, 0, 0, 0)
    assert handle.rm_sdf_eval(ctypes.byref(tape), 0, 4, 0, 0) == -1
    assert b"tape" in handle.rm_last_error()


def test_cpu_tensors_are_rejected_loudly():
    import pytest
    import torch
    from ray_marching_b200 import scene as S
    with pytest.raises(_lib.RmError):
        S.SDFSphere(0.5)(torch.zeros(4, 3))
