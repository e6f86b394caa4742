"""Import the UNMODIFIED reference from /root/reference.  TEST INFRASTRUCTURE ONLY.

Works only in the build container (the GPU box has no /root/reference).  Used by
``oracle/make_golden.py`` and by the optional ``tests/test_oracle_vs_reference.py``.
``control.py`` cannot be imported (pyautogui / pynput are absent), so the two pure
functions needed from it (``RenderLoop``, ``make_reflection_directions``,
control.py:179-258) are compiled from their own source text, unmodified, at run time.
"""
from __future__ import annotations

import ast
import contextlib
import io
import os
import sys
import types

REF = "/root/reference"


def available() -> bool:
    return os.path.isdir(os.path.join(REF, "rendering"))


@contextlib.contextmanager
def _cwd(path):
    old = os.getcwd()
    os.chdir(path)
    try:
        yield
    finally:
        os.chdir(old)


def load():
    """Returns a namespace with the reference modules/classes (cwd-sensitive cmap load handled)."""
    sys.dont_write_bytecode = True
    if REF not in sys.path:
        sys.path.insert(0, REF)
    with _cwd(REF), contextlib.redirect_stdout(io.StringIO()):
        import quaternion as Q                               # noqa
        from rendering import ray_marching as RM             # noqa
        from rendering import shader as SH                   # noqa
        from scene import primitives as P, transformations as T, scene_registry as SR  # noqa
        shader = SH.Shader().float()
    ns = types.SimpleNamespace(Q=Q, RM=RM, SH=SH, P=P, T=T, SR=SR, shader=shader)
    for mod in (P, T):
        for name in dir(mod):
            if name.startswith("SDF"):
                setattr(ns, name, getattr(mod, name))
    # control.py:179-258, executed from the reference's own source text
    src = open(os.path.join(REF, "control.py")).read()
    tree = ast.parse(src)
    keep = [n for n in tree.body
            if isinstance(n, (ast.FunctionDef, ast.ClassDef))
            and n.name in ("make_reflection_directions", "RenderLoop")]
    glb = {"nn": __import__("torch").nn, "Tensor": __import__("torch").Tensor,
           "PinholeCamera": RM.PinholeCamera, "SDFMarcher": RM.SDFMarcher,
           "SDFNormals": RM.SDFNormals, "Shader": SH.Shader, "torch": __import__("torch")}
    exec(compile(ast.Module(body=keep, type_ignores=[]), "control.py", "exec"), glb)
    ns.make_reflection_directions = glb["make_reflection_directions"]
    ref_loop_cls = glb["RenderLoop"]

    def make_render_loop(*a, **k):
        with _cwd(REF), contextlib.redirect_stdout(io.StringIO()):
            return ref_loop_cls(*a, **k).float()
    ns.RenderLoop = make_render_loop

    # control.py:148-165: the two state-update assignments of EventAggregator.get_state, executed
    # from the reference's own source text against a stand-in `self`
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "EventAggregator")
    fn = next(n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name == "get_state")
    upd = [n for n in fn.body if isinstance(n, ast.Assign) and isinstance(n.targets[0], ast.Attribute)
           and n.targets[0].attr in ("position", "orientation")
           and isinstance(n.targets[0].value, ast.Name) and n.targets[0].value.id == "self"]
    assert [n.targets[0].attr for n in upd] == ["position", "orientation"], "control.py:148-165 moved"
    code = compile(ast.Module(body=upd, type_ignores=[]), "control.py", "exec")
    torch = __import__("torch")

    def camera_state_update(state, translation_input, orientation_input):
        """state: namespace with position (N,3), orientation (N,4), translation_sensitivity,
        rotation_sensitivity; updated in place like EventAggregator's attributes."""
        exec(code, {"Q": Q, "F": torch.nn.functional, "self": state,
                    "translation_input": translation_input, "orientation_input": orientation_input})
    ns.camera_state_update = camera_state_update
    return ns
