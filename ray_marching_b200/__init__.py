"""B200-native differentiable sphere tracing behind the kyle-rosa/ray_marching module API.

Public surface mirrors the reference: ``quaternion``, ``scene`` (primitives, transformations,
scene_registry), ``rendering`` (PinholeCamera, SDFMarcher, SDFNormals, Shader) and
``control.RenderLoop``.  All evaluation happens in hand-written sm_100a kernels loaded from
``csrc/librm_b200.so`` through the C ABI in ``include/rm_b200.h``; there is no CPU path.
"""
from . import quaternion, scene, tape  # noqa: F401
from .control import RenderLoop, make_reflection_directions  # noqa: F401
from .frames import CameraPath, FrameLoop, GraphedStep, write_png, write_npy  # noqa: F401
from .rendering.ray_marching import PinholeCamera, SDFMarcher, SDFNormals  # noqa: F401
from .rendering.shader import Shader  # noqa: F401

__all__ = ["quaternion", "scene", "tape", "RenderLoop", "make_reflection_directions",
           "PinholeCamera", "SDFMarcher", "SDFNormals", "Shader", "CameraPath", "FrameLoop", "write_png",
           "write_npy"]
