"""Drop-in leaf primitives (reference: ``scene/primitives.py``).

Same class names, constructor arguments and parameter names as the reference; the
distance formulas themselves live in the CUDA tape interpreter
(``csrc/rm_prims.cuh``) and are cited there.  ``forward`` is inherited from
:class:`SDFNode` and evaluates the lowered tape on the GPU.
"""
from __future__ import annotations

from ._node import SDFNode


class SDFSphere(SDFNode):
    """``|p| - radius``  (reference ``primitives.py:6-17``)."""
    kind = "SDFSphere"

    def __init__(self, radius: float):
        super().__init__()
        self._param(radius, "radius")


class SDFBox(SDFNode):
    """Axis-aligned box with the given half side lengths (``primitives.py:20-33``)."""
    kind = "SDFBox"

    def __init__(self, halfsides: tuple[float, float, float]):
        super().__init__()
        self._param(tuple(halfsides), "halfsides")


class SDFPlane(SDFNode):
    """Half-space ``x < 0``: distance is ``p.x`` (``primitives.py:36-41``)."""
    kind = "SDFPlane"

    def __init__(self):
        super().__init__()


class SDFLine(SDFNode):
    """Capsule around the segment ``start -> end`` (``primitives.py:44-61``)."""
    kind = "SDFLine"

    def __init__(self, start, end, radius: float):
        super().__init__()
        self._param(tuple(start), "start")
        self._param(tuple(end), "end")
        self._param(radius, "radius")


class SDFDisk(SDFNode):
    """Flat disk of the given radius in the yz-plane (``primitives.py:64-82``)."""
    kind = "SDFDisk"

    def __init__(self, radius: float):
        super().__init__()
        self._param(radius, "radius")


class SDFTorus(SDFNode):
    """Torus around the y axis: ring ``radius1``, tube ``radius2`` (``primitives.py:85-102``)."""
    kind = "SDFTorus"

    def __init__(self, radius1: float, radius2: float) -> None:
        super().__init__()
        self._param(radius1, "radius1")
        self._param(radius2, "radius2")
