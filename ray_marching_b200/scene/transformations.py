"""Drop-in SDF combinators (reference: ``scene/transformations.py``).

Constructor signatures, child attribute names (``sdf`` / ``sdfs``) and parameter
names match the reference so that module trees, ``state_dict`` keys and optimisers
carry over.  Evaluation goes through the lowered tape, see :class:`SDFNode`.
"""
from __future__ import annotations

import torch.nn as nn
from torch import Tensor

from ._node import SDFNode


class _CoordsNode(SDFNode):
    """The four value combinators name their ``forward`` argument ``query_coords`` in the reference
    (``transformations.py:67,90,117,131``); keyword callers must keep working."""

    def forward(self, query_coords: Tensor) -> Tensor:     # noqa: D102 - see SDFNode.forward
        return SDFNode.forward(self, query_coords)


class SDFAffineTransformation(SDFNode):
    """``child(rot(p - translation, conj(orientation)))`` (``transformations.py:8-42``).

    The orientation is deliberately *not* normalised (the reference never does), so
    the lowered frame matrix is the same polynomial in ``q`` as ``quaternion.rotation``.
    """
    kind = "SDFAffineTransformation"

    def __init__(self, sdf: nn.Module, orientation, translation):
        super().__init__()
        self.sdf = sdf
        self._param(tuple(translation), "translation")
        self._param(tuple(orientation), "orientation")


class SDFSmoothUnion(_CoordsNode):
    """``-(1/k) logsumexp_i(-k f_i)`` over the children (``transformations.py:45-71``)."""
    kind = "SDFSmoothUnion"

    def __init__(self, sdfs, blend_k: float):
        super().__init__()
        self.sdfs = nn.ModuleList(sdfs)
        self._param(blend_k, "blend_k")


class SDFUnion(_CoordsNode):
    """``min_i f_i`` over the children (``transformations.py:74-94``)."""
    kind = "SDFUnion"

    def __init__(self, sdfs):
        super().__init__()
        self.sdfs = nn.ModuleList(sdfs)


class SDFRounding(_CoordsNode):
    """``child(p) - rounding`` (``transformations.py:97-118``)."""
    kind = "SDFRounding"

    def __init__(self, sdf: nn.Module, rounding: float):
        super().__init__()
        self.sdf = sdf
        self._param(rounding, "rounding")


class SDFOnion(_CoordsNode):
    """``|child(p)| - radius`` (``transformations.py:121-132``)."""
    kind = "SDFOnion"

    def __init__(self, sdf: nn.Module, radius: float):
        super().__init__()
        self.sdf = sdf
        self._param(radius, "radius")
