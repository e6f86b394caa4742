"""Common base of every drop-in SDF scene node.

Each node keeps the reference's constructor arguments and ``nn.Parameter`` names
(so ``state_dict()`` keys such as ``sdfs.0.sdf.sdf.halfsides`` are unchanged), but
its ``forward`` does not run torch arithmetic: the sub-tree rooted at the node is
lowered once to an instruction tape (``ray_marching_b200.tape``) and evaluated by
the sm_100a tape interpreter through ``torch.ops.rm_b200.sdf_eval``.

Reference call contract (``scene/primitives.py:11``, ``scene/transformations.py:33``):
``node(query_positions[..., 3]) -> signed_distance[..., 1]``.
"""
from __future__ import annotations

import torch
import torch.nn as nn
from torch import Tensor


class SDFNode(nn.Module):
    #: filled by subclasses: name understood by ``tape.lower`` (== reference class name)
    kind: str = ""

    def _param(self, value, name: str) -> None:
        self.register_parameter(
            name, nn.Parameter(torch.tensor(value, dtype=torch.get_default_dtype()))
        )

    def tape(self):
        """Lowered instruction tape of the sub-tree rooted here (cached per module)."""
        from .. import tape as _tape

        return _tape.tape_of(self)

    def forward(self, query_positions: Tensor) -> Tensor:
        from .. import ops

        return ops.sdf_eval(self.tape(), query_positions)
