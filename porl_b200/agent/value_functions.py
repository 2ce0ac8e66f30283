"""TwinV value heads (agent/value_functions.py:31-42) over one flat HBM buffer."""
from __future__ import annotations

import torch
from torch import nn

from .. import _C
from ..nets import FlatMLP, InferenceMixin, MLPParams


class TwinV(nn.Module, InferenceMixin):
    """Two independent MLPs [in, H.., 1]; `both()` -> (v1, v2); `forward()` -> elementwise min.

    The two nets live side by side in `flat` and run as one grouped launch per layer.  `flat` /
    `grads` are slices of the owning optimizer group's buffers (None -> private storage, used for
    the target network: deepcopy(vf).requires_grad_(False), agent/por.py:59).
    """

    def __init__(self, state_dim: int, layer_norm: bool = False, hidden_dim: int = 256, n_hidden: int = 2, *,
                 ctx: _C.Context, device, flat: torch.Tensor | None = None, grads: torch.Tensor | None = None,
                 trainable: bool = True, init: bool = True):
        super().__init__()
        dims = [state_dim, *([hidden_dim] * n_hidden), 1]
        self.plan = FlatMLP(ctx, dims, n_nets=2, layer_norm=layer_norm)
        if flat is None:
            flat = torch.zeros(self.plan.n_floats, dtype=torch.float32, device=device)
        self.flat = flat
        self.v1 = MLPParams(self.plan, flat, 0, grads, trainable, init)
        self.v2 = MLPParams(self.plan, flat, 1, grads, trainable, init)

    def both(self, state):
        out = self._infer(self.plan, self.flat, state)
        return out[0, :, 0], out[1, :, 0]

    def forward(self, state):
        v1, v2 = self.both(state)
        return torch.minimum(v1, v2)
