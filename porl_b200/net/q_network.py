"""QNetwork parameter container (src/porl/net/q_network.py:8-30): S -> 64 -> 128 -> 64 -> A, ReLU."""
from __future__ import annotations

import torch
from torch import nn

from .. import _C
from ..nets import FlatMLP, InferenceMixin, MLPParams


class QNetwork(nn.Module, InferenceMixin):
    """state_dict keys `model.{0,2,4,6}.{weight,bias}` as the reference's nn.Sequential.

    `flat` / `grads`: slices of the owning optimizer group's buffers (None -> private storage)."""

    def __init__(self, state_size, action_size, hidden_sizes=(64, 128, 64), *, device="cuda",
                 flat: torch.Tensor | None = None, grads: torch.Tensor | None = None, trainable: bool = True,
                 init: bool = True):
        super().__init__()
        device = torch.device(device)
        ctx = _C.context(device)
        self.plan = FlatMLP(ctx, [state_size, *hidden_sizes, action_size], n_nets=1)
        if flat is None:
            flat = torch.zeros(self.plan.n_floats, dtype=torch.float32, device=device)
        self.flat = flat
        self.model = MLPParams(self.plan, flat, 0, grads, trainable, init)

    def forward(self, x):
        return self._infer(self.plan, self.flat, x)[0]
