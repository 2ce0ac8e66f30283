"""Oracle network containers (eager torch, CPU, fp32 or fp64).  Test infrastructure.

State-dict key layout is the contract, so the module/attribute names below are
the reference's: Sequential indices 0,2,4 (0,3,6 + LayerNorm 1,4 with layer_norm),
`v1`/`v2`, `net`/`log_std`, `model`.

Follows: util/util.py:20-47 (Squeeze, mlp), agent/value_functions.py:31-42 (TwinV),
agent/policy.py:8-49 (GaussianPolicy / BoundedGaussianPolicy, clamp range [-5, 2]),
src/porl/net/q_network.py:8-30 (QNetwork, hidden 64/128/64).
"""
from __future__ import annotations

import torch
from torch import nn
from torch.distributions import MultivariateNormal

LOG_STD_RANGE = (-5.0, 2.0)  # agent/policy.py:8-9


class _DropLast(nn.Module):
    """x[..., 1] -> x[...]  (util/util.py:20-26)."""

    def forward(self, x):
        return x.squeeze(dim=-1)


def build_mlp(widths, *, layer_norm=False, tanh_out=False, scalar_out=False):
    """Linear[+LayerNorm]+ReLU per hidden layer, then Linear[+Tanh][+squeeze]  (util/util.py:29-47)."""
    if len(widths) < 2:
        raise ValueError("an MLP needs an input and an output width")
    seq = []
    for fan_in, fan_out in zip(widths[:-2], widths[1:-1]):
        seq.append(nn.Linear(fan_in, fan_out))
        if layer_norm:
            seq.append(nn.LayerNorm(fan_out))
        seq.append(nn.ReLU())
    seq.append(nn.Linear(widths[-2], widths[-1]))
    if tanh_out:
        seq.append(nn.Tanh())
    if scalar_out:
        if widths[-1] != 1:
            raise ValueError("scalar_out needs a width-1 head")
        seq.append(_DropLast())
    return nn.Sequential(*seq).to(torch.float32)


class TwinV(nn.Module):
    """Two independent scalar-headed MLPs; forward = elementwise min  (value_functions.py:31-42)."""

    def __init__(self, in_dim, hidden, n_hidden, layer_norm=False):
        super().__init__()
        widths = [in_dim] + [hidden] * n_hidden + [1]
        self.v1 = build_mlp(widths, layer_norm=layer_norm, scalar_out=True)
        self.v2 = build_mlp(widths, layer_norm=layer_norm, scalar_out=True)

    def both(self, x):
        return self.v1(x), self.v2(x)

    def forward(self, x):
        a, b = self.both(x)
        return torch.min(a, b)


class GaussianHead(nn.Module):
    """MLP mean + state-independent clamped log_std -> MultivariateNormal(diag)  (policy.py:12-49).

    bounded=True puts Tanh on the mean (BoundedGaussianPolicy, policy.py:35-49); the
    reference's pdb trap on |mean|>1 cannot fire after tanh and is not restated.
    """

    def __init__(self, in_dim, out_dim, hidden, n_hidden, bounded=False):
        super().__init__()
        self.net = build_mlp([in_dim] + [hidden] * n_hidden + [out_dim], tanh_out=bounded)
        self.log_std = nn.Parameter(torch.zeros(out_dim, dtype=torch.float32))

    def forward(self, x):
        mu = self.net(x)
        sigma = self.log_std.clamp(*LOG_STD_RANGE).exp()
        return MultivariateNormal(mu, scale_tril=torch.diag(sigma))


class QNet(nn.Module):
    """S -> 64 -> 128 -> 64 -> A, ReLU  (src/porl/net/q_network.py:8-30)."""

    def __init__(self, state_size, action_size, hidden_sizes=(64, 128, 64)):
        super().__init__()
        seq, width = [], state_size
        for h in hidden_sizes:
            seq += [nn.Linear(width, h), nn.ReLU()]
            width = h
        seq.append(nn.Linear(width, action_size))
        self.model = nn.Sequential(*seq)

    def forward(self, x):
        return self.model(x)
