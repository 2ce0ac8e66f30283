"""Gaussian policy heads (agent/policy.py:12-59) over one flat HBM buffer."""
from __future__ import annotations

import math

import torch
from torch import nn

from .. import _C
from ..nets import FlatMLP, InferenceMixin, MLPParams, pad_to

LOG_STD_MIN = -5.0  # agent/policy.py:8
LOG_STD_MAX = 2.0   # agent/policy.py:9


class DiagGaussian:
    """What `policy(obs)` returns: the reference builds MultivariateNormal(mean, diag(std))
    (agent/policy.py:18-23); only `.mean` (select_action, agent/sorl.py:76) and `.log_prob`
    are used by its callers.  Convenience object, not on the update path."""

    def __init__(self, mean: torch.Tensor, log_std: torch.Tensor):
        self.mean = mean
        self.log_std = log_std.clamp(LOG_STD_MIN, LOG_STD_MAX)
        self.stddev = self.log_std.exp()

    def log_prob(self, value: torch.Tensor) -> torch.Tensor:
        z = (value - self.mean) / self.stddev
        d = self.mean.shape[-1]
        return -0.5 * (z * z).sum(-1) - self.log_std.sum() - 0.5 * d * math.log(2.0 * math.pi)

    def sample(self) -> torch.Tensor:
        return self.mean + self.stddev * torch.randn_like(self.mean)


class GaussianPolicy(nn.Module, InferenceMixin):
    """MLP mean [obs, H.., act] + state-independent log_std clamped to [-5, 2]."""

    _tanh = False

    def __init__(self, obs_dim: int, act_dim: int, hidden_dim: int = 256, n_hidden: int = 2, *, ctx: _C.Context,
                 device, flat: torch.Tensor | None = None, grads: torch.Tensor | None = None, init: bool = True):
        super().__init__()
        dims = [obs_dim, *([hidden_dim] * n_hidden), act_dim]
        self.plan = FlatMLP(ctx, dims, n_nets=1, tanh_out=self._tanh)
        self.act_dim = act_dim
        n = self.group_floats(self.plan, act_dim)
        if flat is None:
            flat = torch.zeros(n, dtype=torch.float32, device=device)
        self.flat = flat
        self.net = MLPParams(self.plan, flat, 0, grads, True, init)
        off = self.plan.n_floats
        self.log_std = nn.Parameter(flat[off:off + act_dim])          # zeros, agent/policy.py:16
        if grads is not None:
            self.log_std.grad = grads[off:off + act_dim]

    @staticmethod
    def group_floats(plan: FlatMLP, act_dim: int) -> int:
        """flat layout of the optimizer group: [mlp parameters | log_std (padded to 4)]"""
        return plan.n_floats + pad_to(act_dim, 4)

    @property
    def log_std_offset(self) -> int:
        return self.plan.n_floats

    def forward(self, obs):
        mean = self._infer(self.plan, self.flat, obs)[0]
        return DiagGaussian(mean, self.log_std.detach())

    def act(self, obs, deterministic: bool = False, enable_grad: bool = False):
        dist = self(obs)
        return dist.mean if deterministic else dist.sample()


class BoundedGaussianPolicy(GaussianPolicy):
    """Same with Tanh on the mean (agent/policy.py:35-49); the reference's `|mean| > 1` pdb trap
    cannot fire after tanh and is dropped."""

    _tanh = True
