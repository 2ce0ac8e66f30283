"""Gaussian policy heads (agent/policy.py:12-59) over one flat HBM buffer."""
from __future__ import annotations

import torch
from torch import nn
from torch.distributions import MultivariateNormal

from .. import _C
from ..nets import FlatMLP, InferenceMixin, MLPParams, pad_to

LOG_STD_MIN = -5.0  # agent/policy.py:8
LOG_STD_MAX = 2.0   # agent/policy.py:9


def policy_distribution(mean: torch.Tensor, log_std: torch.Tensor) -> MultivariateNormal:
    """What `policy(obs)` returns: the reference's MultivariateNormal(mean, scale_tril=diag(std)) with
    std = exp(clamp(log_std, -5, 2)) (agent/policy.py:18-23, :42-49), so `.mean`, `.log_prob`, `.sample`,
    `.rsample`, `.entropy`, `.covariance_matrix` behave as there.  A convenience object for callers of the module
    API (select_action, evaluation); the update path evaluates the same log-likelihood inside porl_awr_gauss_nll."""
    std = torch.exp(log_std.clamp(LOG_STD_MIN, LOG_STD_MAX))
    return MultivariateNormal(mean, scale_tril=torch.diag(std))


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
        return policy_distribution(mean, self.log_std.detach())

    def act(self, obs, deterministic: bool = False, enable_grad: bool = False):
        dist = self(obs)
        return dist.mean if deterministic else dist.sample()


class BoundedGaussianPolicy(GaussianPolicy):
    """Same with Tanh on the mean (agent/policy.py:35-49); the reference's `|mean| > 1` pdb trap
    cannot fire after tanh and is dropped."""

    _tanh = True
