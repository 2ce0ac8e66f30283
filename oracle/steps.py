"""Oracle update steps: POR, SORL, discrete CQL.  Test infrastructure (eager CPU torch).

Each class keeps the reference's attribute names that appear in `state_dict()` so
checkpoints interchange with the reference and with the CUDA product.

Follows:
  POR   agent/por.py:21-71 (ctor), :73-112 (por_residual_update), :15-17 (expectile loss)
  SORL  agent/sorl.py:21-69 (ctor), :71-76 (select_action), :78-128 (update), :130-152 (vf_update)
  CQL   src/porl/train/cql_trainer.py:60-124 (penalty + learn),
        src/porl/train/dqn_trainer.py:66-78 (nets, Adam lr 5e-4, batch 64), :182-204 (train_offline)
  EMA   util/util.py:54-56
"""
from __future__ import annotations

import copy
import math
from types import SimpleNamespace

import numpy as np
import torch
from torch import nn
from torch.optim.lr_scheduler import CosineAnnealingLR

from .nets import GaussianHead, QNet, TwinV

WEIGHT_CAP = 100.0  # EXP_ADV_MAX, agent/por.py:12, agent/sorl.py:12


def expectile_l2(u, tau):
    """mean(|tau - 1[u<0]| * u^2)  (por.py:15-17)."""
    return (torch.abs(tau - (u < 0).to(u.dtype)) * u.square()).mean()


def polyak_(target, online, beta):
    """target <- (1-beta) target + beta online, per parameter  (util/util.py:54-56)."""
    with torch.no_grad():
        for t, s in zip(target.parameters(), online.parameters()):
            t.mul_(1.0 - beta).add_(s, alpha=beta)


def dims(state_size, hidden_dim, n_hidden=2, layer_norm=False, feature_dim=256, action_size=2):
    return SimpleNamespace(state_size=state_size, hidden_dim=hidden_dim, n_hidden=n_hidden,
                           layer_norm=layer_norm, feature_dim=feature_dim, action_size=action_size)


class _ExpectileAgent(nn.Module):
    """Shared value step of POR and SORL (por.py:75-93 == sorl.py:81-98)."""

    _online = _target = ""  # attribute names of the online / target TwinV

    def _value_step(self, feat, next_feat, rewards, terminals):
        online, target = getattr(self, self._online), getattr(self, self._target)
        with torch.no_grad():
            next_v = target(next_feat)
        target_v = rewards + (1.0 - terminals.to(rewards.dtype)) * self.discount * next_v
        v1, v2 = online.both(feat)
        v_loss = (expectile_l2(target_v - v1, self.tau) + expectile_l2(target_v - v2, self.tau)) / 2
        self.v_optimizer.zero_grad(set_to_none=True)
        v_loss.backward()
        self.v_optimizer.step()
        polyak_(target, online, self.beta)
        return target_v, v_loss


class OraclePOR(_ExpectileAgent):
    _online, _target = "vf", "v_target"

    def __init__(self, args, max_steps, tau, alpha, backbone=None, value_lr=1e-4, policy_lr=1e-4,
                 discount=0.99, beta=0.005):
        super().__init__()
        self.backbone = backbone
        width = args.state_size if backbone is None else args.feature_dim
        self.goal_policy = GaussianHead(width, args.state_size, args.hidden_dim, args.n_hidden)
        self.vf = TwinV(width, args.hidden_dim, args.n_hidden, layer_norm=args.layer_norm)
        self.v_target = copy.deepcopy(self.vf).requires_grad_(False)
        self.v_optimizer = torch.optim.Adam(self.vf.parameters(), lr=value_lr)
        self.goal_policy_optimizer = torch.optim.Adam(self.goal_policy.parameters(), lr=policy_lr)
        self.goal_lr_schedule = CosineAnnealingLR(self.goal_policy_optimizer, max_steps)
        self.tau, self.alpha, self.discount, self.beta = tau, alpha, discount, beta

    def por_residual_update(self, observations, next_observations, rewards, terminals):
        feat, next_feat = observations, next_observations
        if self.backbone is not None:
            feat, next_feat = self.backbone(observations), self.backbone(next_observations)
        target_v, v_loss = self._value_step(feat, next_feat, rewards, terminals)
        # goal policy: advantage-weighted likelihood of the RAW next state (por.py:97-110)
        adv = target_v - self.vf(feat.detach())
        weight = torch.exp(adv / self.alpha).clamp_max(WEIGHT_CAP).detach()
        nll = -self.goal_policy(feat.detach()).log_prob(next_observations)
        if nll.detach().min().item() <= 0:  # por.py:104-105 is a pdb trap here
            raise FloatingPointError("non-positive per-sample NLL (reference would drop into pdb)")
        g_loss = (weight * nll).mean()
        self.goal_policy_optimizer.zero_grad(set_to_none=True)
        g_loss.backward()
        self.goal_policy_optimizer.step()
        self.goal_lr_schedule.step()
        return v_loss.item(), g_loss.item()


class OracleSORL(_ExpectileAgent):
    _online, _target = "v_net", "v_tgt"

    def __init__(self, args, max_steps, tau, alpha, backbone=None, value_lr=1e-4, policy_lr=1e-4,
                 discount=0.99, beta=0.005):
        super().__init__()
        self.backbone = backbone
        width = args.state_size if backbone is None else args.feature_dim
        self.v_net = TwinV(width, args.hidden_dim, args.n_hidden, layer_norm=args.layer_norm)
        self.policy = GaussianHead(width, args.action_size, args.hidden_dim, args.n_hidden, bounded=True)
        self.v_tgt = copy.deepcopy(self.v_net).requires_grad_(False)
        self.v_optimizer = torch.optim.Adam(self.v_net.parameters(), lr=value_lr)
        self.policy_optimizer = torch.optim.Adam(self.policy.parameters(), lr=policy_lr)
        self.lr_schedule = CosineAnnealingLR(self.policy_optimizer, max_steps)
        self.tau, self.alpha, self.discount, self.beta = tau, alpha, discount, beta

    def _encode(self, observations, next_observations):
        if self.backbone is None:
            return observations, next_observations
        return self.backbone(observations), self.backbone(next_observations)

    def select_action(self, observations):
        with torch.no_grad():
            feat = observations if self.backbone is None else self.backbone(observations)
            return self.policy(feat).mean.cpu().numpy()

    def vf_update(self, observations, actions, rewards, next_observations, terminals):
        feat, next_feat = self._encode(observations, next_observations)
        _, v_loss = self._value_step(feat, next_feat, rewards, terminals)
        return v_loss.item()

    def update(self, observations, actions, rewards, next_observations, terminals):
        feat, next_feat = self._encode(observations, next_observations)
        target_v, v_loss = self._value_step(feat, next_feat, rewards, terminals)
        # advantage-weighted behaviour cloning; SORL multiplies by alpha (sorl.py:102-113)
        adv = target_v - self.v_net(feat)
        weight = torch.exp(self.alpha * adv).clamp_max(WEIGHT_CAP).detach()
        nll = -self.policy(feat.detach()).log_prob(actions)
        g_loss = (weight * nll).mean()
        self.policy_optimizer.zero_grad(set_to_none=True)
        g_loss.backward()
        self.policy_optimizer.step()
        self.lr_schedule.step()
        return v_loss.item(), g_loss.item()


class OracleCQL:
    """Discrete CQL on a replay buffer (cql_trainer.py:88-124 with the intended dqn_trainer.py ctor).

    The reference's CQLTrainer.__init__ is broken at HEAD (cql_trainer.py:44-55 vs
    dqn_trainer.py:35-51); this restates the evident intent: Adam(lr=5e-4) hard-coded
    (dqn_trainer.py:71), batch 64 (:78).  Hard target copy (dqn_trainer.py:195-196): learn()
    increments self.training_step (cql_trainer.py:122) BEFORE train_offline tests
    `training_step % update_target_freq == 0`, so the copy fires after the k-th update when
    k % freq == 0 (k 1-based), not at step 0 [pinned by tests/golden/cql_*.npz].  Logging is
    left out (host I/O).
    """

    def __init__(self, state_size, action_size, gamma=0.99, update_target_freq=10, alpha=1.0,
                 replay_buffer=None, batch_size=64, lr=5e-4, dtype=torch.float32):
        self.state_size, self.action_size = state_size, action_size
        self.gamma, self.alpha, self.update_target_freq = gamma, alpha, update_target_freq
        self.q_network = QNet(state_size, action_size).to(dtype)
        self.target_network = QNet(state_size, action_size).to(dtype)
        self.target_network.load_state_dict(self.q_network.state_dict())
        self.target_network.eval()
        self.optimizer = torch.optim.Adam(self.q_network.parameters(), lr=lr)
        self.replay_buffer, self.batch_size, self.training_step = replay_buffer, batch_size, 0

    def compute_cql_penalty(self, states, actions):
        q = self.q_network(states)
        lse = torch.logsumexp(q, dim=1) - math.log(self.action_size)
        return (lse - q.gather(1, actions[:, None]).squeeze(1)).mean()

    def learn_on(self, states, actions, rewards, next_states, dones):
        """One gradient update on an explicit batch; returns (loss, td, penalty) floats."""
        with torch.no_grad():
            nq = self.target_network(next_states)
            targets = rewards + self.gamma * nq.max(dim=1).values * (1 - dones)
        q_a = self.q_network(states).gather(1, actions[:, None]).squeeze(1)
        td = torch.nn.functional.mse_loss(q_a, targets)
        penalty = self.compute_cql_penalty(states, actions)
        loss = td + self.alpha * penalty
        self.optimizer.zero_grad()
        loss.backward()
        self.optimizer.step()
        return loss.item(), td.item(), penalty.item()

    def learn(self):
        loss = self.learn_on(*self.replay_buffer.sample(self.batch_size))[0]
        self.training_step += 1  # cql_trainer.py:122 (a logger always exists, dqn_trainer.py:81)
        return loss

    def train_offline(self, num_iterations=10000):
        losses = []
        for self.training_step in range(num_iterations):
            losses.append(self.learn())
            if self.training_step % self.update_target_freq == 0:
                self.target_network.load_state_dict(self.q_network.state_dict())
        return losses

    def get_action(self, state: np.ndarray) -> int:
        with torch.no_grad():
            q = self.q_network(torch.as_tensor(state, dtype=torch.float32)[None])
        return int(q.argmax(dim=1).item())
