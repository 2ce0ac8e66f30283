"""SORL agent (agent/sorl.py:21-152): drop-in constructor, `update`, `vf_update`, `select_action`."""
from __future__ import annotations

import torch

from .. import _C
from ._expectile import ExpectileAgent


class SORL(ExpectileAgent):
    """Same value step as POR; advantage-weighted behaviour cloning of the continuous action with a
    tanh-bounded Gaussian policy.

    Same signature as the reference (agent/sorl.py:21-31; note `device` precedes `backbone`).
    state_dict keys: v_net.v{1,2}.{0,2,4}.*, policy.log_std, policy.net.{0,2,4}.*, v_tgt.v{1,2}.* .
    """

    _value_name, _target_name, _policy_name = "v_net", "v_tgt", "policy"
    _awr_mode = _C.AWR_SORL         # weight = exp(alpha * adv), agent/sorl.py:104
    _bounded = True
    _policy_first = False           # sorl.py:36-45 builds v_net before policy

    def __init__(agent, args, max_steps, tau, alpha, device=torch.device("cuda"), backbone=None, value_lr=1e-4,
                 policy_lr=1e-4, discount=0.99, beta=0.005):
        super().__init__()
        agent._build(args, max_steps, tau, alpha, backbone, device, value_lr, policy_lr, discount, beta,
                     policy_out=args.action_size)
        agent.policy_optimizer = agent._policy_optimizer
        agent.lr_schedule = agent._schedule

    def _policy_target(agent, pb):
        return pb.actions, pb.ld_act          # log_prob(actions), agent/sorl.py:108

    def select_action(agent, observations):
        """agent/sorl.py:71-76: policy mean as a numpy array (B, A)."""
        if agent.backbone is not None:                  # sorl.py:73-74 (train-mode encoder, as the reference)
            observations = agent.backbone(observations)
        return agent.policy(observations).mean.cpu().numpy()

    def update(agent, observations, actions, rewards, next_observations, terminals):
        """agent/sorl.py:78-128.  Returns (v_loss, g_loss) as Python floats (one host sync)."""
        pb = agent.pack(observations, next_observations, rewards, terminals, actions)
        agent._run_update(pb)
        return agent.read_losses()

    def update_async(agent, observations, actions, rewards, next_observations, terminals):
        """`update` without the host synchronisation: enqueues the same work and returns a handle whose `.result()`
        gives (v_loss, g_loss) once THIS update's losses have reached the host.  Lets a training loop keep the next
        minibatch's host->device copy and the next update in flight while it logs the previous losses (extension; the
        reference's update() blocks on .item(), agent/sorl.py:127-128)."""
        pb = agent.pack(observations, next_observations, rewards, terminals, actions)
        agent._run_update(pb)
        return agent.losses_async()

    def vf_update(agent, observations, actions, rewards, next_observations, terminals):
        """agent/sorl.py:130-152: value step + Polyak update only; returns v_loss."""
        pb = agent.pack(observations, next_observations, rewards, terminals, actions)
        agent._run_update(pb, value_only=True)
        return agent.read_losses()[0]
