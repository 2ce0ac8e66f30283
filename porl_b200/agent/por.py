"""POR agent (agent/por.py:21-112): drop-in constructor, `por_residual_update`, `state_dict` layout."""
from __future__ import annotations

import torch

from .. import _C
from ._expectile import ExpectileAgent


class POR(ExpectileAgent):
    """Policy-guided Offline RL: TwinV expectile value learning + an advantage-weighted Gaussian
    "goal policy" that predicts the raw next state.

    Same signature as the reference (agent/por.py:21-31); `args` provides state_size, hidden_dim,
    n_hidden, layer_norm (and feature_dim when a backbone is used).  state_dict keys:
    goal_policy.log_std, goal_policy.net.{0,2,4}.*, vf.v{1,2}.{0,2,4}.*, v_target.v{1,2}.* .
    """

    _value_name, _target_name, _policy_name = "vf", "v_target", "goal_policy"
    _awr_mode = _C.AWR_POR          # weight = exp(adv / alpha), agent/por.py:100
    _bounded = False
    _policy_first = True            # por.py:35-45 builds goal_policy before vf

    def __init__(agent, args, max_steps, tau, alpha, backbone=None, device=torch.device("cuda"), value_lr=1e-4,
                 policy_lr=1e-4, discount=0.99, beta=0.005):
        super().__init__()
        agent._build(args, max_steps, tau, alpha, backbone, device, value_lr, policy_lr, discount, beta,
                     policy_out=args.state_size)
        agent.goal_policy_optimizer = agent._policy_optimizer
        agent.goal_lr_schedule = agent._schedule
        agent.step = 0
        agent.pretrain_step = 0

    def _policy_target(agent, pb):
        return pb.next_obs, pb.ld_obs          # log_prob(next_observations), agent/por.py:103

    def select_action(agent, observations):
        """Mean of the goal policy for a batch of observations, as a numpy array (B, state_size).  The reference's
        evaluation loop calls `agent.select_action(state[None, :])` (test.py:29) but its POR class has no such
        method (only SORL does, agent/sorl.py:71-76); this is the same forward for POR's guide policy."""
        if agent.backbone is not None:
            observations = agent.backbone(observations)
        return agent.goal_policy(observations).mean.cpu().numpy()

    def por_residual_update(agent, observations, next_observations, rewards, terminals):
        """agent/por.py:73-112.  Returns (v_loss, g_loss) as Python floats (one host sync)."""
        pb = agent.pack(observations, next_observations, rewards, terminals)
        agent._run_update(pb)
        return agent.read_losses()

    def por_residual_update_async(agent, observations, next_observations, rewards, terminals):
        """`por_residual_update` without the host synchronisation; returns a handle whose `.result()` gives
        (v_loss, g_loss) of this update (see SORL.update_async)."""
        pb = agent.pack(observations, next_observations, rewards, terminals)
        agent._run_update(pb)
        return agent.losses_async()
