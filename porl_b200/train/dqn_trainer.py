"""DQN base trainer: constructor + offline loop of src/porl/train/dqn_trainer.py:35-91,93-117,182-231."""
from __future__ import annotations

import os
from typing import List, Type

import numpy as np
import torch

from .. import _C, dist as pdist
from .._C import check, lib
from ..buffer.replaybuffer import ReplayBuffer
from ..dataloader.dataset import PackedBatch
from ..net.q_network import QNetwork
from ..nets import ParamGroup, Workspace, mlp_backward, mlp_call, mlp_forward
from ..utils.logger import Logger


class DQNTrainer:
    """Same constructor as the reference (dqn_trainer.py:35-51).  Hard-coded there and kept:
    Adam lr 0.0005 (:71), ReplayBuffer(100000) (:76), batch_size 64 (:78)."""

    cql_alpha = 0.0      # plain DQN: TD loss only (dqn_trainer.py:93-117)

    def __init__(self, state_size: int, action_size: int, gamma: float, epsilon: float, epsilon_min: float,
                 epsilon_decay: float, learning_rate: float, update_target_freq: int, device: torch.device,
                 network: Type[QNetwork] | None = QNetwork, log_dir: str = "logs", replay_buffer=None,
                 transition_learning_step=10000):
        self.state_size, self.action_size = int(state_size), int(action_size)
        self.device = torch.device(device)
        self.ctx = _C.context(self.device)
        self.gamma, self.epsilon, self.epsilon_min = float(gamma), epsilon, epsilon_min
        self.epsilon_decay, self.learning_rate = epsilon_decay, learning_rate
        self.update_target_freq = int(update_target_freq)
        self.training_learning_step = transition_learning_step
        if network is not None:
            if not (isinstance(network, type) and issubclass(network, QNetwork)):
                raise TypeError("porl_b200 trainers take porl_b200.QNetwork (or a subclass) as `network`")
            probe = network(state_size, action_size, device=self.device, init=False)
            self.q_group = ParamGroup(self.ctx, probe.plan.n_floats, self.device, lr=0.0005)   # dqn_trainer.py:71
            self.q_network = network(state_size, action_size, device=self.device, flat=self.q_group.params,
                                     grads=self.q_group.grads)
            self.target_network = network(state_size, action_size, device=self.device, trainable=False)
            self.target_network.load_state_dict(self.q_network.state_dict())
            self.optimizer = self.q_group
        self.replay_buffer = replay_buffer if replay_buffer is not None else ReplayBuffer(
            100000, (self.state_size,), self.device)
        self.batch_size = 64
        self.logger = Logger(log_dir=log_dir)
        self.logger.log_hyperparameters({"learning_rate": self.learning_rate, "gamma": self.gamma,
                                         "batch_size": self.batch_size, "epsilon_decay": self.epsilon_decay,
                                         "update_target_freq": self.update_target_freq})
        self._ws = Workspace(self.device)
        self._host = _C.pinned(4, torch.float32)
        self.training_step = 0

    # ------------------------------------------------------------------ one update on a device batch
    @property
    def fused(self) -> bool:
        """The single-kernel update (porl_cql_fused_grad) serves this network shape and is not switched off."""
        if not hasattr(self, "_fused_ok"):
            self._fused_ok = bool(lib.porl_cql_fused_supported(self.q_network.plan.handle))
        return self._fused_ok and os.environ.get("PORL_CQL_FUSED", "1") != "0"

    def _fused_grad(self, states, next_states, ld, actions, rewards, dones, idx, n_rows: int, B: int) -> None:
        check(lib.porl_cql_fused_grad(self.ctx.handle, self.q_network.plan.handle, self.q_group.params.data_ptr(),
                                      self.target_network.flat.data_ptr(), states.data_ptr(), next_states.data_ptr(),
                                      ld, actions.data_ptr(), rewards.data_ptr(), dones.data_ptr(),
                                      idx.data_ptr() if idx is not None else None, n_rows, self.gamma,
                                      float(self.cql_alpha), B, B * pdist.world_size(), self.q_group.grads.data_ptr(),
                                      self.q_group.loss_ptr, _C.stream_ptr()))

    def _two_launch(self) -> bool:
        """Adam folded into the gradient reduction (porl_cql_fused_step): single process only -- under data
        parallelism the optimizer kernel is the one that sums the ranks' buckets."""
        return pdist.world_size() == 1 and self.q_group.peers is None and os.environ.get("PORL_CQL_FOLD_ADAM", "1") != "0"

    def _fused_step(self, states, next_states, ld, actions, rewards, dones, idx, n_rows: int, B: int, hard_every: int) -> None:
        g = self.q_group
        check(lib.porl_cql_fused_step(self.ctx.handle, self.q_network.plan.handle, g.params.data_ptr(),
                                      self.target_network.flat.data_ptr(), states.data_ptr(), next_states.data_ptr(),
                                      ld, actions.data_ptr(), rewards.data_ptr(), dones.data_ptr(),
                                      idx.data_ptr() if idx is not None else None, n_rows, self.gamma,
                                      float(self.cql_alpha), B, g.grads.data_ptr(), g.loss_ptr, g.m.data_ptr(),
                                      g.v.data_ptr(), g.state.data_ptr(), g.betas[0], g.betas[1], g.eps, int(hard_every),
                                      _C.stream_ptr()))

    def learn_indices(self, idx: torch.Tensor, hard_every: int = 0) -> None:
        """TD (+ CQL) update on rows `idx` of the replay buffer; asynchronous.  With the fused kernel the rows go
        from the HBM-resident buffer straight into shared memory (no gathered copy of the minibatch exists)."""
        rb = self.replay_buffer
        if not self.fused:
            self._pb = rb.gather_packed(idx, getattr(self, "_pb", None))
            return self.learn_packed(self._pb, hard_every)
        if idx.device != self.device or idx.dtype != torch.int64 or not idx.is_contiguous():
            idx = idx.to(self.device, torch.int64).contiguous()
        if self._two_launch():
            return self._fused_step(rb._states, rb._next_states, rb.ld, rb.actions, rb.rewards, rb.dones, idx, rb.capacity,
                                    int(idx.shape[0]), hard_every)
        self._fused_grad(rb._states, rb._next_states, rb.ld, rb.actions, rb.rewards, rb.dones, idx, rb.capacity,
                         int(idx.shape[0]))
        self.q_group.step(target=self.target_network.flat if hard_every > 0 else None, hard_every=hard_every)

    def learn_packed(self, pb: PackedBatch, hard_every: int = 0) -> None:
        """TD (+ CQL) update on a packed minibatch; asynchronous.  losses land in q_group.tail[0:3]."""
        ctx, B, plan = self.ctx, pb.B, self.q_network.plan
        if self.fused and self._two_launch():
            return self._fused_step(pb.obs, pb.next_obs, pb.ld_obs, pb.actions, pb.rewards, pb.dones, None, B, B, hard_every)
        if self.fused:
            self._fused_grad(pb.obs, pb.next_obs, pb.ld_obs, pb.actions, pb.rewards, pb.dones, None, B, B)
            self.q_group.step(target=self.target_network.flat if hard_every > 0 else None, hard_every=hard_every)
            return
        ws = self._ws
        scratch_t = ws.get(("scratch_t", B), plan.saved_floats(B))
        saved_q = ws.get(("saved_q", B), plan.saved_floats(B))
        out_t = ws.get(("out_t", B), B * plan.out_ld)
        out_q = ws.get(("out_q", B), B * plan.out_ld)
        dq = ws.get(("dq", B), B * plan.out_ld)
        call_t = mlp_call(plan, self.target_network.flat, pb.next_obs, pb.ld_obs, scratch_t, out_t, inference=True)
        call_q = mlp_call(plan, self.q_group.params, pb.obs, pb.ld_obs, saved_q, out_q)
        mlp_forward(ctx, [call_t, call_q], B)
        check(lib.porl_cql_head(ctx.handle, out_q.data_ptr(), out_t.data_ptr(), plan.out_ld, pb.actions.data_ptr(),
                                pb.rewards.data_ptr(), pb.dones.data_ptr(), self.gamma, float(self.cql_alpha), B,
                                self.action_size, B * pdist.world_size(), dq.data_ptr(), self.q_group.loss_ptr,
                                _C.stream_ptr()))
        mlp_backward(ctx, call_q, B, dq, self.q_group.grads)
        self.q_group.step(target=self.target_network.flat if hard_every > 0 else None, hard_every=hard_every)

    def read_losses(self) -> tuple[float, float, float]:
        """Synchronising read of (loss, td, penalty) of the last update."""
        self._host[0:3].copy_(self.q_group.tail[0:3], non_blocking=True)
        _C.sync_stream()
        return float(self._host[0]), float(self._host[1]), float(self._host[2])

    def learn(self) -> float:
        """One gradient update on a batch drawn from the replay buffer (dqn_trainer.py:93-117)."""
        self.learn_indices(torch.as_tensor(self.replay_buffer.draw_indices(self.batch_size)))
        return self.read_losses()[0]

    def sync_target(self) -> None:
        """target_network.load_state_dict(q_network.state_dict()) (dqn_trainer.py:196)."""
        self.target_network.flat.copy_(self.q_group.params[:self.target_network.flat.numel()])

    def train_offline(self, policy=None, num_iterations: int = 10000, verbose: bool = True) -> List[float]:
        losses = []
        for self.training_step in range(num_iterations):
            loss = self.learn() if policy is None else policy()
            self.logger.log_loss(self.training_step, loss)
            if self.training_step % self.update_target_freq == 0:
                self.sync_target()
            losses.append(loss)
            if verbose and self.training_step % 10 == 0:
                print(f"step {self.training_step}, Loss: {loss}")
        self.logger.close()
        return losses

    def select_action(self, state: np.ndarray) -> int:
        """Epsilon-greedy action (dqn_trainer.py:206-231)."""
        if np.random.rand() < self.epsilon:
            return int(np.random.randint(self.action_size))
        q = self.q_network(torch.as_tensor(np.asarray(state, dtype=np.float32))[None])
        return int(q.argmax().item())
