"""Discrete CQL trainer (src/porl/train/cql_trainer.py:27-139)."""
from __future__ import annotations

from typing import Type

import numpy as np
import torch

from .._C import check, lib
from .. import _C
from ..net.q_network import QNetwork
from ..nets import mlp_call, mlp_forward, pack_rows
from .dqn_trainer import DQNTrainer


class CQLTrainer(DQNTrainer):
    """Keyword-compatible with scripts/train_cql.py:18-29.  The reference constructor is broken at
    HEAD (cql_trainer.py:44-55 passes 10 positionals into an 11-positional base, dqn_trainer.py:35-51);
    this forwards by keyword with the hard-coded lr the base ignores anyway (dqn_trainer.py:71)."""

    def __init__(self, state_size: int, action_size: int, gamma: float, epsilon: float, epsilon_min: float,
                 epsilon_decay: float, update_target_freq: int, device: torch.device,
                 network: Type[QNetwork] = QNetwork, log_dir: str = "logs", num_epochs: int = 1000,
                 threshold: float = 0.1, alpha: int = 1):
        super().__init__(state_size=state_size, action_size=action_size, gamma=gamma, epsilon=epsilon,
                         epsilon_min=epsilon_min, epsilon_decay=epsilon_decay, learning_rate=0.0005,
                         update_target_freq=update_target_freq, device=device, network=network, log_dir=log_dir)
        self.num_epochs = num_epochs
        self.threshold = threshold
        self.alpha = alpha

    @property
    def cql_alpha(self) -> float:
        return float(self.alpha)

    def compute_cql_penalty(self, states: torch.Tensor, actions: torch.Tensor) -> torch.Tensor:
        """mean(logsumexp_a Q(s,.) - log|A| - Q(s,a_data)) (cql_trainer.py:60-86), as a 0-d device tensor.
        Evaluated by the same head kernel the update uses (the TD half is ignored)."""
        B = int(states.shape[0])
        plan, ws = self.q_network.plan, self._ws
        x = ws.get(("pen_x", B), B * plan.in_ld)
        pack_rows(self.ctx, states.to(self.device), self.state_size, x, plan.in_ld)
        scratch = ws.get(("pen_s", B), plan.saved_floats(B))
        out = ws.get(("pen_o", B), B * plan.out_ld)
        dq = ws.get(("pen_dq", B), B * plan.out_ld)
        zeros = ws.get(("pen_z", B), B)
        res = ws.get(("pen_r", 0), 4)
        mlp_forward(self.ctx, [mlp_call(plan, self.q_group.params, x, plan.in_ld, scratch, out, inference=True)], B)
        acts = actions.to(self.device, torch.int64).contiguous()
        check(lib.porl_cql_head(self.ctx.handle, out.data_ptr(), out.data_ptr(), plan.out_ld, acts.data_ptr(),
                                zeros.data_ptr(), zeros.data_ptr(), 0.0, 1.0, B, self.action_size, B,
                                dq.data_ptr(), res.data_ptr(), _C.stream_ptr()))
        return res[2].clone()

    def learn(self) -> float:
        """cql_trainer.py:88-124: TD MSE + alpha * conservative penalty, one Adam step."""
        self.learn_indices(torch.as_tensor(self.replay_buffer.draw_indices(self.batch_size)))
        loss, td, pen = self.read_losses()
        if hasattr(self, "logger"):                                     # cql_trainer.py:116-122
            self.logger.writer.add_scalar("Loss/TD", td, self.training_step)
            self.logger.writer.add_scalar("Loss/CQLPenalty", pen, self.training_step)
            self.logger.writer.add_scalar("Loss/Total", loss, self.training_step)
            self.training_step += 1
        return loss

    def get_action(self, state: np.ndarray) -> int:
        """Greedy action for evaluation (cql_trainer.py:126-139)."""
        q = self.q_network(torch.as_tensor(np.asarray(state, dtype=np.float32))[None])
        return int(q.argmax(dim=1).item())
