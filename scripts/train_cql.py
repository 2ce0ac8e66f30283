#!/usr/bin/env python
"""Discrete CQL offline training on B200: scripts/train_cql.py:14-40 of the reference with the same trainer
keywords.  The reference collects its dataset by random roll-outs in gymnasium's LunarLander
(policy/collect_dataset.py:5-18; gymnasium is absent here), so the replay buffer is filled with a synthetic
LaserScan-shaped dataset instead: --state_size 362 --action_size 5 (the Gazebo fixture) or 8/4
(the LunarLander shape)."""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from porl_b200 import CQLTrainer, QNetwork  # noqa: E402
from porl_b200.dataloader.dataset import synthetic_rows  # noqa: E402


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--state_size", type=int, default=362)
    p.add_argument("--action_size", type=int, default=5)
    p.add_argument("--dataset_rows", type=int, default=100_000)       # ReplayBuffer capacity, dqn_trainer.py:76
    p.add_argument("--batch_size", type=int, default=64)              # dqn_trainer.py:78
    p.add_argument("--num_iterations", type=int, default=10_000)      # dqn_trainer.py:182
    p.add_argument("--seed", type=int, default=0)
    p.add_argument("--log_dir", type=str, default="logs")
    a = p.parse_args()
    device = torch.device("cuda")
    torch.manual_seed(1000 + a.seed)
    np.random.seed(2000 + a.seed)
    trainer = CQLTrainer(state_size=a.state_size, action_size=a.action_size, gamma=0.99, epsilon=1.0,
                         epsilon_min=0.05, epsilon_decay=0.99, update_target_freq=10, device=device,
                         network=QNetwork, log_dir=a.log_dir)          # scripts/train_cql.py:18-29
    trainer.batch_size = a.batch_size
    S = a.state_size
    rows = synthetic_rows(min(a.dataset_rows, 100_000), S, 20260103 + a.seed, 2)
    acts = np.random.default_rng(a.seed).integers(0, a.action_size, len(rows))
    trainer.replay_buffer.push_batch(rows[:, :S], acts, rows[:, S], rows[:, S + 1:2 * S + 1], rows[:, 2 * S + 1])
    losses = trainer.train_offline(num_iterations=a.num_iterations)    # scripts/train_cql.py:36
    print({"iterations": len(losses), "first_loss": losses[0], "last_loss": losses[-1]})


if __name__ == "__main__":
    main()
