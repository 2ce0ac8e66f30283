"""Oracle minibatch sources (numpy / torch CPU).  Test infrastructure.

Follows:
  src/porl/buffer/replaybuffer.py:14-76  SoA ring buffer, sample = np.random.choice(size, B, replace=False)
  por_train.py:59-63,71-78 / sorl_train.py:42-46,52-58  DataLoader(shuffle=True) over AoS rows
      [obs(S) | r | next_obs(S) | done | action(A)], drop_last=False, caller slices columns
"""
from __future__ import annotations

import numpy as np
import torch


class RingBuffer:
    """Pre-allocated SoA ring (replaybuffer.py:24-48); `sample` returns CPU tensors."""

    def __init__(self, capacity, state_shape):
        self.capacity, self.size, self.position = capacity, 0, 0
        self.states = np.zeros((capacity, *state_shape), np.float32)
        self.next_states = np.zeros((capacity, *state_shape), np.float32)
        self.actions = np.zeros(capacity, np.int64)
        self.rewards = np.zeros(capacity, np.float32)
        self.dones = np.zeros(capacity, np.float32)

    def push(self, state, action, reward, next_state, done):
        p = self.position
        self.states[p], self.actions[p], self.rewards[p] = state, action, reward
        self.next_states[p], self.dones[p] = next_state, float(done)
        self.position = (p + 1) % self.capacity
        self.size = min(self.size + 1, self.capacity)

    def draw_indices(self, batch_size):
        return np.random.choice(self.size, batch_size, replace=False)  # replaybuffer.py:61

    def take(self, idx):
        return (torch.from_numpy(self.states[idx]), torch.from_numpy(self.actions[idx]),
                torch.from_numpy(self.rewards[idx]), torch.from_numpy(self.next_states[idx]),
                torch.from_numpy(self.dones[idx]))

    def sample(self, batch_size):
        return self.take(self.draw_indices(batch_size))

    def __len__(self):
        return self.size


def epoch_batches(n_rows, batch_size, generator):
    """Index stream of DataLoader(shuffle=True, drop_last=False): one randperm per epoch,
    chopped into consecutive chunks, the last one ragged (torch RandomSampler semantics)."""
    perm = torch.randperm(n_rows, generator=generator)
    return [perm[i:i + batch_size] for i in range(0, n_rows, batch_size)]


def slice_row(rows, state_size, action_size=2):
    """Column split the trainers apply to a collated AoS batch (por_train.py:74-78)."""
    s, a = state_size, action_size
    return dict(observations=rows[:, :s], rewards=rows[:, s], next_observations=rows[:, s + 1:-a - 1],
                terminals=rows[:, -a - 1], actions=rows[:, -a:])
