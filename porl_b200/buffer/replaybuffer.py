"""HBM-resident replay buffer (src/porl/buffer/replaybuffer.py:5-76, same class in buffer/replay_buffer.py).

Same constructor, `push`, `sample`, `__len__`; the SoA arrays live on the device and `sample` is one
gather kernel (porl_gather_soa) instead of five numpy fancy-index gathers + pageable H2D copies.
The index draw is kept identical to the reference -- `np.random.choice(size, B, replace=False)`
(replaybuffer.py:61) -- so a seeded run samples the same transitions; `sample_indices` / `gather_packed`
take caller-provided (e.g. device-generated) indices and are what the throughput path uses, since
the O(size) numpy draw costs 35 ms at 1 M rows.
"""
from __future__ import annotations

import numpy as np
import torch

from .. import _C
from .._C import check, lib
from ..dataloader.dataset import PackedBatch
from ..nets import pad_to


class ReplayBuffer:
    def __init__(self, capacity: int, state_shape: tuple, device: torch.device):
        if len(state_shape) != 1:
            raise ValueError("porl_b200.ReplayBuffer stores flat state vectors (state_shape=(S,))")
        self.capacity = int(capacity)
        self.state_shape = tuple(state_shape)
        self.device = torch.device(device)
        self.ctx = _C.context(self.device)
        self.size = 0
        self.position = 0
        self.S = int(state_shape[0])
        self.ld = pad_to(self.S)
        z = lambda *s, dt=torch.float32: torch.zeros(*s, dtype=dt, device=self.device)
        self._states, self._next_states = z(self.capacity, self.ld), z(self.capacity, self.ld)
        self.actions = z(self.capacity, dt=torch.int64)
        self.rewards, self.dones = z(self.capacity), z(self.capacity)

    # logical views with the reference's attribute names
    @property
    def states(self) -> torch.Tensor:
        return self._states[:, :self.S]

    @property
    def next_states(self) -> torch.Tensor:
        return self._next_states[:, :self.S]

    def push(self, state, action, reward, next_state, done) -> None:
        """Add one experience (replaybuffer.py:30-48).  Host-driven, one small copy per field;
        bulk ingestion should use push_batch."""
        p = self.position
        self._states[p, :self.S] = torch.as_tensor(np.asarray(state, dtype=np.float32))
        self._next_states[p, :self.S] = torch.as_tensor(np.asarray(next_state, dtype=np.float32))
        self.actions[p] = int(action)
        self.rewards[p] = float(reward)
        self.dones[p] = float(done)
        self.position = (p + 1) % self.capacity
        self.size = min(self.size + 1, self.capacity)

    def push_batch(self, states, actions, rewards, next_states, dones) -> None:
        """Ring-insert n experiences at once (same final contents as n push() calls)."""
        n = int(len(actions))
        if n > self.capacity:
            raise ValueError("batch larger than the buffer")
        t = lambda a, dt: (a.to(self.device, dt) if torch.is_tensor(a)
                           else torch.as_tensor(np.asarray(a), dtype=dt).to(self.device))   # device tensors stay put
        pos = (self.position + torch.arange(n, device=self.device)) % self.capacity
        self._states[pos, :self.S] = t(states, torch.float32)
        self._next_states[pos, :self.S] = t(next_states, torch.float32)
        self.actions[pos] = t(actions, torch.int64)
        self.rewards[pos] = t(rewards, torch.float32)
        self.dones[pos] = t(dones, torch.float32)
        self.position = (self.position + n) % self.capacity
        self.size = min(self.size + n, self.capacity)

    def new_batch(self, batch: int) -> PackedBatch:
        return PackedBatch(batch, self.S, 1, self.device, discrete_actions=True)

    def gather_packed(self, idx: torch.Tensor, out: PackedBatch | None = None) -> PackedBatch:
        if idx.device != self.device or idx.dtype != torch.int64 or not idx.is_contiguous():
            idx = idx.to(self.device, torch.int64).contiguous()
        B = idx.shape[0]
        if out is None or out.B != B:
            out = self.new_batch(B)
        check(lib.porl_gather_soa(self.ctx.handle, self._states.data_ptr(), self._next_states.data_ptr(),
                                  self.actions.data_ptr(), self.rewards.data_ptr(), self.dones.data_ptr(),
                                  self.capacity, self.S, self.ld, idx.data_ptr(), B, out.obs.data_ptr(),
                                  out.next_obs.data_ptr(), out.ld_obs, out.actions.data_ptr(),
                                  out.rewards.data_ptr(), out.dones.data_ptr(), _C.stream_ptr()))
        return out

    def draw_indices(self, batch_size: int) -> np.ndarray:
        return np.random.choice(self.size, batch_size, replace=False)     # replaybuffer.py:61

    def sample_indices(self, idx) -> tuple:
        pb = self.gather_packed(torch.as_tensor(idx))
        return pb.observations, pb.actions, pb.rewards, pb.next_observations, pb.dones

    def sample(self, batch_size: int) -> tuple:
        """(states, actions[int64], rewards, next_states, dones) on the device (replaybuffer.py:50-72)."""
        return self.sample_indices(self.draw_indices(batch_size))

    def __len__(self) -> int:
        return self.size
