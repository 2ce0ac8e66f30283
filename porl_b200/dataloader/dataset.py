"""HBM-resident offline dataset and its minibatch gather (kernel K1).

Stands in for dataloader/dataloader.py:10-55 (`CustomDataset`) + torch DataLoader(shuffle=True)
(por_train.py:59-63, sorl_train.py:42-46).  The reference keeps AoS rows
    [obs(S) | reward | next_obs(S) | done | action(A)]        (2S+2+A floats; dataloader.py:13)
and the trainers slice the collated batch (por_train.py:74-78).  Here the whole table lives in HBM
(1e7 rows x 2.9 KB = 29 GB fits a B200) and one kernel gathers rows straight into the padded SoA
minibatch the update kernels read; payload bytes are copied bit-exactly.
"""
from __future__ import annotations

import numpy as np
import torch

from .. import _C
from .._C import check, lib
from ..nets import pad_to


class PackedBatch:
    """Device minibatch in the layout the contractions need (feature rows padded to 32 floats)."""

    def __init__(self, batch: int, state_size: int, action_size: int, device, discrete_actions: bool = False):
        self.B, self.S, self.A = int(batch), int(state_size), int(action_size)
        self.ld_obs = pad_to(self.S)
        self.ld_act = pad_to(max(self.A, 1), 4)
        dev = torch.device(device)
        z = lambda *s, dt=torch.float32: torch.zeros(*s, dtype=dt, device=dev)
        self.obs, self.next_obs = z(self.B, self.ld_obs), z(self.B, self.ld_obs)
        self.rewards, self.dones = z(self.B), z(self.B)
        self.actions = z(self.B, dt=torch.int64) if discrete_actions else z(self.B, self.ld_act)
        self.discrete = discrete_actions

    # logical (unpadded) views, for tests and API compatibility
    @property
    def observations(self):
        return self.obs[:, :self.S]

    @property
    def next_observations(self):
        return self.next_obs[:, :self.S]

    @property
    def action_values(self):
        return self.actions if self.discrete else self.actions[:, :self.A]


class TransitionDataset:
    """AoS transition table resident in HBM.

    `rows`: (N, 2S+2+A) float32 (numpy or torch, any device).  Rows are stored with a stride padded
    to a multiple of 4 floats so that 16-byte vector loads stay aligned (S=365 rows are 2936 B).
    """

    def __init__(self, rows, state_size: int, action_size: int = 2, device="cuda"):
        self.device = torch.device(device)
        self.ctx = _C.context(self.device)
        self.S, self.A = int(state_size), int(action_size)
        self.row_len = 2 * self.S + 2 + self.A
        rows = torch.as_tensor(rows) if not torch.is_tensor(rows) else rows
        if rows.dim() != 2 or rows.shape[1] != self.row_len or rows.dtype != torch.float32:
            raise ValueError(f"expected float32 rows of length {self.row_len}, got {tuple(rows.shape)} {rows.dtype}")
        self.n_rows = int(rows.shape[0])
        self.row_stride = pad_to(self.row_len, 4)
        self.table = torch.zeros(self.n_rows, self.row_stride, dtype=torch.float32, device=self.device)
        self.table[:, :self.row_len].copy_(rows)          # one-off ingest (H2D + pad), not on the step path

    def __len__(self) -> int:                              # dataloader.py:27-32
        return self.n_rows

    def __getitem__(self, i: int) -> torch.Tensor:         # dataloader.py:34-52: one (2S+2+A,) row
        return self.table[i, :self.row_len]

    def gather(self, idx: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
        """(B, 2S+2+A) collated batch, the tensor the reference's DataLoader yields (bit-exact)."""
        idx = idx.to(self.device, torch.int64).contiguous()
        B = idx.shape[0]
        if out is None:
            out = torch.empty(B, self.row_len, dtype=torch.float32, device=self.device)
        if B == 0:
            return out
        check(lib.porl_gather_rows_f32(self.ctx.handle, self.table.data_ptr(), self.n_rows, self.row_stride,
                                       self.row_len, idx.data_ptr(), B, out.data_ptr(), out.stride(0),
                                       _C.stream_ptr()))
        return out

    def new_batch(self, batch: int) -> PackedBatch:
        return PackedBatch(batch, self.S, self.A, self.device)

    def gather_packed(self, idx: torch.Tensor, out: PackedBatch | None = None) -> PackedBatch:
        """Gather + column split (por_train.py:74-78) in one kernel, into the padded SoA minibatch."""
        if idx.device != self.device or idx.dtype != torch.int64 or not idx.is_contiguous():
            idx = idx.to(self.device, torch.int64).contiguous()
        B = idx.shape[0]
        if out is None or out.B != B:
            out = self.new_batch(B)
        check(lib.porl_gather_transitions(self.ctx.handle, self.table.data_ptr(), self.n_rows, self.row_stride,
                                          self.S, self.A, idx.data_ptr(), B, out.obs.data_ptr(),
                                          out.next_obs.data_ptr(), out.ld_obs, out.rewards.data_ptr(),
                                          out.dones.data_ptr(), out.actions.data_ptr(), out.ld_act,
                                          _C.stream_ptr()))
        return out


class EpochSampler:
    """Index stream of DataLoader(shuffle=True, drop_last=False): one permutation per epoch chopped
    into consecutive batches, the last one ragged (torch RandomSampler semantics).  The permutation
    is drawn with the same call the reference's sampler makes (torch.randperm with a CPU generator),
    so a seeded run visits rows in the same order; indices are then kept on the device."""

    def __init__(self, n_rows: int, batch_size: int, device="cuda", generator: torch.Generator | None = None,
                 drop_last: bool = False):
        self.n, self.bs, self.device = int(n_rows), int(batch_size), torch.device(device)
        self.generator, self.drop_last = generator, drop_last

    def __len__(self) -> int:
        return self.n // self.bs if self.drop_last else (self.n + self.bs - 1) // self.bs

    def state_dict(self) -> dict:
        """What a resumed run needs to continue the same index stream: the generator state BEFORE the next epoch's
        permutation is drawn (checkpoints are written between epochs)."""
        g = self.generator if self.generator is not None else torch.default_generator
        return {"kind": "epoch", "rng_state": g.get_state()}

    def load_state_dict(self, sd: dict) -> None:
        g = self.generator if self.generator is not None else torch.default_generator
        g.set_state(sd["rng_state"].cpu())

    def __iter__(self):
        perm = torch.randperm(self.n, generator=self.generator).to(self.device, non_blocking=True)
        end = self.n - (self.n % self.bs) if self.drop_last else self.n
        for i in range(0, end, self.bs):
            yield perm[i:min(i + self.bs, end)]


class DeviceSampler:
    """Shuffled minibatch indices computed ON THE DEVICE (porl_sample_permutation): position i of epoch e is
    P_e(i) for a keyed pseudo-random bijection P_e of [0, n_rows), so a batch costs O(B), needs no permutation
    table and no host round-trip.  Same contract as DataLoader(shuffle=True) / RandomSampler (every row once per
    epoch, consecutive batches, ragged last batch unless drop_last) and as
    np.random.choice(size, B, replace=False) (distinct rows), not the same index stream -- the parity tests use
    EpochSampler / recorded indices instead."""

    def __init__(self, n_rows: int, batch_size: int, device="cuda", seed: int = 0, drop_last: bool = False):
        self.n, self.bs, self.device = int(n_rows), int(batch_size), torch.device(device)
        self.seed, self.drop_last = int(seed), drop_last
        self.ctx = _C.context(self.device)
        self.epoch = 0

    def __len__(self) -> int:
        return self.n // self.bs if self.drop_last else (self.n + self.bs - 1) // self.bs

    def indices(self, epoch: int, first: int, count: int, out: torch.Tensor | None = None) -> torch.Tensor:
        """P_epoch(first .. first + count - 1) as an int64 device tensor."""
        if out is None:
            out = torch.empty(count, dtype=torch.int64, device=self.device)
        check(lib.porl_sample_permutation(self.ctx.handle, self.n, self.seed, epoch, first, count, out.data_ptr(),
                                          _C.stream_ptr()))
        return out

    def state_dict(self) -> dict:
        return {"kind": "device", "seed": self.seed, "epoch": self.epoch}

    def load_state_dict(self, sd: dict) -> None:
        self.seed, self.epoch = int(sd["seed"]), int(sd["epoch"])

    def draw(self, batch_size: int | None = None, out: torch.Tensor | None = None) -> torch.Tensor:
        """One without-replacement draw of `batch_size` rows (the ReplayBuffer.sample contract): every call uses a
        fresh key."""
        self.epoch += 1
        return self.indices(self.epoch, 0, batch_size or self.bs, out)

    def __iter__(self):
        self.epoch += 1
        end = self.n - (self.n % self.bs) if self.drop_last else self.n
        for first in range(0, end, self.bs):
            yield self.indices(self.epoch, first, min(self.bs, end - first))


def synthetic_rows(n: int, state_size: int, seed: int, action_size: int = 2) -> np.ndarray:
    """Synthetic LaserScan+goal transitions: scan beams in {U(0.13,3.5), 3.5, 10.0},
    goal within 3.5 m, done ~ Bernoulli(0.01), unit-scale rewards, actions U(-1,1)^A.  Vectorised numpy;
    used by the trainers' --synthetic mode and by bench.py (tests use oracle.synth, an independent copy)."""
    rng = np.random.default_rng(seed)
    n_scan = min(360, state_size - 2)
    scan = rng.uniform(0.13, 3.5, (n, n_scan)).astype(np.float32)
    u = rng.random((n, n_scan), dtype=np.float32)
    scan[u < 0.25] = 3.5
    scan[u > 0.97] = 10.0
    extra = state_size - n_scan
    rad = rng.uniform(0.2, 3.5, n)
    th = rng.uniform(-np.pi, np.pi, n)
    tail = np.zeros((n, extra), np.float32)
    tail[:, -2], tail[:, -1] = rad * np.cos(th), rad * np.sin(th)
    if extra > 2:
        tail[:, :-2] = rng.uniform(-3.0, 3.0, (n, extra - 2))
    nscan = np.clip(scan + rng.normal(0, 0.05, scan.shape).astype(np.float32), 0.13, 3.5)
    nscan[scan == 3.5] = 3.5
    nscan[scan == 10.0] = 10.0
    ntail = tail + rng.normal(0, 0.05, tail.shape).astype(np.float32)
    done = (rng.random(n) < 0.01).astype(np.float32)
    rew = (rng.normal(0, 0.1, n) + done * rng.choice([-5.0, 5.0], n)).astype(np.float32)
    act = rng.uniform(-1, 1, (n, action_size)).astype(np.float32)
    return np.concatenate([scan, tail, rew[:, None], nscan, ntail, done[:, None], act], 1).astype(np.float32)
