"""Deterministic synthetic inputs and weights for oracle / golden / parity tests (CPU, numpy RNG).

Test infrastructure.  Shapes and value ranges follow SURVEY.md 8d: scan beams in
{U(0.13,3.5), 3.5 ("inf"), 10.0 ("nan/inf" sentinel of env/gazebo.py:80-81)}, goal within 3.5 m,
done ~ Bernoulli(0.01), unit-scale rewards, actions U(-1,1)^2 or randint(A).
Weights are a pure function of (key, shape, seed) so that the reference, the oracle and the
CUDA product can be given identical parameters through load_state_dict without sharing files.
"""
from __future__ import annotations

import zlib

import numpy as np
import torch


def laser_rows(n, state_size, seed, action_size=2):
    """(n, 2S+2+A) float32 AoS rows [obs | r | next_obs | done | action]  (por_train.py:74-78)."""
    rng = np.random.RandomState(seed)
    n_scan = min(360, state_size - 2)
    scan = rng.uniform(0.13, 3.5, (n, n_scan))
    u = rng.uniform(size=(n, n_scan))
    scan[u < 0.25] = 3.5
    scan[u > 0.97] = 10.0
    extra = state_size - n_scan
    rad, th = rng.uniform(0.2, 3.5, n), rng.uniform(-np.pi, np.pi, n)
    tail = np.zeros((n, extra))
    tail[:, -2], tail[:, -1] = rad * np.cos(th), rad * np.sin(th)
    if extra > 2:
        tail[:, :-2] = rng.uniform(-3.0, 3.0, (n, extra - 2))
    obs = np.concatenate([scan, tail], 1)
    nscan = np.clip(scan + rng.normal(0, 0.05, scan.shape), 0.13, 3.5)
    nscan[scan == 3.5], nscan[scan == 10.0] = 3.5, 10.0
    nxt = np.concatenate([nscan, tail + rng.normal(0, 0.05, tail.shape)], 1)
    done = (rng.uniform(size=n) < 0.01).astype(np.float64)
    rew = rng.normal(0, 0.1, n) + done * rng.choice([-5.0, 5.0], n)
    act = rng.uniform(-1, 1, (n, action_size))
    rows = np.concatenate([obs, rew[:, None], nxt, done[:, None], act], 1)
    return rows.astype(np.float32)


def discrete_actions(n, n_actions, seed):
    return np.random.RandomState(seed).randint(0, n_actions, n).astype(np.int64)


def det_state_dict(module, seed, dtype=torch.float32):
    """Deterministic replacement for every floating tensor of module.state_dict().

    weight (>=2-D): U(+-1/sqrt(fan_in));  bias / 1-D: U(+-0.1) (log_std: U(+-0.2));
    BatchNorm/LayerNorm weight: 1 + U(+-0.1); running_var: U(0.5,1.5); integer buffers untouched.
    """
    out = {}
    for key, ref in module.state_dict().items():
        if not ref.is_floating_point():
            out[key] = ref.clone()
            continue
        rng = np.random.RandomState((zlib.crc32(key.encode()) ^ (seed * 2654435761)) & 0x7FFFFFFF)
        shape = tuple(ref.shape)
        if key.endswith("running_var"):
            val = rng.uniform(0.5, 1.5, shape)
        elif ref.ndim >= 2:
            fan_in = int(np.prod(shape[1:]))
            val = rng.uniform(-1, 1, shape) / np.sqrt(fan_in)
        elif key.endswith("log_std"):
            val = rng.uniform(-0.2, 0.2, shape)
        elif key.endswith("weight"):  # norm scale
            val = 1.0 + rng.uniform(-0.1, 0.1, shape)
        else:
            val = rng.uniform(-0.1, 0.1, shape)
        out[key] = torch.from_numpy(np.asarray(val, np.float64)).to(dtype)
    return out
