"""porl_b200 -- B200-native (sm_100a) offline-RL update step of hadleyhzy34/porl.

Drop-in for the reference's agent/trainer API on ONE path: minibatch gather -> TwinV / policy /
Q-network MLPs -> POR / SORL / discrete-CQL losses -> Adam (+ Polyak / hard target update).
All arithmetic runs in hand-written CUDA kernels behind a C ABI (include/porl_b200.h,
porl_b200/libporl_b200.so); there is no CPU or eager-PyTorch fallback.
"""
import sys as _sys

if "porl_b200.build" not in getattr(_sys, "orig_argv", []):   # `python -m porl_b200.build` must run without the library
    from . import _C  # noqa: F401  (raises ImportError when the library has not been built)
    from .agent.por import POR  # noqa: F401
    from .agent.sorl import SORL  # noqa: F401
    from .buffer.replaybuffer import ReplayBuffer  # noqa: F401
    from .dataloader.dataset import TransitionDataset, EpochSampler  # noqa: F401
    from .train.cql_trainer import CQLTrainer  # noqa: F401
    from .train.dqn_trainer import DQNTrainer  # noqa: F401
    from .net.q_network import QNetwork  # noqa: F401

__version__ = "0.1.0"
