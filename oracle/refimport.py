"""Import the UNMODIFIED reference from /root/reference (build container only).  Test infrastructure.

Used by tests/golden/make_golden.py to generate fixtures and by the live-parity tests, which
skip when the mount is absent (it does not exist on the GPU box).  Nothing is copied: modules
are imported through sys.path with stubs for absent, off-path imports:
matplotlib (util/costmap.py:3), gymnasium + ipdb (src/porl/train/dqn_trainer.py:1,4).
pdb.set_trace is replaced by a raiser so a reference debugger trap fails instead of hanging.
"""
from __future__ import annotations

import os
import sys
import tempfile
import types

REF_ROOT = os.environ.get("PORL_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "agent", "por.py"))


def _stub(name):
    if name not in sys.modules:
        sys.modules[name] = types.ModuleType(name)
    return sys.modules[name]


def prepare():
    if not available():
        raise RuntimeError(f"reference not mounted at {REF_ROOT}")
    sys.dont_write_bytecode = True
    for p in (REF_ROOT, os.path.join(REF_ROOT, "src")):
        if p not in sys.path:
            sys.path.insert(0, p)
    try:
        import matplotlib  # noqa: F401
    except Exception:
        _stub("matplotlib").pyplot = _stub("matplotlib.pyplot")
    for name in ("gymnasium", "ipdb"):
        try:
            __import__(name)
        except Exception:
            _stub(name)
    import pdb

    def _trap(*a, **k):
        raise RuntimeError("reference hit pdb.set_trace()")

    pdb.set_trace = _trap


def por_cls():
    prepare()
    from agent.por import POR
    return POR


def sorl_cls():
    prepare()
    from agent.sorl import SORL
    return SORL


def pointmlp_module():
    """agent.feature_pcl of the reference (PointMLPModel and its helpers)."""
    prepare()
    import agent.feature_pcl as m
    return m


def replay_buffer_cls():
    prepare()
    from porl.buffer.replaybuffer import ReplayBuffer
    return ReplayBuffer


def make_cql_trainer(state_size, action_size, gamma=0.99, update_target_freq=10, alpha=1, log_dir=None):
    """Construct the reference CQLTrainer through the shim of SURVEY.md 8c: its own __init__
    passes 10 positionals into an 11-positional base (cql_trainer.py:44-55 vs dqn_trainer.py:35-51)."""
    prepare()
    import torch
    from porl.net.q_network import QNetwork
    from porl.train.cql_trainer import CQLTrainer
    from porl.train.dqn_trainer import DQNTrainer

    t = CQLTrainer.__new__(CQLTrainer)
    DQNTrainer.__init__(t, state_size=state_size, action_size=action_size, gamma=gamma, epsilon=1.0,
                        epsilon_min=0.05, epsilon_decay=0.99, learning_rate=5e-4,
                        update_target_freq=update_target_freq, device=torch.device("cpu"),
                        network=QNetwork, log_dir=log_dir or tempfile.mkdtemp(prefix="porl_ref_logs_"))
    t.num_epochs, t.threshold, t.alpha, t.training_step = 1000, 0.1, alpha, 0
    return t


def costmap_fn():
    prepare()
    from util.costmap import state2costmap
    return state2costmap


def fasternet_cls():
    prepare()
    from agent.fasternet import FasterNet
    return FasterNet
