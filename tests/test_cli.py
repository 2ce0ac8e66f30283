"""Trainer CLIs: reference flag names / defaults are kept, and the scripts run end to end."""
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# flag -> default of the reference parsers (por_train.py:124-150, sorl_train.py:84-104)
POR_FLAGS = {"namespace": "tb3", "state_size": 365, "action_size": 2, "episodes": 5000, "device": "cuda", "batch_size": 32,
             "replay_buffer_size": 10_000, "episode_step": 500, "lr": 0.001, "rank_update_interval": 200,
             "learning_starts": 1000, "target_update_interval": 1, "total_timesteps": 100_000, "hidden_dim": 1024,
             "n_hidden": 2, "normalize": False, "layer_norm": False, "train_steps": 1000, "feature_dim": 256,
             "discount": 0.99, "tau": 0.9, "value_lr": 1e-4, "policy_lr": 1e-4, "alpha": 10.0}
SORL_FLAGS = {"namespace": "tb3", "state_size": 362, "test_episodes": 10, "episode_step": 500, "action_size": 2,
              "episodes": 5000, "device": "cuda", "batch_size": 32, "lr": 0.001, "hidden_dim": 512, "n_hidden": 2,
              "normalize": False, "layer_norm": False, "train_steps": 1000, "feature_dim": 256, "discount": 0.99,
              "tau": 0.9, "value_lr": 1e-4, "policy_lr": 1e-4, "alpha": 10.0, "evaluate": False}


def _declared(script):
    """{flag: default literal} parsed from the add_argument lines of a script (no import: needs no GPU)."""
    src = open(os.path.join(ROOT, script)).read()
    out = {}
    for m in re.finditer(r'add_argument\("--(\w+)"([^\n]*)\)', src):
        rest = m.group(2)
        d = re.search(r"default=([^,)]+)", rest)
        if "store_true" in rest:
            out[m.group(1)] = False
        elif d:
            out[m.group(1)] = eval(d.group(1))
    return out


@pytest.mark.parametrize("script,flags", [("por_train.py", POR_FLAGS), ("sorl_train.py", SORL_FLAGS)])
def test_reference_flags_and_defaults_are_kept(script, flags):
    got = _declared(script)
    for k, v in flags.items():
        assert k in got, f"{script}: reference flag --{k} is missing"
        assert got[k] == v, (script, k, got[k], v)


@pytest.mark.gpu
@pytest.mark.parametrize("cmd", [
    ["por_train.py", "--episodes", "1", "--updates", "12", "--dataset_rows", "2048", "--batch_size", "64",
     "--hidden_dim", "128", "--log_every", "4"],
    ["sorl_train.py", "--episodes", "2", "--dataset_rows", "1000", "--batch_size", "96", "--hidden_dim", "64",
     "--log_every", "5", "--evaluate", "--weights_dir", "/tmp/porl_b200_cli_weights"],
    ["scripts/train_cql.py", "--dataset_rows", "4096", "--num_iterations", "25", "--log_dir", "/tmp/porl_b200_cli_logs"],
    # the literal sorl_train.py agent (FasterNet scan encoder, sorl_train.py:29-32) with the device-side shuffle
    ["sorl_train.py", "--episodes", "1", "--updates", "6", "--dataset_rows", "512", "--batch_size", "32", "--hidden_dim", "64",
     "--log_every", "3", "--backbone", "fasternet", "--device_sampler"],
])
def test_cli_runs_end_to_end(cmd):
    r = subprocess.run([sys.executable, os.path.join(ROOT, cmd[0]), *cmd[1:]], cwd=ROOT, capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "loss" in r.stdout
    if cmd[0] == "sorl_train.py":   # ragged last batch (1000 % 96) and the checkpoint cadence of sorl_train.py:73-74
        import torch
        sd = torch.load("/tmp/porl_b200_cli_weights/model_1.pt")
        assert "v_net.v1.0.weight" in sd and "policy.log_std" in sd and "v_tgt.v2.4.bias" in sd


def test_sweep_grid_assignment():
    """por_sweep.py (BASELINE config 5): replica -> (seed, tau), row-major over seeds x taus, wrapping."""
    import por_sweep
    taus, seeds = [0.5, 0.7, 0.9], [0, 1]
    got = [por_sweep.grid_point(r, taus, seeds) for r in range(8)]
    assert got == [(0, 0.5), (0, 0.7), (0, 0.9), (1, 0.5), (1, 0.7), (1, 0.9), (0, 0.5), (0, 0.7)]
    a = por_sweep.parse(["--taus", "0.6,0.8", "--seeds", "3,4,5", "--rows", "1000"])
    assert a.taus == [0.6, 0.8] and a.seeds == [3, 4, 5] and a.state_size == 365 and a.hidden_dim == 1024
    import pytest
    with pytest.raises(ValueError):
        por_sweep.grid_point(0, [], [1])
