"""Minimal stand-in for src/porl/utils/logger.py:10-137 (host-side I/O, off the update path):
same method names; CSV always, TensorBoard only when it is importable."""
from __future__ import annotations

import csv
import os


class _NullWriter:
    def add_scalar(self, *a, **k):
        pass

    def add_hparams(self, *a, **k):
        pass

    def close(self):
        pass


class Logger:
    def __init__(self, log_dir: str = "logs", tensorboard: bool = True):
        os.makedirs(log_dir, exist_ok=True)
        self.log_dir = log_dir
        self.writer = _NullWriter()
        if tensorboard:
            try:
                from torch.utils.tensorboard import SummaryWriter

                self.writer = SummaryWriter(log_dir)
            except Exception:
                self.writer = _NullWriter()
        self._csv = open(os.path.join(log_dir, "metrics.csv"), "w", newline="")
        self._rows = csv.writer(self._csv)
        self._rows.writerow(["kind", "step", "value"])

    def log_hyperparameters(self, hparams: dict) -> None:
        for k, v in hparams.items():
            self._rows.writerow([f"hparam/{k}", 0, v])

    def log_loss(self, step: int, loss: float) -> None:
        self.writer.add_scalar("Loss/Step", loss, step)
        self._rows.writerow(["loss", step, loss])

    def log_step(self, step: int, reward: float) -> None:
        self.writer.add_scalar("Reward/Step", reward, step)
        self._rows.writerow(["reward_step", step, reward])

    def log_episode(self, episode: int, reward: float, *extra) -> None:
        self.writer.add_scalar("Reward/Episode", reward, episode)
        self._rows.writerow(["reward_episode", episode, reward])

    def close(self) -> None:
        self.writer.close()
        self._csv.close()
