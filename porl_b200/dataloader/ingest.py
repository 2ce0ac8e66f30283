"""Loaders for the reference's on-disk dataset formats into the (N, 2S+2+A) float32 row table that
TransitionDataset keeps in HBM.  Host-side I/O, run once before training.

Formats (all written by the reference's own scripts):
  * CSV shards  checkpoint/dataset_{k}.csv : 100 rows per file, flattened, 728 (S=362) or 734 (S=365) values per
    row [obs(S) | reward | next_obs(S) | done | action(2)]; read by dataloader/dataloader_bk.py:14-24 with
    np.loadtxt(delimiter=',').reshape(100, -1) and cast to float32 (dataloader_bk.py:32).
  * dict shards checkpoint/dataset_{rank}_{k}.pt : torch.save of a list of 100 dicts
    {'state', 'next_state', 'reward', 'done', 'truncated'} (collect.py:46-56).  No action is stored; the action
    columns are filled with zeros (usable by POR, which does not read actions: por_train.py:79).
  * A*-labelled shards checkpoint/a_star/dataset_{k}.pt : (100, 361) float32 tensors [scan(360) | label]
    (preprocess.py:76-91), returned as they are (they are not transition rows).
"""
from __future__ import annotations

import glob
import os
import re

import numpy as np
import torch


def _natural(path: str):
    return [int(t) if t.isdigit() else t for t in re.split(r"(\d+)", os.path.basename(path))]


def load_csv_shards(directory: str, state_size: int, action_size: int = 2, pattern: str = "dataset_*.csv") -> np.ndarray:
    """(N, 2S+2+A) float32 rows from the CSV shards, files in natural order, rows in file order."""
    row_len = 2 * state_size + 2 + action_size
    files = sorted(glob.glob(os.path.join(directory, pattern)), key=_natural)
    if not files:
        raise FileNotFoundError(f"no {pattern} under {directory}")
    out = []
    for f in files:
        data = np.loadtxt(f, delimiter=",", dtype=float)              # dataloader_bk.py:18
        if data.size % row_len:
            raise ValueError(f"{f}: {data.size} values is not a multiple of the row length {row_len}")
        out.append(data.reshape(-1, row_len).astype(np.float32))      # dataloader_bk.py:23,32
    return np.concatenate(out, 0)


def load_dict_shards(directory: str, state_size: int, action_size: int = 2, pattern: str = "dataset_*_*.pt") -> np.ndarray:
    """(N, 2S+2+A) float32 rows from collect.py's lists of transition dicts (action columns zero)."""
    files = sorted(glob.glob(os.path.join(directory, pattern)), key=_natural)
    if not files:
        raise FileNotFoundError(f"no {pattern} under {directory}")
    rows = []
    for f in files:
        for d in torch.load(f, weights_only=False):
            s = np.asarray(d["state"], dtype=np.float32).reshape(-1)
            s2 = np.asarray(d["next_state"], dtype=np.float32).reshape(-1)
            if s.size != state_size or s2.size != state_size:
                raise ValueError(f"{f}: state of {s.size} values, expected {state_size}")
            rows.append(np.concatenate([s, [np.float32(d["reward"])], s2, [np.float32(bool(d["done"]))],
                                        np.zeros(action_size, np.float32)]))
    return np.stack(rows).astype(np.float32)


def load_astar_shards(directory: str, pattern: str = "dataset_*.pt") -> torch.Tensor:
    """(N, 361) float32 [scan(360) | A* label] rows (preprocess.py:85-91)."""
    files = sorted(glob.glob(os.path.join(directory, pattern)), key=_natural)
    if not files:
        raise FileNotFoundError(f"no {pattern} under {directory}")
    return torch.cat([torch.load(f, weights_only=False).reshape(-1, 361).to(torch.float32) for f in files], 0)


def load_directory(directory: str, state_size: int, action_size: int = 2) -> np.ndarray:
    """Transition rows from whichever of the two transition formats the directory holds."""
    if glob.glob(os.path.join(directory, "dataset_*.csv")):
        return load_csv_shards(directory, state_size, action_size)
    return load_dict_shards(directory, state_size, action_size)
