"""Epoch loop shared by por_train.py / sorl_train.py (the `for data in train_dataloader` loops of
por_train.py:67-103 and sorl_train.py:49-80), with the dataset resident in HBM.

Differences from the reference loop, all additive: the dataset is synthetic or an .npy / .pt
row table given with --dataset (the reference's CustomDataset reads 100-row files per item and has pdb traps,
dataloader/dataloader.py:18-52); losses are read back every --log_every updates instead of every update; the
Gazebo evaluation (test.py:8-53, needs ROS) is reported as unavailable; with WORLD_SIZE > 1 the minibatch is
sharded across ranks and gradients are all-reduced (porl_b200/dist.py).
"""
from __future__ import annotations

import os
import statistics
import time

import numpy as np
import torch

from .. import dist as pdist
from ..dataloader.dataset import EpochSampler, TransitionDataset, synthetic_rows


def add_common_flags(parser, default_backbone: str = "none") -> None:
    """Flags that do not exist in the reference (new names only; every reference flag keeps its name and default)."""
    parser.add_argument("--dataset", type=str, default="",
                        help=".npy or .pt file with (N, 2S+2+A) float32 rows, or a directory of the reference's "
                             "CSV / dict shards (checkpoint/)")
    parser.add_argument("--device_sampler", action="store_true",
                        help="shuffle on the device (Feistel permutation kernel) instead of torch.randperm on the host")
    parser.add_argument("--dataset_rows", type=int, default=100_000, help="rows of the synthetic dataset")
    parser.add_argument("--seed", type=int, default=0)
    parser.add_argument("--log_every", type=int, default=100, help="read the losses back every k updates")
    parser.add_argument("--updates", type=int, default=0, help="stop after this many updates (0: run all episodes)")
    parser.add_argument("--gemm_mode", type=int, default=1, help="0 fp32 FFMA, 1 tcgen05 3xTF32 contractions")
    parser.add_argument("--weights_dir", type=str, default="weights")
    parser.add_argument("--resume", type=str, default="",
                        help="training checkpoint (train_{i}.pt: weights + optimizer + LR clocks) to continue from")
    parser.add_argument("--backbone", type=str, default=default_backbone, choices=["none", "fasternet"],
                        help="scan encoder in front of the heads (sorl_train.py:29-32 always builds FasterNet; "
                             "por_train.py never does)")


def load_rows(args, row_len: int) -> np.ndarray | torch.Tensor:
    if args.dataset and os.path.isdir(args.dataset):       # the reference's shard directories (dataloader/ingest.py)
        from ..dataloader.ingest import load_directory
        return torch.from_numpy(load_directory(args.dataset, args.state_size, args.action_size))
    if args.dataset:
        rows = np.load(args.dataset) if args.dataset.endswith(".npy") else torch.load(args.dataset)
        rows = torch.as_tensor(rows, dtype=torch.float32)
        if rows.dim() != 2 or rows.shape[1] != row_len:
            raise ValueError(f"--dataset must hold (N, {row_len}) rows, got {tuple(rows.shape)}")
        return rows
    return synthetic_rows(args.dataset_rows, args.state_size, 20260101 + args.seed, args.action_size)


def run(args, agent, step_fn, save_every: int, name: str) -> dict:
    """step_fn(agent, packed_batch) enqueues one update; returns summary statistics."""
    rank, _, world = pdist.init_from_env()
    device = agent.device
    S, A = args.state_size, args.action_size
    ds = TransitionDataset(load_rows(args, 2 * S + 2 + A), S, A, device)
    if args.batch_size % world:
        raise ValueError(f"--batch_size {args.batch_size} is not divisible by the world size {world}")
    gen = torch.Generator().manual_seed(2000 + args.seed)           # same permutation on every rank
    if getattr(args, "device_sampler", False):               # same key on every rank -> same global batch
        from ..dataloader.dataset import DeviceSampler
        sampler = DeviceSampler(len(ds), args.batch_size, device, seed=2000 + args.seed)
    else:
        sampler = EpochSampler(len(ds), args.batch_size, device, generator=gen)
    agent.ctx.set_gemm_mode(args.gemm_mode)
    done, ep0 = 0, 0
    if getattr(args, "resume", ""):
        ck = torch.load(args.resume, map_location=device, weights_only=False)
        agent.load_training_state_dict(ck)
        loop = ck.get("loop")
        if loop is not None:          # data order continues where the checkpointed run stopped (saved at episode ends)
            sampler.load_state_dict(loop["sampler"])
            done, ep0 = int(loop["updates"]), int(loop["episode"]) + 1
    pb = None
    t0, done0 = time.perf_counter(), done
    last = (float("nan"), float("nan"))
    for ep in range(ep0, args.episodes):
        ep_v, ep_g = [], []
        for idx in sampler:
            if idx.numel() % world:                                   # ragged tail: drop the remainder rows
                idx = idx[: idx.numel() - idx.numel() % world]
            if idx.numel() == 0:
                continue
            local = pdist.shard_indices(idx, rank, world)
            pb = ds.gather_packed(local, pb if pb is not None and pb.B == local.numel() else None)
            step_fn(agent, pb)
            done += 1
            if done % args.log_every == 0:
                last = agent.read_losses()
                ep_v.append(last[0]); ep_g.append(last[1])
                if rank == 0:
                    dt = time.perf_counter() - t0
                    print(f"{name} episode:{ep}||update:{done}||v_loss: {statistics.fmean(ep_v):.4f}||"
                          f"g_loss: {statistics.fmean(ep_g):.4f}||{(done - done0) / dt:.1f} updates/s", flush=True)
            if args.updates and done >= args.updates:
                break
        if save_every and (ep + 1) % save_every == 0 and rank == 0:
            os.makedirs(args.weights_dir, exist_ok=True)
            torch.save(agent.state_dict(), os.path.join(args.weights_dir, f"model_{ep}.pt"))   # por_train.py:97
            ck = agent.training_state_dict()
            ck["loop"] = {"sampler": sampler.state_dict(), "updates": done, "episode": ep}
            torch.save(ck, os.path.join(args.weights_dir, f"train_{ep}.pt"))
            print("evaluate_policy: unavailable on this box (Gazebo/ROS roll-outs, test.py:8-53)", flush=True)
        if args.updates and done >= args.updates:
            break
    last = agent.read_losses()
    dt = time.perf_counter() - t0
    return {"updates": done, "seconds": dt, "updates_per_s": (done - done0) / max(dt, 1e-9), "v_loss": last[0],
            "g_loss": last[1]}
