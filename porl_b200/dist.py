"""Data-parallel plumbing: one process per GPU, torch.distributed (NCCL over NVLink) for the
gradient-bucket all-reduce.  The reference has no distributed code; every loss
on the path is a mean over independent samples (agent/por.py:17,106; cql_trainer.py:86,102), so the
global minibatch is split into `world` contiguous chunks, per-sample gradients are scaled by
1/B_global inside the loss kernels, and ONE sum all-reduce per optimizer group (gradients with the
loss scalar in the bucket's tail) reproduces the single-process update.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def is_dp() -> bool:
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


def world_size() -> int:
    return dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1


def rank() -> int:
    return dist.get_rank() if (dist.is_available() and dist.is_initialized()) else 0


def init_from_env(backend: str | None = None) -> tuple[int, int, int]:
    """Initialise the default process group from torchrun's environment; returns (rank, local_rank, world)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rk = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, rank=rk, world_size=world, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend, rank=rk, world_size=world)
    return rk, local, world


def allreduce_sum_(bucket: torch.Tensor) -> None:
    """In-place sum over ranks of one flat fp32 bucket (no-op single process)."""
    if is_dp():
        dist.all_reduce(bucket, op=dist.ReduceOp.SUM)


def shard_indices(global_idx: torch.Tensor, rk: int | None = None, world: int | None = None) -> torch.Tensor:
    """This rank's contiguous chunk of a global minibatch index vector (length divisible by world)."""
    rk = rank() if rk is None else rk
    world = world_size() if world is None else world
    n = global_idx.shape[0]
    if n % world != 0:
        raise ValueError(f"global batch {n} is not divisible by world size {world}")
    per = n // world
    return global_idx[rk * per:(rk + 1) * per]


def broadcast_(t: torch.Tensor, src: int = 0) -> None:
    if is_dp():
        dist.broadcast(t, src)
