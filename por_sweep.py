#!/usr/bin/env python
"""Seed x expectile-tau sweep of independent POR agents, one per GPU (BASELINE.json config 5; SURVEY.md 8e "sweeps:
replicas only").  Every process owns one B200: its own synthetic dataset replica in HBM, its own seed and tau, its
own optimizer -- there is NO data-path collective; torch.distributed (if launched under torchrun) is used once at
the end to print the table on rank 0.

    python por_sweep.py --updates 2000                                   # one replica on cuda:0
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29540 \
        por_sweep.py --rows 10000000 --updates 2000 --taus 0.5,0.6,0.7,0.8,0.9 --seeds 0,1

Replica r runs grid point r of the row-major (seed, tau) grid (wrapping around when there are more GPUs than points).
Each update = device-side shuffled index window (porl_sample_permutation) + gather + POR.por_residual_update's kernels,
replayed as one CUDA graph; the agent API, hyper-parameters and state_dict are those of agent/por.py / por_train.py.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time
from types import SimpleNamespace

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def grid_point(replica: int, taus: list[float], seeds: list[int]) -> tuple[int, float]:
    """(seed, tau) of replica `replica`: row-major over seeds x taus, wrapping."""
    if not taus or not seeds:
        raise ValueError("empty grid")
    k = replica % (len(taus) * len(seeds))
    return seeds[k // len(taus)], taus[k % len(taus)]


def parse(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--taus", type=str, default="0.5,0.6,0.7,0.8,0.9,0.9,0.9,0.9")
    p.add_argument("--seeds", type=str, default="0")
    p.add_argument("--rows", type=int, default=1_000_000, help="transitions of each replica's dataset (config 5: 10000000)")
    p.add_argument("--state_size", type=int, default=365)          # por_train.py:127
    p.add_argument("--hidden_dim", type=int, default=1024)         # por_train.py:139
    p.add_argument("--n_hidden", type=int, default=2)
    p.add_argument("--batch_size", type=int, default=256)
    p.add_argument("--alpha", type=float, default=10.0)
    p.add_argument("--updates", type=int, default=2000)
    p.add_argument("--warmup", type=int, default=20)
    p.add_argument("--gemm_mode", type=int, default=1)
    p.add_argument("--save_dir", type=str, default="", help="write weights/sweep_seed{s}_tau{t}.pt per replica")
    a = p.parse_args(argv)
    a.taus = [float(x) for x in a.taus.split(",") if x]
    a.seeds = [int(x) for x in a.seeds.split(",") if x]
    return a


def main(argv=None):
    import torch
    import bench
    from porl_b200 import POR, TransitionDataset, dist as pdist
    from porl_b200.dataloader.dataset import DeviceSampler
    a = parse(argv)
    rank, local, world = pdist.init_from_env()
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    seed, tau = grid_point(rank, a.taus, a.seeds)
    # porl_b200.dist is initialised only to gather the report: hide it from the agent so that no gradient is ever summed
    real = pdist.world_size, pdist.is_dp
    pdist.world_size, pdist.is_dp = (lambda: 1), (lambda: False)
    try:
        rows = torch.empty(a.rows, 2 * a.state_size + 4, dtype=torch.float32, device=device)
        for c, first in enumerate(range(0, a.rows, 1_000_000)):    # generated in 1M-row pieces: 10M rows are 29 GB
            n = min(1_000_000, a.rows - first)
            rows[first:first + n] = bench.synth_rows_torch(n, device, 20260105 + 1000 * seed + c, state_size=a.state_size)
        ds = TransitionDataset(rows, a.state_size, 2, device)
        del rows
        torch.cuda.empty_cache()
        args = SimpleNamespace(state_size=a.state_size, hidden_dim=a.hidden_dim, n_hidden=a.n_hidden, layer_norm=False,
                               feature_dim=256, action_size=2)
        torch.manual_seed(1000 + seed)
        agent = POR(args, max_steps=max(a.updates, 1), tau=tau, alpha=a.alpha, device=device)
        agent.ctx.set_gemm_mode(a.gemm_mode)
        sampler = DeviceSampler(len(ds), a.batch_size, device, seed=2000 + seed, drop_last=True)
        pb = ds.new_batch(a.batch_size)
        idx = torch.zeros(a.batch_size, dtype=torch.int64, device=device)
        per_epoch = len(sampler)

        def update():
            ds.gather_packed(idx, pb)
            agent.update_packed(pb)

        graph = bench.graph_of(update)

        def run(first, count):
            for u in range(first, first + count):
                sampler.indices(1 + u // per_epoch, (u % per_epoch) * a.batch_size, a.batch_size, out=idx)
                graph.replay()

        run(0, a.warmup)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        run(a.warmup, a.updates)
        e1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        v_loss, g_loss = agent.read_losses()
        if a.save_dir:
            os.makedirs(a.save_dir, exist_ok=True)
            torch.save(agent.state_dict(), os.path.join(a.save_dir, f"sweep_seed{seed}_tau{tau}.pt"))
    finally:
        pdist.world_size, pdist.is_dp = real
    mine = {"replica": rank, "gpu": local, "seed": seed, "tau": tau, "updates": a.updates,
            "updates_per_s": a.updates / (e0.elapsed_time(e1) * 1e-3), "wall_s": wall, "v_loss": v_loss, "g_loss": g_loss}
    table = [mine]
    if world > 1:
        import torch.distributed as dist
        table = [None] * world
        dist.all_gather_object(table, mine)
    if rank == 0:
        for row in table:
            print(json.dumps(row))
        print(json.dumps({"metric": "offline-RL gradient updates/sec", "config": "POR seed x tau sweep, independent "
                          f"replicas, S={a.state_size} H={a.hidden_dim} batch {a.batch_size}, {a.rows} rows per replica",
                          "replicas": world, "value": sum(r["updates_per_s"] for r in table), "unit": "updates/s",
                          "scaling": "replicas only (no collective)"}))
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
