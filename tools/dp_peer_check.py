"""2+ GPU check of the peer-memory gradient sum (porl_adam_step_peer) against the NCCL all-reduce path.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/dp_peer_check.py [--updates N] [--batch B] [--no-timing]

Both agents start from the same weights and see the same sharded minibatches.  Checks (asserted): (1) the peer path is
active, (2) replicas stay BITWISE identical across ranks, (3) parameters / losses agree with the NCCL path to
summation-order accuracy, (4) the data-parallel run reproduces a SINGLE-PROCESS agent that takes the same global
minibatches (batch x world rows) -- 1e-6 over the first updates, the Adam noise bound of oracle/compare.py after all
of them, (5) no timeout flag; then times both (eager and, for the peer path, graph replay).  Prints "OK" on success.
"""
from __future__ import annotations

import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from porl_b200 import TransitionDataset, dist as pdist  # noqa: E402


def build(a, device, rank, world, table, peer):
    os.environ["PORL_PEER_REDUCE"] = "1" if peer else "0"
    return bench.Runner(a, a.batch, device, rank, world, table)


def single_process(fn):
    """Run fn() with the process group hidden: a single-process agent inside the DP job."""
    real = pdist.world_size, pdist.is_dp
    pdist.world_size, pdist.is_dp = (lambda: 1), (lambda: False)
    try:
        return fn()
    finally:
        pdist.world_size, pdist.is_dp = real


def main():
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("steps", nargs="?", type=int, default=30)
    ap.add_argument("--updates", type=int, default=0)
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--no-timing", action="store_true")
    cli = ap.parse_args()
    steps = cli.updates or cli.steps
    sys.argv = sys.argv[:1]
    a = bench.parse()
    a.rows = 200000
    a.batch = cli.batch
    rank, local, world = pdist.init_from_env()
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    rows = bench.synth_rows_torch(a.rows, device, 20260102)
    table = TransitionDataset(rows, bench.S, bench.A, device)
    rp = build(a, device, rank, world, table, True)
    rn = build(a, device, rank, world, table, False)
    groups = lambda r: (r.agent.v_group, r.agent.p_group)
    assert all(g.peers is not None for g in groups(rp)), "peer path not active"
    assert all(g.peers is None for g in groups(rn))
    # the single-process yardstick: one agent, the GLOBAL minibatch (every rank computes it redundantly)
    one = single_process(lambda: bench.Runner(a, a.batch * world, device, 0, 1, table))
    assert all(g.peers is None for g in groups(one))
    idx = rp.indices(steps, 77)
    idx_global = rp.indices_global(steps, 77)                # (steps, world * batch): rank r's rows are chunk r
    flat = lambda r: torch.cat([g.params.double() for g in groups(r)])
    early = None
    for i in range(steps):
        if i == 3:      # summation order only, before a ReLU unit of some sample lands on different sides
            early = float((flat(rp) - flat(one)).norm() / flat(one).norm())
            assert early <= 1e-6, early
        rp.step(idx[i])
        single_process(lambda: one.step(idx_global[i]))
    rn.run(idx)
    vs_single = float((flat(rp) - flat(one)).norm() / flat(one).norm())
    vs_single_abs = float((flat(rp) - flat(one)).abs().max())
    assert vs_single_abs <= 2 * 1e-4 * steps, vs_single_abs
    l1 = one.agent.read_losses()
    lp, ln = rp.agent.read_losses(), rn.agent.read_losses()
    worst, worst_abs = 0.0, 0.0
    for gp, gn in zip(groups(rp), groups(rn)):
        rel = float((gp.params.double() - gn.params.double()).norm() / gn.params.double().norm())
        worst = max(worst, rel)
        worst_abs = max(worst_abs, float((gp.params - gn.params).abs().max()))
        gathered = [torch.empty_like(gp.params) for _ in range(world)]
        dist.all_gather(gathered, gp.params)
        assert all(torch.equal(gathered[0], t) for t in gathered), "replicas diverged bitwise"
    # two correct summation orders differ by rounding noise, which Adam turns into +-lr steps on weights whose
    # gradient is zero up to rounding (oracle/compare.py): every element stays within 2 * lr * steps
    assert worst_abs <= 2 * 1e-4 * steps, worst_abs
    assert abs(lp[0] - ln[0]) <= 1e-4 * abs(ln[0]) and abs(lp[1] - ln[1]) <= 1e-4 * abs(ln[1]), (lp, ln)

    def rate(r, n=200):
        ii = r.indices(n, 78)
        ms, _ = bench.timed(r, ii[:10], ii, True)
        return world * n / (ms * 1e-3)

    assert abs(lp[0] - l1[0]) <= 1e-3 * abs(l1[0]) and abs(lp[1] - l1[1]) <= 1e-3 * abs(l1[1]), (lp, l1)
    if rank == 0:
        print(f"world {world}: after {steps} updates params rel-L2 peer vs nccl {worst:.2e} (max abs {worst_abs:.2e}); losses {lp} vs {ln}; "
              f"replicas bitwise identical")
        print(f"world {world}: data parallel vs ONE process on the same global batches ({a.batch * world} rows): params rel-L2 "
              f"{early:.2e} after 3 updates (asserted <= 1e-6), {vs_single:.2e} after {steps} (max abs {vs_single_abs:.2e}, "
              f"bound 2*lr*n = {2e-4 * steps:.1e}); losses {lp} vs {l1}")
    if not cli.no_timing:
        eager_peer, eager_nccl = rate(rp), rate(rn)
        rp.capture()
        graph_peer = rate(rp)
        rp.agent.read_losses()                                   # raises on PORL_FLAG_PEER_TIMEOUT
        if rank == 0:
            print(f"updates/s (batch {a.batch}/GPU): nccl eager {eager_nccl:.0f} | peer eager {eager_peer:.0f} | "
                  f"peer graph {graph_peer:.0f}")
    rp.agent.read_losses()
    if rank == 0:
        print("OK dp_peer_check")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
