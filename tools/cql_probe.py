"""A few fused CQL updates at the bench shape (QNetwork 362-64-128-64-5, 1M-row buffer), for ncu / timing.

    python -m tools.cql_probe [batch] [updates] [mode: 0 FFMA tiles, 1 mma.sync 3xTF32 tiles]
"""
import sys

import torch

sys.path.insert(0, ".")


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    n_upd = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    mode = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    from porl_b200 import CQLTrainer, QNetwork
    from porl_b200.buffer.replaybuffer import ReplayBuffer
    dev = torch.device("cuda:0")
    S, n = 362, 200_000
    t = CQLTrainer(state_size=S, action_size=5, gamma=0.99, epsilon=1.0, epsilon_min=0.05, epsilon_decay=0.99,
                   update_target_freq=10, device=dev, network=QNetwork, log_dir="/tmp/porl_b200_probe_logs")
    g = torch.Generator(device=dev).manual_seed(3)
    t.replay_buffer = ReplayBuffer(n, (S,), dev)
    st = 3.5 * torch.rand(n, S, generator=g, device=dev)
    t.replay_buffer.push_batch(st, torch.randint(0, 5, (n,), generator=g, device=dev),
                               0.1 * torch.randn(n, generator=g, device=dev),
                               st + 0.05 * torch.randn(n, S, generator=g, device=dev),
                               (torch.rand(n, generator=g, device=dev) < 0.01).float())
    t.ctx.set_gemm_mode(mode)
    t.ctx.set_pdl(False)                       # per-launch event times must not include the wait for the predecessor
    t.ctx.profile(True)
    for _ in range(n_upd):
        t.learn_indices(torch.randint(0, n, (B,), generator=g, device=dev), hard_every=10)
    torch.cuda.synchronize()
    pr = t.ctx.profile_read()
    print({k: (v["launches"], round(v["ms"] / max(v["launches"], 1) * 1e3, 2)) for k, v in pr.items() if v["launches"]},
          "us per launch; losses", t.read_losses())


if __name__ == "__main__":
    main()
