"""A few eager CQL updates (BASELINE config 3: Q-net 362-64-128-64-5, 1M-row replay buffer), for ncu launch lists.

    python tools/cql_probe.py [BATCH] [STEPS]          (PORL_CQL_FOLD_ADAM=0: separate optimizer launch)
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from porl_b200 import CQLTrainer, QNetwork  # noqa: E402
from porl_b200.buffer.replaybuffer import ReplayBuffer  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
dev = torch.device("cuda:0")
S, A, N = 362, 5, 200_000
torch.manual_seed(3)
t = CQLTrainer(state_size=S, action_size=A, gamma=0.99, epsilon=1.0, epsilon_min=0.05, epsilon_decay=0.99,
               update_target_freq=10, device=dev, network=QNetwork, log_dir="/tmp/porl_b200_logs")
t.ctx.set_gemm_mode(1)
t.replay_buffer = ReplayBuffer(N, (S,), dev)
g = torch.Generator().manual_seed(1)
t.replay_buffer.push_batch(torch.rand(N, S, generator=g).numpy(), torch.randint(0, A, (N,), generator=g).numpy(),
                           torch.rand(N, generator=g).numpy(), torch.rand(N, S, generator=g).numpy(),
                           (torch.rand(N, generator=g) < 0.05).float().numpy())
idx = torch.randint(0, N, (B,), generator=g).to(dev)
l0 = t.ctx.launches
for _ in range(2 + steps):
    t.learn_indices(idx, hard_every=10)
torch.cuda.synchronize()
print(f"cql B={B}: {(t.ctx.launches - l0) // (2 + steps)} kernels per update, losses {t.read_losses()}")
