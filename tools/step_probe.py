"""A few eager updates at one batch size, for `ncu --metrics gpu__time_duration.sum --cache-control none` launch lists.

    python tools/step_probe.py [sorl|por] BATCH [STEPS] [GEMM_MODE]

Builds the bench's dataset (100k rows here) and agent, runs 2 warm-up updates and STEPS eager ones (gather + update),
prints how many kernels one update launches."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from porl_b200 import POR, SORL, TransitionDataset  # noqa: E402
from types import SimpleNamespace  # noqa: E402

kind = sys.argv[1] if len(sys.argv) > 1 else "sorl"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
mode = int(sys.argv[4]) if len(sys.argv) > 4 else 1
dev = torch.device("cuda:0")
S, H = (362, 512) if kind == "sorl" else (365, 1024)
rows = bench.synth_rows_torch(100_000, dev, 20260102, state_size=S)
ds = TransitionDataset(rows, S, 2, dev)
args = SimpleNamespace(state_size=S, hidden_dim=H, n_hidden=2, layer_norm=False, feature_dim=256, action_size=2)
torch.manual_seed(1000)
agent = (SORL if kind == "sorl" else POR)(args, max_steps=1_000_000, tau=0.9, alpha=10.0, device=dev)
agent.ctx.set_gemm_mode(mode)
pb = ds.new_batch(B)
g = torch.Generator(device=dev).manual_seed(1)
idx = torch.zeros(B, dtype=torch.int64, device=dev)
per = 0
for i in range(2 + steps):
    torch.randint(0, len(ds), (B,), generator=g, device=dev, out=idx)
    l0 = agent.ctx.launches
    ds.gather_packed(idx, pb)
    agent.update_packed(pb)
    per = agent.ctx.launches - l0
torch.cuda.synchronize()
print(f"{kind} B={B}: {per} kernels per update, losses {agent.read_losses()}")
