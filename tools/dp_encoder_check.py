"""torchrun --nproc-per-node 2 tools/dp_encoder_check.py: SORL + FasterNet, one update of a global batch split over
the ranks (NCCL), against the same update in a single process on rank 0.  Prints the loss / parameter agreement."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from types import SimpleNamespace  # noqa: E402

from porl_b200 import SORL, dist as pdist  # noqa: E402
from porl_b200.agent.fasternet import FasterNet  # noqa: E402
from porl_b200.dataloader.dataset import synthetic_rows  # noqa: E402

rank, local, world = pdist.init_from_env()
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
B, S, A = 64, 362, 2
args = SimpleNamespace(state_size=S, hidden_dim=128, n_hidden=2, layer_norm=False, feature_dim=256, action_size=A)
rows = torch.from_numpy(synthetic_rows(B, S, 5, A)).to(dev)


def agent_and_masks():
    torch.manual_seed(7)
    agent = SORL(args, max_steps=1000, tau=0.8, alpha=3.0, device=dev, backbone=FasterNet(3, 256))
    g = torch.Generator().manual_seed(3)
    keeps = [[None] + [torch.bernoulli(torch.full((B,), 1 - p), generator=g) / (1 - p) for p in (0.05, 0.1)] for _ in range(2)]
    return agent, keeps


def update(agent, keeps, sl):
    agent.backbone.replay_keep_scales([[None if k is None else k[sl] for k in ks] for ks in keeps])
    r = rows[sl]
    return agent.update(r[:, :S], r[:, -A:], r[:, S], r[:, S + 1:2 * S + 1], r[:, 2 * S + 1])


agent, keeps = agent_and_masks()
per = B // world
got = update(agent, keeps, slice(rank * per, (rank + 1) * per))
torch.cuda.synchronize()
dist.barrier()
if rank == 0:
    # the same update in one process: take this rank out of the group's view by overriding both reductions
    single, keeps1 = agent_and_masks()
    single.backbone.sync_batch_stats(1, None)
    import porl_b200.agent._expectile as ex
    saved = (ex.pdist.allreduce_sum_, ex.pdist.world_size)
    ex.pdist.allreduce_sum_, ex.pdist.world_size = (lambda t: None), (lambda: 1)
    want = update(single, keeps1, slice(0, B))
    ex.pdist.allreduce_sum_, ex.pdist.world_size = saved
    err = max(float((a - b).abs().max() / b.abs().max().clamp_min(1e-12))
              for (k, a), b in zip(agent.state_dict().items(), single.state_dict().values()) if a.is_floating_point())
    print(f"DP{world} losses {got} vs single-process {want}; max relative parameter difference {err:.3e}", flush=True)
    assert abs(got[0] - want[0]) <= 2e-5 * abs(want[0]) and abs(got[1] - want[1]) <= 2e-5 * abs(want[1]) and err < 1e-4
dist.barrier()
dist.destroy_process_group()
