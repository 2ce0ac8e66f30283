#!/usr/bin/env python
"""POR offline training on B200: same flags and loop shape as the reference por_train.py:20-152
(agent.por_residual_update on minibatches of [obs | r | next_obs | done | action] rows), with the dataset
resident in HBM and the update running in libporl_b200.so.  New flags: see porl_b200/train/offline_loop.py."""
import argparse

import torch

from porl_b200 import POR
from porl_b200.train.offline_loop import add_common_flags, run


def train(args):
    device = torch.device(args.device)
    if device.type != "cuda":
        raise SystemExit("porl_b200 runs on CUDA (sm_100a) only; use the reference for --device cpu")
    # torchrun: the process group and THIS rank's device come first -- the agent allocates its parameter and
    # gradient buffers (and maps its peers' gradient buckets) at construction
    from porl_b200 import dist as pdist
    _, local, world = pdist.init_from_env()
    if world > 1 or device.index is None:
        device = torch.device("cuda", local if world > 1 else torch.cuda.current_device())
    torch.cuda.set_device(device)
    torch.manual_seed(1000 + args.seed)
    backbone = None
    if args.backbone == "fasternet":                    # BASELINE config 4: POR with the FasterNet scan encoder
        from porl_b200.agent.fasternet import FasterNet
        backbone = FasterNet(3, args.feature_dim)
    agent = POR(args, max_steps=args.train_steps, tau=args.tau, alpha=args.alpha, discount=args.discount,
                value_lr=args.value_lr, policy_lr=args.policy_lr, device=device, backbone=backbone)  # por_train.py:48-57
    out = run(args, agent, lambda a, pb: a.update_packed(pb), save_every=25, name="por")   # :96-97 every 25 episodes
    print(out)
    return out


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    # reference flags, names and defaults of por_train.py:124-150 (several are unused there too)
    parser.add_argument("--namespace", type=str, default="tb3")
    parser.add_argument("--state_size", type=int, default=365)
    parser.add_argument("--action_size", type=int, default=2)
    parser.add_argument("--episodes", type=int, default=5000)
    parser.add_argument("--device", type=str, default="cuda")
    parser.add_argument("--batch_size", type=int, default=32)
    parser.add_argument("--replay_buffer_size", type=int, default=10_000)
    parser.add_argument("--episode_step", type=int, default=500)
    parser.add_argument("--lr", type=float, default=0.001)
    parser.add_argument("--rank_update_interval", type=int, default=200)
    parser.add_argument("--learning_starts", type=int, default=1000)
    parser.add_argument("--target_update_interval", type=int, default=1)
    parser.add_argument("--total_timesteps", type=int, default=100_000)
    parser.add_argument("--hidden_dim", type=int, default=1024)
    parser.add_argument("--n_hidden", type=int, default=2)
    parser.add_argument("--normalize", action="store_true")
    parser.add_argument("--layer_norm", action="store_true")
    parser.add_argument("--train_steps", type=int, default=1_000)
    parser.add_argument("--feature_dim", type=int, default=256)
    parser.add_argument("--discount", type=float, default=0.99)
    parser.add_argument("--tau", type=float, default=0.9)
    parser.add_argument("--value_lr", type=float, default=1e-4)
    parser.add_argument("--policy_lr", type=float, default=1e-4)
    parser.add_argument("--alpha", type=float, default=10.0)
    add_common_flags(parser)
    train(parser.parse_args())
