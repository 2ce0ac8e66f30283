#!/usr/bin/env python
"""SORL offline training on B200: same flags and loop shape as the reference sorl_train.py:20-106
(agent.update(obs, actions, rewards, next_obs, dones)).  The reference always wraps the heads in a FasterNet
encoder (sorl_train.py:29-32), so --backbone fasternet is the default here and model_{i}.pt carries the
reference's backbone.* keys; --backbone none is the MLP-only variant of sorl_train_v0.py.  New flags: see porl_b200/train/offline_loop.py."""
import argparse

import torch

from porl_b200 import SORL
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
    if args.backbone == "fasternet":
        from porl_b200.agent.fasternet import FasterNet
        backbone = FasterNet(3, args.feature_dim)                                          # sorl_train.py:29
    agent = SORL(args, max_steps=args.train_steps, tau=args.tau, alpha=args.alpha, discount=args.discount,
                 value_lr=args.value_lr, policy_lr=args.policy_lr, device=device, backbone=backbone)  # sorl_train.py:31-40
    out = run(args, agent, lambda a, pb: a.update_packed(pb), save_every=2 if args.evaluate else 0, name="sorl")
    print(out)
    return out


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    # reference flags, names and defaults of sorl_train.py:84-104
    parser.add_argument("--namespace", type=str, default="tb3")
    parser.add_argument("--state_size", type=int, default=362)
    parser.add_argument("--test_episodes", type=int, default=10)
    parser.add_argument("--episode_step", type=int, default=500)
    parser.add_argument("--action_size", type=int, default=2)
    parser.add_argument("--episodes", type=int, default=5000)
    parser.add_argument("--device", type=str, default="cuda")
    parser.add_argument("--batch_size", type=int, default=32)
    parser.add_argument("--lr", type=float, default=0.001)
    parser.add_argument("--hidden_dim", type=int, default=512)
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
    parser.add_argument("--evaluate", action="store_true")
    add_common_flags(parser, default_backbone="fasternet")   # sorl_train.py:29-40: SORL(backbone=FasterNet(3, feature_dim))
    train(parser.parse_args())
