"""CPU oracle of the scan encoder: state2costmap + FasterNet.forward_cls.  TEST INFRASTRUCTURE ONLY.

Restates, as a function of a reference-layout state_dict (no nn.Module), the forward pass of
    util/costmap.py:7-64          state2costmap (in-place clamp, one-hot scan raster, 5-pixel goal cross)
    agent/fasternet.py:428-438    FasterNet.forward_cls
    agent/fasternet.py:251-261    PatchEmbed  (Conv2d k4 s4 no bias + BatchNorm2d)
    agent/fasternet.py:141-195    MLPBlock    (x + DropPath(Conv1x1(ReLU(BN(Conv1x1(Partial_conv3(x)))))))
    agent/fasternet.py:110-138    Partial_conv3.forward_split_cat (3x3 conv on the first dim//4 channels)
    agent/fasternet.py:264-277    PatchMerging (Conv2d k2 s2 no bias + BatchNorm2d)
    agent/fasternet.py:76-93      drop_path (per-sample Bernoulli(keep)/keep scale; the mask is an INPUT here)
in train mode: BatchNorm uses batch statistics (biased variance), and updates running_mean / running_var
(unbiased variance, momentum 0.1) and num_batches_tracked -- the agents never call .eval().

The dtype follows the state_dict (float32 = the reference's arithmetic, float64 = the yardstick).
Pinned against the imported reference by tests/golden/make_golden.py -> tests/golden/encoder_*.npz.
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

from .synth import laser_rows

DIST_INC = (4.0 + 1e-4) / 256                     # costmap.py:19
ANGLE_INC = (2 * np.pi + 1e-4) / 360              # costmap.py:20
BN_EPS, BN_MOMENTUM = 1e-5, 0.1                   # nn.BatchNorm2d defaults (fasternet.py:292)
DEPTHS = (1, 2)                                   # fasternet.py:285
DROP_PATH_RATE = 0.1                              # fasternet.py:294


def drop_probs() -> list[float]:
    """fasternet.py:325-326: linspace(0, rate, sum(depths)) -> [0, 0.05, 0.1]."""
    return [x.item() for x in torch.linspace(0, DROP_PATH_RATE, sum(DEPTHS))]


def state2costmap(state: torch.Tensor) -> torch.Tensor:
    """costmap.py:7-64.  Mutates `state` (values > 8 become 0, costmap.py:17).  Returns the
    (B,3,360,256) view of a (B,360,256,3) buffer, like the reference (costmap.py:64)."""
    b = state.shape[0]
    state[state > 8.0] = 0.0
    cm = torch.zeros((b, 360, 256, 3), dtype=state.dtype)
    idx = (state[:, :360] / DIST_INC).to(torch.long)                     # trunc toward zero
    idx = torch.roll(idx, 180, 1)
    if int(idx.min()) < 0 or int(idx.max()) > 255:
        raise RuntimeError("scan value outside [0, 4.0001): the reference's scatter_ raises here")
    rows = torch.arange(360)[None, :].expand(b, 360)
    cm[torch.arange(b)[:, None].expand(b, 360), rows, idx, 0] = 1.0
    cm[:, :, 0, 0] = 0.0                                                 # costmap.py:30
    deg = torch.atan2(state[:, -1], state[:, -2])
    lim = (2 * np.pi + 2e-4) / 360
    deg = torch.clamp(deg, min=-math.pi + lim, max=math.pi - lim)
    deg = ((deg + math.pi) / ANGLE_INC).to(torch.long)
    dist = torch.linalg.norm(state[:, -2:], dim=-1)
    dist = (torch.clamp(dist, max=4 - 4.0 / 256) / DIST_INC).to(torch.long)
    ar = torch.arange(b)
    for d in (-1, 0, 1):                                                 # negative indices wrap (costmap.py:52-55)
        cm[ar, deg, (dist + d) % 256, :] = 1.0
        cm[ar, (deg + d) % 360, dist, :] = 1.0
    return cm.permute(0, 3, 1, 2)


def _bn_train(x, sd, prefix, update):
    """BatchNorm2d.forward in training mode on an NCHW tensor (batch stats; running stats updated in `sd`)."""
    w, b = sd[prefix + ".weight"], sd[prefix + ".bias"]
    mean = x.mean(dim=(0, 2, 3))
    var = x.var(dim=(0, 2, 3), unbiased=False)
    if update:
        n = x.numel() // x.shape[1]
        sd[prefix + ".running_mean"] = (1 - BN_MOMENTUM) * sd[prefix + ".running_mean"] + BN_MOMENTUM * mean
        sd[prefix + ".running_var"] = (1 - BN_MOMENTUM) * sd[prefix + ".running_var"] + BN_MOMENTUM * var * n / max(n - 1, 1)
        sd[prefix + ".num_batches_tracked"] = sd[prefix + ".num_batches_tracked"] + 1
    y = (x - mean[None, :, None, None]) / torch.sqrt(var[None, :, None, None] + BN_EPS)
    return y * w[None, :, None, None] + b[None, :, None, None]


def _block(x, sd, p, keep_scale):
    """MLPBlock.forward (fasternet.py:183-187); keep_scale: (B,) per-sample DropPath factor or None."""
    wp = sd[p + ".spatial_mixing.partial_conv3.weight"]
    c = wp.shape[0]
    mixed = torch.cat((F.conv2d(x[:, :c], wp, padding=1), x[:, c:]), 1)
    h = F.conv2d(mixed, sd[p + ".mlp.0.weight"])
    h = torch.relu(_bn_train(h, sd, p + ".mlp.1", True))
    h = F.conv2d(h, sd[p + ".mlp.3.weight"])
    if keep_scale is not None:
        h = h * keep_scale.to(h.dtype)[:, None, None, None]
    return x + h


def fasternet_forward(sd: dict, state: torch.Tensor, keep_scales=None, intermediates: dict | None = None):
    """FasterNet.forward_cls (fasternet.py:428-438).  `sd`: reference-layout state_dict (tensors of one
    float dtype; running stats and num_batches_tracked are UPDATED in place in the dict).
    `keep_scales`: [None, (B,), (B,)] per-block DropPath factors mask/keep_prob (None: all kept)."""
    dt = sd["head.weight"].dtype
    x = state2costmap(state).to(dt)
    if intermediates is not None:
        intermediates["costmap"] = x
    ks = keep_scales or [None, None, None]
    x = _bn_train(F.conv2d(x, sd["patch_embed.proj.weight"], stride=4), sd, "patch_embed.norm", True)
    x = _block(x, sd, "stages.0.blocks.0", ks[0])
    x = _bn_train(F.conv2d(x, sd["stages.1.reduction.weight"], stride=2), sd, "stages.1.norm", True)
    x = _block(x, sd, "stages.2.blocks.0", ks[1])
    x = _block(x, sd, "stages.2.blocks.1", ks[2])
    if intermediates is not None:
        intermediates["stage_out"] = x
    x = x.mean(dim=(2, 3), keepdim=True)
    x = torch.relu(F.conv2d(x, sd["avgpool_pre_head.1.weight"]))
    return F.linear(x.flatten(1), sd["head.weight"], sd["head.bias"])


def det_encoder_state_dict(seed: int, feature_dim: int = 256, dtype=torch.float32) -> dict:
    """Deterministic FasterNet(3, feature_dim) parameters (numpy RNG, independent of torch's global RNG):
    conv/linear weights N(0, 0.05) clipped at 2 sigma (larger than the reference's trunc-normal(0.02) init so
    that every BatchNorm sees O(1) variances), BN affine near (1, 0), running stats at their init values."""
    rng = np.random.default_rng(seed)
    shapes = [("patch_embed.proj.weight", (96, 3, 4, 4))]
    bns = [("patch_embed.norm", 96)]
    for stage, dim, depth in ((0, 96, 1), (2, 192, 2)):
        for blk in range(depth):
            p = f"stages.{stage}.blocks.{blk}"
            shapes += [(p + ".mlp.0.weight", (2 * dim, dim, 1, 1)), (p + ".mlp.3.weight", (dim, 2 * dim, 1, 1)),
                       (p + ".spatial_mixing.partial_conv3.weight", (dim // 4, dim // 4, 3, 3))]
            bns.append((p + ".mlp.1", 2 * dim))
        if stage == 0:
            shapes.append(("stages.1.reduction.weight", (192, 96, 2, 2)))
            bns.append(("stages.1.norm", 192))
    shapes += [("avgpool_pre_head.1.weight", (1280, 192, 1, 1)), ("head.weight", (feature_dim, 1280))]
    sd = {}
    for name, shp in shapes:
        w = np.clip(rng.normal(0.0, 0.05, shp), -0.1, 0.1)
        sd[name] = torch.tensor(w, dtype=dtype)
    sd["head.bias"] = torch.tensor(rng.normal(0.0, 0.05, (feature_dim,)), dtype=dtype)
    for name, c in bns:
        sd[name + ".weight"] = torch.tensor(1.0 + 0.1 * rng.normal(size=c), dtype=dtype)
        sd[name + ".bias"] = torch.tensor(0.1 * rng.normal(size=c), dtype=dtype)
        sd[name + ".running_mean"] = torch.zeros(c, dtype=dtype)
        sd[name + ".running_var"] = torch.ones(c, dtype=dtype)
        sd[name + ".num_batches_tracked"] = torch.tensor(0, dtype=torch.int64)
    return sd


def encoder_inputs(seed, B):
    """States (B, 362) and DropPath factors for the encoder fixtures: synthetic scan rows plus the goal-bound
    cases of the reference's own CostmapTestCase (util/costmap.py:66-150) in the first five rows."""
    st = laser_rows(B, 362, seed)[:, :362].copy()
    goals = [(-4.0, 0.0), (-4.0, 1e-4), (-4.0, -1e-4), (0.0, -4.0), (0.0, 4.0)]
    for i, (gx, gy) in enumerate(goals[:B]):
        st[i, -2], st[i, -1] = gx, gy
    g = torch.Generator().manual_seed(seed)
    keeps = [None] + [torch.bernoulli(torch.full((B,), 1 - p), generator=g) / (1 - p) for p in drop_probs()[1:]]
    keeps[1][B - 1] = 0.0                      # at least one dropped path in either block
    keeps[2][0] = 0.0
    return torch.from_numpy(st), keeps


def order_like(sd: dict, keys) -> dict:
    return {k: sd[k] for k in keys}
