"""CPU oracle of the PointMLP scan encoder (agent/feature_pcl.py).  TEST INFRASTRUCTURE ONLY.

Restates, as a function of a reference-layout state_dict (no nn.Module), PointMLPModel.forward in train mode:
    agent/feature_pcl.py:407-420   PointMLPModel.forward (embedding, 4 x [grouper, pre, pos], max-pool, classifier)
    agent/feature_pcl.py:69-89     farthest_point_sample (the random start index is an INPUT here)
    agent/feature_pcl.py:28-48     square_distance  (-2 a.b + |a|^2 + |b|^2, in that order)
    agent/feature_pcl.py:163-174   knn_point (top-k smallest, unsorted: only the SET of neighbours matters downstream)
    agent/feature_pcl.py:204-229   LocalGrouper.forward (gather, center / anchor normalisation by the per-sample
                                   std over all groups, affine, concat with the repeated anchor feature)
    agent/feature_pcl.py:263-301   ConvBNReLU1D / ConvBNReLURes1D (groups = 1)
    agent/feature_pcl.py:304-355   PreExtraction (per-group MLP + max over the k neighbours), PosExtraction
    agent/feature_pcl.py:392-402   classifier (Linear-BN-ReLU-Dropout x2, Linear; dropout multipliers are INPUTS)
BatchNorm1d in training mode: batch statistics over (batch, length), running stats updated in the dict.
`scan_to_points` states the intent of batch_farthest_point_sample's preamble (agent/feature_pcl.py:113-138, which
does not run as written): beams -> (r cos, r sin, 0), goal -> (gx, gy, 1); beams above 8 (no return) collapse to the
origin like util/costmap.py:17.

The dtype follows the state_dict.  Pinned against the imported reference by tests/golden/make_golden.py.
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

BN_EPS, BN_MOMENTUM = 1e-5, 0.1


def scan_to_points(state: torch.Tensor) -> torch.Tensor:
    """(B, 362) -> (B, 3, 361) point set: 360 beams (degree i) and the goal, flag channel 1 for the goal."""
    b = state.shape[0]
    r = torch.where(state[:, :360] > 8.0, torch.zeros_like(state[:, :360]), state[:, :360])
    ang = torch.arange(360, dtype=state.dtype) * (math.pi / 180)
    pts = torch.zeros(b, 3, 361, dtype=state.dtype)
    pts[:, 0, :360] = torch.cos(ang)[None] * r
    pts[:, 1, :360] = torch.sin(ang)[None] * r
    pts[:, 0, 360], pts[:, 1, 360], pts[:, 2, 360] = state[:, -2], state[:, -1], 1.0
    return pts


def fps(xyz: torch.Tensor, npoint: int, start: torch.Tensor) -> torch.Tensor:
    """feature_pcl.py:69-89 with `farthest` initialised from `start` (B,) instead of torch.randint."""
    b, n, _ = xyz.shape
    cent = torch.zeros(b, npoint, dtype=torch.long)
    distance = torch.ones(b, n, dtype=xyz.dtype) * 1e10
    farthest = start.clone().long()
    ar = torch.arange(b)
    for i in range(npoint):
        cent[:, i] = farthest
        c = xyz[ar, farthest, :].view(b, 1, 3)
        distance = torch.min(distance, torch.sum((xyz - c) ** 2, -1))
        farthest = torch.max(distance, -1)[1]
    return cent


def square_distance(src, dst):
    d = -2 * torch.matmul(src, dst.permute(0, 2, 1))
    d += torch.sum(src ** 2, -1)[:, :, None]
    d += torch.sum(dst ** 2, -1)[:, None, :]
    return d


def knn(k: int, xyz, new_xyz) -> torch.Tensor:
    return torch.topk(square_distance(new_xyz, xyz), k, dim=-1, largest=False, sorted=False)[1]


def knn_margin(k: int, xyz, new_xyz) -> float:
    """Smallest relative gap between the k-th and (k+1)-th neighbour distance: fixtures keep it away from zero so
    the neighbour SET does not hinge on the last bit of a distance."""
    d = torch.sort(square_distance(new_xyz, xyz), dim=-1)[0]
    if d.shape[-1] <= k:
        return float("inf")
    return float(((d[..., k] - d[..., k - 1]) / d[..., k].abs().clamp_min(1e-12)).min())


def _gather(points, idx):
    b = points.shape[0]
    ar = torch.arange(b).view(b, *([1] * (idx.dim() - 1))).expand_as(idx)
    return points[ar, idx, :]


def _bn(x, sd, p):
    """BatchNorm1d training forward on (N, C, L) or (N, C); updates the running stats in `sd`."""
    dims = (0, 2) if x.dim() == 3 else (0,)
    mean, var = x.mean(dims), x.var(dims, unbiased=False)
    n = x.numel() // x.shape[1]
    sd[p + ".running_mean"] = (1 - BN_MOMENTUM) * sd[p + ".running_mean"] + BN_MOMENTUM * mean
    sd[p + ".running_var"] = (1 - BN_MOMENTUM) * sd[p + ".running_var"] + BN_MOMENTUM * var * n / max(n - 1, 1)
    sd[p + ".num_batches_tracked"] = sd[p + ".num_batches_tracked"] + 1
    shp = (1, -1, 1) if x.dim() == 3 else (1, -1)
    return (x - mean.view(shp)) / torch.sqrt(var.view(shp) + BN_EPS) * sd[p + ".weight"].view(shp) + sd[p + ".bias"].view(shp)


def _conv_bn(x, sd, p, relu=True):
    y = _bn(F.conv1d(x, sd[p + ".0.weight"], sd.get(p + ".0.bias")), sd, p + ".1")
    return torch.relu(y) if relu else y


def _res(x, sd, p):
    return torch.relu(_conv_bn(_conv_bn(x, sd, p + ".net1"), sd, p + ".net2", relu=False) + x)


def _count(sd, prefix):
    n = 0
    while f"{prefix}.{n}.net1.0.weight" in sd:
        n += 1
    return n


def group(sd, i, xyz, points, groups, k, start, use_xyz, normalize, record=None):
    """LocalGrouper.forward (feature_pcl.py:204-229) for stage i."""
    b = xyz.shape[0]
    fidx = fps(xyz, groups, start)
    new_xyz, new_points = _gather(xyz, fidx), _gather(points, fidx)
    idx = knn(k, xyz, new_xyz)
    if record is not None:
        record.append({"fps": fidx, "knn": idx, "margin": knn_margin(k, xyz, new_xyz)})
    gp = _gather(points, idx)
    if use_xyz:
        gp = torch.cat([gp, _gather(xyz, idx)], dim=-1)
    if normalize:
        if normalize == "center":
            mean = gp.mean(dim=2, keepdim=True)
        else:
            mean = (torch.cat([new_points, new_xyz], dim=-1) if use_xyz else new_points).unsqueeze(-2)
        std = torch.std((gp - mean).reshape(b, -1), dim=-1, keepdim=True)[:, :, None, None]
        gp = (gp - mean) / (std + 1e-5)
        gp = sd[f"local_grouper_list.{i}.affine_alpha"] * gp + sd[f"local_grouper_list.{i}.affine_beta"]
    out = torch.cat([gp, new_points.view(b, groups, 1, -1).repeat(1, 1, k, 1)], dim=-1)
    return new_xyz, out


def pointmlp_forward(sd: dict, x: torch.Tensor, cfg: dict, fps_start: torch.Tensor, drops=None, record=None):
    """PointMLPModel.forward (feature_pcl.py:407-420), train mode.  x: (B, 3, N); fps_start: (stages, B) long;
    drops: [(B,512), (B,256)] multipliers mask/(1-p) or None; cfg: points, k_neighbors, reducers, use_xyz, normalize."""
    xyz = x.permute(0, 2, 1)
    b = x.shape[0]
    h = _conv_bn(x, sd, "embedding.net")
    anchors = cfg["points"]
    for i, (k, red) in enumerate(zip(cfg["k_neighbors"], cfg["reducers"])):
        anchors //= red
        xyz, g = group(sd, i, xyz, h.permute(0, 2, 1), anchors, k, fps_start[i], cfg["use_xyz"], cfg["normalize"], record)
        bb, n, s, d = g.shape
        t = g.permute(0, 1, 3, 2).reshape(-1, d, s)
        t = _conv_bn(t, sd, f"pre_blocks_list.{i}.transfer.net")
        for j in range(_count(sd, f"pre_blocks_list.{i}.operation")):
            t = _res(t, sd, f"pre_blocks_list.{i}.operation.{j}")
        h = F.adaptive_max_pool1d(t, 1).view(bb * n, -1).reshape(bb, n, -1).permute(0, 2, 1)
        for j in range(_count(sd, f"pos_blocks_list.{i}.operation")):
            h = _res(h, sd, f"pos_blocks_list.{i}.operation.{j}")
    h = F.adaptive_max_pool1d(h, 1).squeeze(-1)
    if record is not None:
        record.append({"pooled": h})
    for lin, bn, d in (("classifier.0", "classifier.1", 0), ("classifier.4", "classifier.5", 1)):
        h = torch.relu(_bn(F.linear(h, sd[lin + ".weight"], sd[lin + ".bias"]), sd, bn))
        if drops is not None:
            h = h * drops[d].to(h.dtype)
    return F.linear(h, sd["classifier.8.weight"], sd["classifier.8.bias"])


def randomize_(sd: dict, seed: int) -> dict:
    """Make every BatchNorm / affine parameter non-trivial (the default init is weight 1, bias 0)."""
    rng = np.random.default_rng(seed)
    for kname, v in sd.items():
        if kname.endswith("num_batches_tracked") or v.dim() == 0:
            continue
        if ".1.weight" in kname or ".5.weight" in kname or "affine_alpha" in kname:
            if v.dim() == 1 or "affine" in kname:
                sd[kname] = torch.tensor(1.0 + 0.2 * rng.normal(size=tuple(v.shape)), dtype=v.dtype)
        if (".1.bias" in kname or ".5.bias" in kname or "affine_beta" in kname) and (v.dim() == 1 or "affine" in kname):
            sd[kname] = torch.tensor(0.1 * rng.normal(size=tuple(v.shape)), dtype=v.dtype)
    return sd
