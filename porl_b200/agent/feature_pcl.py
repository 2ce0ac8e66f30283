"""PointMLP scan encoder with the reference's module tree and state_dict, evaluated by libporl_b200.so.

Drop-in for `agent/feature_pcl.py` (PointMLPModel, :358-420): same constructor arguments, same parameter names,
shapes and initialisation order (so the same torch seed gives the same weights), train-mode forward
`model(x)` with x (B, 3, points) on a CUDA device.  The modules below only HOLD parameters; the arithmetic --
farthest-point sampling, kNN grouping, the Conv1d/BatchNorm1d/ReLU residual stacks, max-pooling and the classifier --
runs in porl_pointmlp_forward (include/porl_b200.h).  `scan_to_points` turns (B, 362) laser-scan states into the
(B, 3, 361) point sets the model reads (the intent of feature_pcl.py:113-138, which does not run as written).

Not built: groups > 1, activations other than ReLU, eval-mode (running-statistics) forward, cross-rank BatchNorm.
"""
from __future__ import annotations

import ctypes as C

import torch
from torch import nn

from .. import _C
from .._C import check, lib


def _pad4(n: int) -> int:
    return (n + 3) // 4 * 4


class ConvBNReLU1D(nn.Module):
    """feature_pcl.py:263-275: Sequential(Conv1d(k=1), BatchNorm1d, ReLU) under `net`."""

    def __init__(self, in_channels, out_channels, kernel_size=1, bias=True, activation="relu"):
        super().__init__()
        _only_relu(activation)
        self.act = nn.ReLU(inplace=True)
        self.net = nn.Sequential(nn.Conv1d(in_channels, out_channels, kernel_size, bias=bias),
                                 nn.BatchNorm1d(out_channels), self.act)


class ConvBNReLURes1D(nn.Module):
    """feature_pcl.py:278-301 with groups = 1: net1 = Conv-BN-ReLU, net2 = Conv-BN, out = ReLU(net2(net1(x)) + x)."""

    def __init__(self, channel, kernel_size=1, groups=1, res_expansion=1.0, bias=True, activation="relu"):
        super().__init__()
        _only_relu(activation)
        if groups != 1:
            raise NotImplementedError("porl_b200 PointMLP: grouped convolutions (groups > 1) are not built")
        mid = int(channel * res_expansion)
        self.act = nn.ReLU(inplace=True)
        self.net1 = nn.Sequential(nn.Conv1d(channel, mid, kernel_size, groups=groups, bias=bias), nn.BatchNorm1d(mid), self.act)
        self.net2 = nn.Sequential(nn.Conv1d(mid, channel, kernel_size, bias=bias), nn.BatchNorm1d(channel))


class LocalGrouper(nn.Module):
    """feature_pcl.py:177-202: holds affine_alpha / affine_beta of the normalised groups."""

    def __init__(self, channel, groups, kneighbors, use_xyz=True, normalize="center", **kwargs):
        super().__init__()
        self.groups, self.kneighbors, self.use_xyz = groups, kneighbors, use_xyz
        self.normalize = normalize.lower() if normalize is not None else None
        if self.normalize not in ("center", "anchor"):
            self.normalize = None
        if self.normalize is not None:
            width = channel + (3 if use_xyz else 0)
            self.affine_alpha = nn.Parameter(torch.ones([1, 1, 1, width]))
            self.affine_beta = nn.Parameter(torch.zeros([1, 1, 1, width]))


class PreExtraction(nn.Module):
    def __init__(self, channels, out_channels, blocks=1, groups=1, res_expansion=1, bias=True, activation="relu",
                 use_xyz=True):
        super().__init__()
        self.transfer = ConvBNReLU1D(2 * channels + (3 if use_xyz else 0), out_channels, bias=bias, activation=activation)
        self.operation = nn.Sequential(*[ConvBNReLURes1D(out_channels, groups=groups, res_expansion=res_expansion, bias=bias,
                                                         activation=activation) for _ in range(blocks)])


class PosExtraction(nn.Module):
    def __init__(self, channels, blocks=1, groups=1, res_expansion=1, bias=True, activation="relu"):
        super().__init__()
        self.operation = nn.Sequential(*[ConvBNReLURes1D(channels, groups=groups, res_expansion=res_expansion, bias=bias,
                                                         activation=activation) for _ in range(blocks)])


def _only_relu(activation: str) -> None:
    if activation.lower() in ("gelu", "rrelu", "selu", "silu", "hardswish", "leakyrelu"):
        raise NotImplementedError(f"porl_b200 PointMLP: activation '{activation}' is not built (ReLU only)")


class PointMLPModel(nn.Module):
    """agent/feature_pcl.py:358-420."""

    def __init__(self, points=1024, class_num=40, embed_dim=64, groups=1, res_expansion=1.0, activation="relu", bias=True,
                 use_xyz=True, normalize="center", dim_expansion=[2, 2, 2, 2], pre_blocks=[2, 2, 2, 2],
                 pos_blocks=[2, 2, 2, 2], k_neighbors=[32, 32, 32, 32], reducers=[2, 2, 2, 2], **kwargs):
        super().__init__()
        self.stages = len(pre_blocks)
        self.class_num, self.points = class_num, points
        self.embedding = ConvBNReLU1D(3, embed_dim, bias=bias, activation=activation)
        assert len(pre_blocks) == len(k_neighbors) == len(reducers) == len(pos_blocks) == len(dim_expansion), \
            "Please check stage number consistent for pre_blocks, pos_blocks k_neighbors, reducers."
        if self.stages > _C.PM_MAX_STAGES or max(list(pre_blocks) + list(pos_blocks)) > _C.PM_MAX_BLOCKS:
            raise NotImplementedError(f"porl_b200 PointMLP: at most {_C.PM_MAX_STAGES} stages of {_C.PM_MAX_BLOCKS} blocks")
        self.local_grouper_list, self.pre_blocks_list, self.pos_blocks_list = nn.ModuleList(), nn.ModuleList(), nn.ModuleList()
        last, anchors = embed_dim, points
        for i in range(self.stages):
            out = last * dim_expansion[i]
            anchors = anchors // reducers[i]
            self.local_grouper_list.append(LocalGrouper(last, anchors, k_neighbors[i], use_xyz, normalize))
            self.pre_blocks_list.append(PreExtraction(last, out, pre_blocks[i], groups=groups, res_expansion=res_expansion,
                                                      bias=bias, activation=activation, use_xyz=use_xyz))
            self.pos_blocks_list.append(PosExtraction(out, pos_blocks[i], groups=groups, res_expansion=res_expansion,
                                                      bias=bias, activation=activation))
            last = out
        self.act = nn.ReLU(inplace=True)
        self.classifier = nn.Sequential(nn.Linear(last, 512), nn.BatchNorm1d(512), self.act, nn.Dropout(0.5),
                                        nn.Linear(512, 256), nn.BatchNorm1d(256), self.act, nn.Dropout(0.5),
                                        nn.Linear(256, class_num))
        self.use_xyz = use_xyz
        self._cache: dict = {}
        self._ws = None

    # ------------------------------------------------------------------ parameters -> porl_pointmlp_t
    def _bns(self):
        return [m for m in self.modules() if isinstance(m, nn.BatchNorm1d)]

    def _padded(self, name: str, w: torch.Tensor, rows: int, cols: int) -> torch.Tensor:
        """(rows, cols) view of a weight with rows padded to a multiple of 4 floats; cached per parameter version."""
        key = (w.data_ptr(), w._version)
        hit = self._cache.get(name)
        if hit is not None and hit[0] == key:
            return hit[1]
        ld = _pad4(cols)
        buf = torch.zeros(rows, ld, dtype=torch.float32, device=w.device)
        buf[:, :cols] = w.detach().reshape(rows, cols)
        self._cache[name] = (key, buf)
        return buf

    def _fill_conv(self, dst, seq, name: str) -> None:
        conv, bn = seq[0], seq[1]
        for t in (conv.weight, bn.weight, bn.bias, bn.running_mean, bn.running_var):
            if t.dtype != torch.float32 or not t.is_cuda:
                raise TypeError("porl_b200 PointMLP parameters must be float32 CUDA tensors (there is no CPU path)")
        cout, cin = conv.weight.shape[0], conv.weight.shape[1]
        dst.weight = self._padded(name + ".w", conv.weight, cout, cin).data_ptr()
        dst.bias = self._padded(name + ".b", conv.bias, 1, cout).data_ptr() if conv.bias is not None else None
        dst.bn.weight, dst.bn.bias = bn.weight.data_ptr(), bn.bias.data_ptr()
        dst.bn.running_mean, dst.bn.running_var = bn.running_mean.data_ptr(), bn.running_var.data_ptr()

    def _params(self, drops) -> "_C.PointMlpParams":
        p = _C.PointMlpParams()
        first = self.local_grouper_list[0]
        p.points, p.embed_dim, p.n_stages = self.points, self.embedding.net[0].weight.shape[0], self.stages
        p.use_xyz = int(bool(self.use_xyz))
        p.normalize = {None: 0, "center": 1, "anchor": 2}[first.normalize]
        p.class_num = self.class_num
        self._fill_conv(p.embedding, self.embedding.net, "embedding")
        for i in range(self.stages):
            g, pre, pos, s = self.local_grouper_list[i], self.pre_blocks_list[i], self.pos_blocks_list[i], p.stage[i]
            s.anchors, s.k = g.groups, g.kneighbors
            s.c_out = pre.transfer.net[0].weight.shape[0]
            s.c_in = (pre.transfer.net[0].weight.shape[1] - (3 if self.use_xyz else 0)) // 2
            if g.normalize is not None:
                s.affine_alpha, s.affine_beta = g.affine_alpha.data_ptr(), g.affine_beta.data_ptr()
            self._fill_conv(s.transfer, pre.transfer.net, f"pre{i}.transfer")
            s.n_pre, s.n_pos = len(pre.operation), len(pos.operation)
            for kind, ops, dst in (("pre", pre.operation, s.pre), ("pos", pos.operation, s.pos)):
                for j, blk in enumerate(ops):
                    self._fill_conv(dst[j].net1, blk.net1, f"{kind}{i}.{j}.net1")
                    self._fill_conv(dst[j].net2, blk.net2, f"{kind}{i}.{j}.net2")
                    dst[j].c_mid = blk.net1[0].weight.shape[0]
        c = self.classifier
        p.fc1_w, p.fc1_b = c[0].weight.data_ptr(), c[0].bias.data_ptr()
        p.fc2_w, p.fc2_b = c[4].weight.data_ptr(), c[4].bias.data_ptr()
        p.fc3_w = c[8].weight.data_ptr()
        p.fc3_b = self._padded("fc3.b", c[8].bias, 1, self.class_num).data_ptr()
        for dst, bn in ((p.bn1, c[1]), (p.bn2, c[5])):
            dst.weight, dst.bias = bn.weight.data_ptr(), bn.bias.data_ptr()
            dst.running_mean, dst.running_var = bn.running_mean.data_ptr(), bn.running_var.data_ptr()
        if drops is not None:
            p.drop1, p.drop2 = drops[0].data_ptr(), drops[1].data_ptr()
        return p

    # ------------------------------------------------------------------ forward
    def draw(self, B: int, device):
        """The forward's two random inputs, drawn like the reference does: the FPS start index of every stage
        (torch.randint, feature_pcl.py:81) and the classifier's Dropout(0.5) multipliers."""
        starts, n = [], self.points
        for g in self.local_grouper_list:
            starts.append(torch.randint(0, n, (B,), dtype=torch.long, device=device))
            n = g.groups
        keep = lambda w: torch.bernoulli(torch.full((B, w), 0.5, device=device)) / 0.5
        return torch.stack(starts), [keep(512), keep(256)]

    def forward(self, x: torch.Tensor, fps_start: torch.Tensor | None = None, drops=None) -> torch.Tensor:
        """x: (B, 3, points) float32 CUDA -> (B, class_num).  `fps_start` (stages, B) int64 and `drops`
        [(B,512), (B,256)] replay recorded random draws (tests); by default they are drawn on the device."""
        if not self.training:
            raise NotImplementedError("porl_b200 PointMLP: only the train-mode forward (batch statistics) is built")
        if not x.is_cuda:
            raise _C.PorlError("porl_b200 PointMLP runs on CUDA (sm_100a) devices only; there is no CPU fallback")
        if x.dim() != 3 or x.shape[1] != 3 or x.shape[2] != self.points:
            raise ValueError(f"expected x of shape (B, 3, {self.points}), got {tuple(x.shape)}")
        B, dev = int(x.shape[0]), x.device
        x = x.contiguous().float()
        if fps_start is None or drops is None:
            s, d = self.draw(B, dev)
            fps_start = s if fps_start is None else fps_start
            drops = d if drops is None else drops
        fps_start = fps_start.to(dev, torch.long).contiguous()
        drops = [t.to(dev, torch.float32).contiguous() for t in drops]
        ctx = _C.context(dev)
        p = self._params(drops)
        need = int(lib.porl_pointmlp_workspace_floats(C.byref(p), B))
        if need < 0:
            raise _C.PorlError(lib.porl_last_error().decode(errors="replace"))
        if self._ws is None or self._ws.numel() < need or self._ws.device != dev:
            self._ws = torch.empty(need, dtype=torch.float32, device=dev)
        out = torch.empty(B, self.class_num, dtype=torch.float32, device=dev)
        check(lib.porl_pointmlp_forward(ctx.handle, C.byref(p), x.data_ptr(), fps_start.data_ptr(), B, out.data_ptr(),
                                        self.class_num, self._ws.data_ptr(), self._ws.numel(), _C.stream_ptr()))
        torch._foreach_add_([m.num_batches_tracked for m in self._bns()], 1)
        return out


def scan_to_points(state: torch.Tensor) -> torch.Tensor:
    """(B, 362) laser-scan states (360 beams + goal x, y) -> (B, 3, 361) point sets for PointMLPModel(points=361)."""
    if not state.is_cuda:
        raise _C.PorlError("porl_b200 runs on CUDA (sm_100a) devices only; there is no CPU fallback")
    st = state.contiguous().float()
    out = torch.empty(st.shape[0], 3, 361, dtype=torch.float32, device=st.device)
    check(lib.porl_scan_to_points(_C.context(st.device).handle, st.data_ptr(), st.stride(0), st.shape[0], out.data_ptr(),
                                  _C.stream_ptr()))
    return out
