"""Parameter containers with the reference's state_dict layout, backed by flat HBM buffers.

Every network of the hot path is an MLP (util/util.py:29-47).  The C library works on ONE flat fp32
buffer per optimizer group (so Adam, the Polyak update and the DP all-reduce are single launches),
with weight rows padded to 32 floats.  The classes below expose that buffer as `nn.Parameter`
views under the reference's names -- `v1.0.weight`, `net.4.bias`, `log_std`, `model.2.weight` -- so
`state_dict()` / `load_state_dict()` interchange with the reference classes key for key
(agent/value_functions.py:31-42, agent/policy.py:12-49, src/porl/net/q_network.py:8-30).

PyTorch here is plumbing (allocation, views, checkpoint I/O); all arithmetic runs in libporl_b200.so.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Sequence

import torch
from torch import nn

from . import _C
from . import dist as pdist
from ._C import check, lib

GRAD_TAIL = 8  # floats after the gradients in a group's bucket: [loss, aux...] ride the DP all-reduce


def pad_to(x: int, m: int = _C.PAD) -> int:
    return (x + m - 1) // m * m


class FlatMLP:
    """The C-side plan of `n_nets` structurally identical MLPs plus the geometry of their flat buffer."""

    def __init__(self, ctx: _C.Context, dims: Sequence[int], n_nets: int = 1, layer_norm: bool = False,
                 tanh_out: bool = False):
        if len(dims) < 2:
            raise ValueError("MLP requires at least two dims (input and output)")  # util/util.py:31
        if len(dims) - 1 > _C.MAX_LAYERS:
            raise ValueError(f"at most {_C.MAX_LAYERS} linear layers")
        self.ctx, self.dims, self.n_nets = ctx, [int(d) for d in dims], int(n_nets)
        self.layer_norm, self.tanh_out = bool(layer_norm), bool(tanh_out)
        desc = _C.MlpDesc()
        desc.n_nets, desc.n_layers = self.n_nets, len(dims) - 1
        for i, d in enumerate(self.dims):
            desc.dims[i] = d
        desc.layer_norm, desc.out_act = int(self.layer_norm), int(self.tanh_out)
        h = C.c_void_p()
        check(lib.porl_mlp_create(ctx.handle, C.byref(desc), C.byref(h)))
        self.handle = h
        self.n_layers = len(dims) - 1
        self.n_floats = int(lib.porl_mlp_param_floats(h))
        self.out_ld = int(lib.porl_mlp_out_ld(h))
        self.in_ld = pad_to(self.dims[0])

    def __del__(self):
        try:
            lib.porl_mlp_destroy(self.handle)
        except Exception:
            pass

    def saved_floats(self, batch: int) -> int:
        return int(lib.porl_mlp_saved_floats(self.handle, batch))

    def layout(self, net: int, layer: int, kind: int):
        off, ld = C.c_int64(), C.c_int64()
        check(lib.porl_mlp_param_layout(self.handle, net, layer, kind, C.byref(off), C.byref(ld)))
        return off.value, ld.value

    def view(self, flat: torch.Tensor, net: int, layer: int, kind: int) -> torch.Tensor:
        """Logical tensor (weight (out, in) strided over padded rows; vectors (out,)) inside `flat`."""
        off, ld = self.layout(net, layer, kind)
        out_f, in_f = self.dims[layer + 1], self.dims[layer]
        if kind == _C.PARAM_WEIGHT:
            return flat[off:off + out_f * ld].view(out_f, ld)[:, :in_f]
        return flat[off:off + out_f]


class ParamGroup:
    """One optimizer group: flat parameters, gradient bucket, Adam moments and the device-side clock.

    Mirrors torch.optim.Adam(params, lr) [+ CosineAnnealingLR(T_max)] of agent/por.py:61-65.
    """

    def __init__(self, ctx: _C.Context, n_floats: int, device, lr: float, t_max: int = 0,
                 betas=(0.9, 0.999), eps: float = 1e-8, with_target: bool = False, sharded: bool = False):
        """with_target: the group owns a target-net buffer of the same geometry (`self.target`); sharded: in a
        data-parallel job take the two-shot step (every rank updates one shard and stores it into all ranks' buffers,
        porl_peer_group_t.params) -- parameters and target then live in peer-mappable memory."""
        self.ctx, self.device = ctx, torch.device(device)
        self.n = pad_to(n_floats, 4)
        z = lambda k: torch.zeros(k, dtype=torch.float32, device=self.device)
        self.m, self.v = z(self.n), z(self.n)
        self.peers = None
        self.bucket = self.params = self.target = None
        if pdist.is_dp() and self.device.type == "cuda" and os.environ.get("PORL_PEER_REDUCE", "1") != "0":
            two_shot = sharded and os.environ.get("PORL_PEER_MODE", "2") != "1"
            self._init_peers(two_shot, with_target)
        if self.params is None:
            self.params = z(self.n)
        if with_target and self.target is None:
            self.target = z(self.n)
        if self.bucket is None:
            self.bucket = z(self.n + GRAD_TAIL)      # gradients | loss tail
        self.grads = self.bucket[:self.n]
        self._local_tail = self.bucket[self.n:]
        self.betas, self.eps = betas, eps
        self.base_lr, self.t_max = float(lr), int(t_max)
        self.state = torch.zeros(C.sizeof(_C.OptState), dtype=torch.uint8, device=self.device)
        self.reset_clock(step=0, lr=float(lr))

    def _init_peers(self, two_shot: bool, with_target: bool) -> None:
        """Collective (every rank builds its groups in the same order): put the bucket in peer-mappable memory and
        map the other ranks' buckets, so that `step()` sums the gradients inside the Adam kernel
        (porl_adam_step_peer).  two_shot: parameters and target net are peer-mapped too and every rank updates one shard
        (traffic ~2-3x the bucket whatever the world size).  One-shot reads cost (world-1)x the bucket per rank: big
        buckets then fall back to the NCCL all-reduce, as does a node without a P2P path."""
        from . import peer

        world = pdist.world_size()
        nbytes = (self.n + GRAD_TAIL) * 4
        if world > _C.MAX_PEERS or (not two_shot and (world - 1) * nbytes > peer.MAX_PEER_READ_BYTES):
            return
        bucket = peer.PeerBuffer(self.ctx, nbytes)
        pad = peer.PeerBuffer(self.ctx, _C.PEER_PAD_WORDS * 4)
        params = peer.PeerBuffer(self.ctx, self.n * 4) if two_shot else None
        target = peer.PeerBuffer(self.ctx, self.n * 4) if (two_shot and with_target) else None
        self.peers = peer.exchange(self.ctx, bucket, pad, GRAD_TAIL, params=params, target=target)
        if self.peers is not None:
            self.bucket = bucket.tensor(torch.float32)
            if two_shot:
                self.params = params.tensor(torch.float32)
                if target is not None:
                    self.target = target.tensor(torch.float32)

    @property
    def tail(self) -> torch.Tensor:
        """[loss, aux...] of the last update, summed over ranks once `step()` has run."""
        return self.peers.tail_out if self.peers is not None else self._local_tail

    def reset_clock(self, step: int, lr: float) -> None:
        st = _C.OptState(step=step, lr=lr, base_lr=self.base_lr, t_max=self.t_max, blocks_done=0, pad_=0)
        host = torch.frombuffer(bytearray(bytes(st)), dtype=torch.uint8)
        self.state.copy_(host)

    def read_clock(self) -> _C.OptState:
        """Synchronising read of the device-side optimizer clock."""
        raw = bytes(self.state.cpu().numpy().tobytes())
        return _C.OptState.from_buffer_copy(raw)

    @property
    def loss_ptr(self) -> int:
        return self._local_tail.data_ptr()

    def step(self, target: torch.Tensor | None = None, ema_beta: float = 0.0, hard_every: int = 0) -> None:
        """Sum the bucket over the data-parallel ranks, then Adam (+ target update).  With peer-mapped buckets both
        happen in one kernel; otherwise one NCCL all-reduce precedes the single-rank kernel."""
        tgt = target.data_ptr() if target is not None else None
        if self.peers is not None and getattr(self.peers, "sharded", False) and target is not None \
                and target.data_ptr() != (self.target.data_ptr() if self.target is not None else 0):
            raise ValueError("two-shot data-parallel step: the target net must be the group's own peer-mapped buffer "
                             "(ParamGroup(with_target=True).target)")
        if self.peers is not None:
            check(lib.porl_adam_step_peer(self.ctx.handle, self.params.data_ptr(), C.byref(self.peers.struct),
                                          self.m.data_ptr(), self.v.data_ptr(), self.n, self.state.data_ptr(),
                                          self.betas[0], self.betas[1], self.eps, tgt, ema_beta, hard_every,
                                          _C.stream_ptr()))
            return
        pdist.allreduce_sum_(self.bucket)
        check(lib.porl_adam_step(self.ctx.handle, self.params.data_ptr(), self.grads.data_ptr(), self.m.data_ptr(),
                                 self.v.data_ptr(), self.n, self.state.data_ptr(), self.betas[0], self.betas[1],
                                 self.eps, target.data_ptr() if target is not None else None, ema_beta,
                                 hard_every, _C.stream_ptr()))

    def c_group(self) -> _C.OptGroup:
        """porl_opt_group_t view of this group for the whole-update entry points (porl_sorl_step / porl_por_step)."""
        g = _C.OptGroup()
        g.params, g.grads, g.m, g.v = self.params.data_ptr(), self.grads.data_ptr(), self.m.data_ptr(), self.v.data_ptr()
        g.n, g.state, g.loss = self.n, self.state.data_ptr(), self.loss_ptr
        g.beta1, g.beta2, g.eps = self.betas[0], self.betas[1], self.eps
        g.peers = C.pointer(self.peers.struct) if self.peers is not None else None
        return g

    # --- torch.optim-like surface used by trainers / checkpoint extensions -------------------
    @property
    def param_groups(self):
        return [{"lr": self.read_clock().lr, "betas": self.betas, "eps": self.eps, "initial_lr": self.base_lr}]

    def _full_moments(self):
        """(exp_avg, exp_avg_sq) of the whole group.  In the two-shot data-parallel step every rank keeps the moments of
        its own shard only: the full vectors are assembled with one sum all-reduce of the masked shards (collective)."""
        if self.peers is None or not getattr(self.peers, "sharded", False):
            return self.m.clone(), self.v.clone()
        from . import peer
        s0, s1 = peer.shard_range(self.n, self.peers.world, self.peers.rank)
        out = []
        for t in (self.m, self.v):
            full = torch.zeros_like(t)
            full[s0:s1] = t[s0:s1]
            pdist.allreduce_sum_(full)
            out.append(full)
        return out[0], out[1]

    def state_dict(self):
        c = self.read_clock()
        m, v = self._full_moments()
        return {"step": c.step, "lr": c.lr, "base_lr": self.base_lr, "t_max": self.t_max, "exp_avg": m, "exp_avg_sq": v}

    def load_state_dict(self, sd) -> None:
        self.m.copy_(sd["exp_avg"])
        self.v.copy_(sd["exp_avg_sq"])
        self.reset_clock(int(sd["step"]), float(sd["lr"]))

    def zero_grad(self, set_to_none: bool = True) -> None:  # gradients are fully overwritten every update
        pass


class CosineSchedule:
    """Stand-in for torch CosineAnnealingLR (agent/por.py:65): the recursion runs on the device
    inside the Adam kernel, so `step()` is a no-op kept for trainer compatibility."""

    def __init__(self, group: ParamGroup):
        self.group = group
        self.T_max = group.t_max

    def step(self) -> None:
        pass

    def get_last_lr(self):
        return [self.group.read_clock().lr]

    @property
    def last_epoch(self) -> int:
        return int(self.group.read_clock().step)


class _Tensors(nn.Module):
    """A parameter-only node (Linear / LayerNorm slot of the reference's Sequential)."""

    def __init__(self, weight: torch.Tensor, bias: torch.Tensor, trainable: bool):
        super().__init__()
        self.weight = nn.Parameter(weight, requires_grad=trainable)
        self.bias = nn.Parameter(bias, requires_grad=trainable)

    def forward(self, *_):
        raise RuntimeError("porl_b200 layers are evaluated by the fused CUDA MLP kernels, not layer by layer")


class MLPParams(nn.Module):
    """Views of one net of a FlatMLP under the reference Sequential's child indices
    (Linear at 0,2,4 -- or 0,3,6 with LayerNorm at 1,4; util/util.py:34-41)."""

    def __init__(self, plan: FlatMLP, flat: torch.Tensor, net: int, grads: torch.Tensor | None = None,
                 trainable: bool = True, init: bool = True):
        super().__init__()
        self._slots = []
        idx = 0
        for layer in range(plan.n_layers):
            hidden = layer < plan.n_layers - 1
            w = plan.view(flat, net, layer, _C.PARAM_WEIGHT)
            b = plan.view(flat, net, layer, _C.PARAM_BIAS)
            if init:
                ref = nn.Linear(plan.dims[layer], plan.dims[layer + 1])   # torch default init, global RNG
                w.copy_(ref.weight.detach())
                b.copy_(ref.bias.detach())
            node = _Tensors(w, b, trainable)
            self.add_module(str(idx), node)
            self._slots.append((node, layer, _C.PARAM_WEIGHT, _C.PARAM_BIAS))
            idx += 1
            if hidden and plan.layer_norm:
                g = plan.view(flat, net, layer, _C.PARAM_LN_WEIGHT)
                bb = plan.view(flat, net, layer, _C.PARAM_LN_BIAS)
                if init:
                    g.fill_(1.0)
                    bb.zero_()
                ln = _Tensors(g, bb, trainable)
                self.add_module(str(idx), ln)
                self._slots.append((ln, layer, _C.PARAM_LN_WEIGHT, _C.PARAM_LN_BIAS))
                idx += 1
            if hidden:
                idx += 1  # the ReLU slot
        if grads is not None and trainable:
            for node, layer, kw, kb in self._slots:
                node.weight.grad = plan.view(grads, net, layer, kw)
                node.bias.grad = plan.view(grads, net, layer, kb)

    def __getitem__(self, i: int):
        return getattr(self, str(i))


class Workspace:
    """Per-(agent, batch size) device scratch, allocated once and reused every update."""

    def __init__(self, device):
        self.device = torch.device(device)
        self._bufs: dict = {}

    def get(self, key, n_floats: int, dtype=torch.float32) -> torch.Tensor:
        k = (key, dtype)
        t = self._bufs.get(k)
        if t is None or t.numel() < n_floats:
            t = torch.zeros(max(int(n_floats), 4), dtype=dtype, device=self.device)
            self._bufs[k] = t
        return t


def mlp_call(plan: FlatMLP, params: torch.Tensor, x: torch.Tensor, ldx: int, saved: torch.Tensor,
             out: torch.Tensor, inference: bool = False) -> _C.MlpCall:
    return _C.MlpCall(mlp=plan.handle.value, params=params.data_ptr(), x=x.data_ptr(), ldx=ldx,
                      saved=saved.data_ptr(), out=out.data_ptr(), inference=int(inference), pad_=0)


def mlp_forward(ctx: _C.Context, calls: Sequence[_C.MlpCall], batch: int) -> None:
    arr = (_C.MlpCall * len(calls))(*calls)
    check(lib.porl_mlp_forward(ctx.handle, len(calls), arr, batch, _C.stream_ptr()))


def mlp_backward(ctx: _C.Context, call: _C.MlpCall, batch: int, dout: torch.Tensor, grads: torch.Tensor) -> None:
    check(lib.porl_mlp_backward(ctx.handle, C.byref(call), batch, dout.data_ptr(), grads.data_ptr(),
                                _C.stream_ptr()))


def pack_rows(ctx: _C.Context, src: torch.Tensor, n: int, dst: torch.Tensor, ld_dst: int) -> None:
    """(B, n) float32 rows with arbitrary row stride -> padded (B, ld_dst) rows (pad columns zeroed)."""
    if src.dtype != torch.float32:
        src = src.float()
    if src.dim() == 1:
        src = src.unsqueeze(1)
    if src.stride(1) != 1:
        src = src.contiguous()
    check(lib.porl_pack_rows(ctx.handle, src.data_ptr(), src.stride(0), src.shape[0], n, dst.data_ptr(), ld_dst,
                             _C.stream_ptr()))


class InferenceMixin:
    """Forward of a parameter container through the fused kernels, for API compatibility
    (TwinV.forward / policy.forward are callable in the reference)."""

    def _infer(self, plan: FlatMLP, flat: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
        ctx = plan.ctx
        if x.dim() == 1:
            x = x.unsqueeze(0)
        x = x.to(flat.device, torch.float32)
        batch = x.shape[0]
        xin = torch.zeros(batch, plan.in_ld, dtype=torch.float32, device=flat.device)
        pack_rows(ctx, x, plan.dims[0], xin, plan.in_ld)
        saved = torch.empty(plan.saved_floats(batch), dtype=torch.float32, device=flat.device)
        out = torch.empty(plan.n_nets, batch, plan.out_ld, dtype=torch.float32, device=flat.device)
        mlp_forward(ctx, [mlp_call(plan, flat, xin, plan.in_ld, saved, out, inference=True)], batch)
        return out[:, :, :plan.dims[-1]]
