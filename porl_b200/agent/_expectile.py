"""Shared update step of POR and SORL: twin expectile value step, Polyak target update, then an
advantage-weighted Gaussian log-likelihood policy step (agent/por.py:73-112 == agent/sorl.py:78-128
up to the weight formula, the regression target and the tanh head).

Kernel sequence of one update (all on the caller's current CUDA stream, no host sync inside):
   1  grouped MLP forward : target TwinV(next_obs) + online TwinV(obs)          [K2, 3 launches]
   2  expectile loss      : TD target, twin loss, dL/dv                          [K4]
   3  MLP backward        : dW/db of both value nets into the group's bucket     [K3, 5 launches]
      (DP: all-reduce of the bucket -- gradients + loss scalar)
   4  Adam + Polyak       : value params, moments, target net in one pass       [K7]
   5  grouped MLP forward : online TwinV(obs) with the NEW weights + policy(obs) [K2]
   6  AWR Gaussian NLL    : weight, NLL, dL/dmean (tanh' folded), dL/dlog_std    [K5]
   7  MLP backward        : policy dW/db                                         [K3]
      (DP: all-reduce)
   8  Adam                : policy params (+ device-side cosine LR recursion)    [K7]
"""
from __future__ import annotations

import ctypes as C
import os

import torch
from torch import nn

from .. import _C, dist as pdist
from .._C import check, lib
from ..dataloader.dataset import PackedBatch
from ..nets import (CosineSchedule, ParamGroup, Workspace, mlp_backward, mlp_call, mlp_forward, pack_rows, pad_to)
from .policy import BoundedGaussianPolicy, GaussianPolicy
from .value_functions import TwinV


class ExpectileAgent(nn.Module):
    # subclasses set these
    _value_name = _target_name = _policy_name = ""
    _awr_mode = _C.AWR_POR
    _bounded = False

    def _build(self, args, max_steps, tau, alpha, backbone, device, value_lr, policy_lr, discount, beta,
               policy_out: int):
        self.device = torch.device(device)
        self.ctx = _C.context(self.device)
        S, H, n_hidden = int(args.state_size), int(args.hidden_dim), int(args.n_hidden)
        self.state_size, self.policy_out = S, int(policy_out)
        dev = self.device
        # agent/por.py:34,46 / agent/sorl.py:34,47: with a backbone the heads read its (B, feature_dim) output.
        # Registered first, like the reference, so its keys lead the state_dict.
        self.backbone = None
        if backbone is not None:
            if not hasattr(backbone, "forward_features"):
                raise TypeError("backbone must be a porl_b200.agent.fasternet.FasterNet (kernel-backed encoder)")
            if S != 362:
                raise ValueError(f"the scan encoder reads 362-float states (util/costmap.py:11), got state_size={S}")
            self.backbone = backbone.to(dev)
        F_in = int(args.feature_dim) if backbone is not None else S
        self.in_features = F_in
        S_heads = F_in

        # --- value group: online TwinV parameters in the group's flat buffer; private target buffer
        probe = TwinV(S_heads, args.layer_norm, H, n_hidden, ctx=self.ctx, device=dev, init=False)  # geometry only
        self.v_group = ParamGroup(self.ctx, probe.plan.n_floats, dev, lr=value_lr, with_target=True, sharded=True)
        policy_cls = BoundedGaussianPolicy if self._bounded else GaussianPolicy
        # construction order follows the reference so that the global torch RNG is consumed alike:
        # POR builds goal_policy then vf (por.py:35-45); SORL builds v_net then policy (sorl.py:36-45)
        def make_value():
            return TwinV(S_heads, args.layer_norm, H, n_hidden, ctx=self.ctx, device=dev, flat=self.v_group.params,
                         grads=self.v_group.grads)

        def make_policy():
            pol_probe = policy_cls(S_heads, self.policy_out, H, n_hidden, ctx=self.ctx, device=dev, init=False)
            n = policy_cls.group_floats(pol_probe.plan, self.policy_out)
            self.p_group = ParamGroup(self.ctx, n, dev, lr=policy_lr, t_max=int(max_steps), sharded=True)
            return policy_cls(S_heads, self.policy_out, H, n_hidden, ctx=self.ctx, device=dev, flat=self.p_group.params,
                              grads=self.p_group.grads)

        if self._policy_first:
            policy, value = make_policy(), make_value()
        else:
            value, policy = make_value(), make_policy()
        target = TwinV(S_heads, args.layer_norm, H, n_hidden, ctx=self.ctx, device=dev, trainable=False, init=False,
                       flat=self.v_group.target)        # the group's own buffer: peer-mapped in the two-shot DP step
        target.flat.copy_(self.v_group.params[:target.flat.numel()])       # deepcopy(vf).requires_grad_(False)
        # registration order = the reference's attribute order, so state_dict() key order matches
        first, second = ((self._policy_name, policy), (self._value_name, value))
        if not self._policy_first:
            first, second = second, first
        setattr(self, first[0], first[1])
        setattr(self, second[0], second[1])
        setattr(self, self._target_name, target)
        # aliases that bypass nn.Module registration (state_dict must hold the reference's keys only)
        self.__dict__.update(_value=value, _policy=policy, _target=target)

        self.v_optimizer = self.v_group
        self._policy_optimizer = self.p_group
        self._schedule = CosineSchedule(self.p_group)
        self.tau, self.alpha, self.discount, self.beta = float(tau), float(alpha), float(discount), float(beta)
        self._ws = Workspace(dev)
        self._host = _C.pinned(4, torch.float32)
        self._host_status = _C.pinned(1, torch.int32)
        self._status = self.ctx.status_tensor()
        self._packed: dict[int, PackedBatch] = {}
        self._features: dict[int, tuple[torch.Tensor, torch.Tensor]] = {}
        self._graphs: dict[tuple, object] = {}

    # ------------------------------------------------------------------ batch packing
    def pack(self, observations, next_observations, rewards, terminals, actions=None) -> PackedBatch:
        """Copy caller tensors (any strides, e.g. column views of a collated row batch) into the
        padded device minibatch.  Inputs are not modified."""
        B = int(observations.shape[0])
        pb = self._packed.get(B)
        if pb is None:
            pb = PackedBatch(B, self.state_size, 2 if actions is None else int(actions.shape[-1]), self.device)
            self._packed[B] = pb
        dev = self.device
        pack_rows(self.ctx, observations.to(dev), self.state_size, pb.obs, pb.ld_obs)
        pack_rows(self.ctx, next_observations.to(dev), self.state_size, pb.next_obs, pb.ld_obs)
        pack_rows(self.ctx, rewards.to(dev).reshape(B), 1, pb.rewards, 1)
        pack_rows(self.ctx, terminals.to(dev).reshape(B), 1, pb.dones, 1)
        if actions is not None:
            pack_rows(self.ctx, actions.to(dev), pb.A, pb.actions, pb.ld_act)
        return pb

    # ------------------------------------------------------------------ the update
    def _head_inputs(self, pb: PackedBatch):
        """(x_obs, x_next, ld): what the value / policy heads read.  With a backbone: its features of obs, then of
        next_obs (that order: agent/sorl.py:81-83, agent/por.py:75-77), each call a train-mode forward that also
        clamps the packed states in place (util/costmap.py:17)."""
        if self.backbone is None:
            return pb.obs, pb.next_obs, pb.ld_obs
        ld = pad_to(self.in_features)
        f = self._features.get(pb.B)
        if f is None:
            f = tuple(torch.zeros(pb.B, ld, dtype=torch.float32, device=self.device) for _ in range(2))
            self._features[pb.B] = f
        self.backbone.forward_features(pb.obs, pb.ld_obs, pb.B, f[0])
        self.backbone.forward_features(pb.next_obs, pb.ld_obs, pb.B, f[1])
        return f[0], f[1], ld

    def _value_step(self, pb: PackedBatch, x_obs, x_next, ld_x) -> None:
        ctx, B = self.ctx, pb.B
        plan = self._value.plan
        ws = self._ws
        scratch_t = ws.get(("scratch_t", B), plan.saved_floats(B))
        saved_v = ws.get(("saved_v", B), plan.saved_floats(B))
        out_t = ws.get(("out_t", B), 2 * B * plan.out_ld)
        out_v = ws.get(("out_v", B), 2 * B * plan.out_ld)
        dv = ws.get(("dv", B), 2 * B * plan.out_ld)
        target_v = ws.get(("target_v", B), B)
        call_t = mlp_call(plan, self._target.flat, x_next, ld_x, scratch_t, out_t, inference=True)
        call_v = mlp_call(plan, self.v_group.params, x_obs, ld_x, saved_v, out_v)
        mlp_forward(ctx, [call_t, call_v], B)
        ld, net = plan.out_ld, B * plan.out_ld * 4
        b_norm = B * pdist.world_size()
        check(lib.porl_expectile_twin_loss(
            ctx.handle, out_v.data_ptr(), out_v.data_ptr() + net, out_t.data_ptr(), out_t.data_ptr() + net, ld,
            pb.rewards.data_ptr(), pb.dones.data_ptr(), self.discount, self.tau, B, b_norm, target_v.data_ptr(),
            dv.data_ptr(), dv.data_ptr() + net, self.v_group.loss_ptr, _C.stream_ptr()))
        mlp_backward(ctx, call_v, B, dv, self.v_group.grads)
        self.v_group.step(target=self._target.flat, ema_beta=self.beta)

    def _policy_step(self, pb: PackedBatch, x_obs, ld_obs, x: torch.Tensor, ld_x: int) -> None:
        ctx, B = self.ctx, pb.B
        vplan, pplan = self._value.plan, self._policy.plan
        ws = self._ws
        scratch_v = ws.get(("scratch_v2", B), vplan.saved_floats(B))
        out_v2 = ws.get(("out_v2", B), 2 * B * vplan.out_ld)
        saved_p = ws.get(("saved_p", B), pplan.saved_floats(B))
        out_p = ws.get(("out_p", B), B * pplan.out_ld)
        dz_p = ws.get(("dz_p", B), B * pplan.out_ld)
        target_v = ws.get(("target_v", B), B)
        call_v = mlp_call(vplan, self.v_group.params, x_obs, ld_obs, scratch_v, out_v2, inference=True)
        call_p = mlp_call(pplan, self.p_group.params, x_obs, ld_obs, saved_p, out_p)
        mlp_forward(ctx, [call_v, call_p], B)
        off = self._policy.log_std_offset * 4
        net = B * vplan.out_ld * 4
        b_norm = B * pdist.world_size()
        check(lib.porl_awr_gauss_nll(
            ctx.handle, out_p.data_ptr(), pplan.out_ld, self.p_group.params.data_ptr() + off, x.data_ptr(), ld_x,
            target_v.data_ptr(), out_v2.data_ptr(), out_v2.data_ptr() + net, vplan.out_ld, self.alpha,
            self._awr_mode, int(self._bounded), B, self.policy_out, b_norm, dz_p.data_ptr(),
            self.p_group.grads.data_ptr() + off, self.p_group.loss_ptr, _C.stream_ptr()))
        mlp_backward(ctx, call_p, B, dz_p, self.p_group.grads)
        self.p_group.step()

    def _policy_target(self, pb: PackedBatch):
        raise NotImplementedError

    def _whole_step(self) -> bool:
        """The update runs through the C ABI's whole-step entry point (porl_sorl_step / porl_por_step) unless a host-side
        NCCL all-reduce has to be issued between backward and optimizer (data parallel without peer-mapped buckets)."""
        if os.environ.get("PORL_WHOLE_STEP", "1") == "0":
            return False
        return not pdist.is_dp() or (self.v_group.peers is not None and self.p_group.peers is not None)

    def _step_desc(self) -> _C.AgentStep:
        d = self.__dict__.get("_desc")
        if d is None:
            d = _C.AgentStep()
            d.value_mlp, d.policy_mlp = self._value.plan.handle.value, self._policy.plan.handle.value
            d.value, d.policy = self.v_group.c_group(), self.p_group.c_group()
            d.target_params = self._target.flat.data_ptr()
            d.log_std_offset = self._policy.log_std_offset
            d.awr_mode, d.tanh_out, d.policy_out = self._awr_mode, int(self._bounded), self.policy_out
            self.__dict__["_desc"] = d
        d.discount, d.tau, d.alpha, d.ema_beta = self.discount, self.tau, self.alpha, self.beta   # plain attributes
        d.world = pdist.world_size()
        return d

    def _c_step(self, pb: PackedBatch, value_only: bool) -> None:
        x_obs, x_next, ld = self._head_inputs(pb)
        d = self._step_desc()
        n = int(lib.porl_step_workspace_floats(C.byref(d), pb.B))
        ws = self._ws.get(("step", pb.B), n)
        x, ld_x = self._policy_target(pb)
        if self._awr_mode == _C.AWR_SORL:
            check(lib.porl_sorl_step(self.ctx.handle, C.byref(d), pb.B, x_obs.data_ptr(), x_next.data_ptr(), ld,
                                     x.data_ptr(), ld_x, pb.rewards.data_ptr(), pb.dones.data_ptr(), ws.data_ptr(), n,
                                     int(value_only), _C.stream_ptr()))
        elif value_only:
            raise NotImplementedError("the reference's POR has no value-only update (SORL.vf_update, agent/sorl.py:130)")
        else:
            check(lib.porl_por_step(self.ctx.handle, C.byref(d), pb.B, x_obs.data_ptr(), x_next.data_ptr(), ld,
                                    x.data_ptr(), ld_x, pb.rewards.data_ptr(), pb.dones.data_ptr(), ws.data_ptr(), n,
                                    _C.stream_ptr()))

    def update_packed(self, pb: PackedBatch) -> None:
        """One gradient update on a packed device minibatch; asynchronous, returns nothing."""
        if self._whole_step():
            return self._c_step(pb, False)
        x_obs, x_next, ld = self._head_inputs(pb)
        self._value_step(pb, x_obs, x_next, ld)
        x, ld_x = self._policy_target(pb)
        self._policy_step(pb, x_obs, ld, x, ld_x)

    def value_update_packed(self, pb: PackedBatch) -> None:
        if self._whole_step() and self._awr_mode == _C.AWR_SORL:
            return self._c_step(pb, True)
        self._value_step(pb, *self._head_inputs(pb))

    def _graphable(self) -> bool:
        """The update may be replayed from a CUDA graph: no profiling, and no host-driven collective inside the step
        (single process, or the peer-memory gradient sum).  With a scan encoder in front of the heads the step is
        capturable in a single process: its DropPath factors come from torch's device generator, whose Philox offsets
        torch.cuda.graph advances per replay; under data parallelism the encoder's BatchNorm sums go through a
        torch.distributed callback, and a test that feeds recorded DropPath factors (`keep_queue`) consumes them on the
        host -- both stay eager."""
        if self.device.type != "cuda" or os.environ.get("PORL_UPDATE_GRAPH", "1") == "0":
            return False
        if self.backbone is not None and (pdist.is_dp() or self.backbone._cache.get("keep_queue") or
                                          os.environ.get("PORL_ENCODER_GRAPH", "1") == "0"):
            return False
        if getattr(self.ctx, "profiling", False) or torch.cuda.is_current_stream_capturing():
            return False
        return not pdist.is_dp() or (self.v_group.peers is not None and self.p_group.peers is not None)

    def _run_update(self, pb: PackedBatch, value_only: bool = False) -> None:
        """`update_packed(pb)` on the agent's own packed minibatch (stable addresses) as ONE CUDA graph launch from the
        third call at a batch size on: the reference-facing `update()` issues ~25 kernels, whose launch cost is what a
        small-batch update is made of.  First call eager (allocates the workspaces), second call captures.  Graphs are
        keyed by everything the launches bake in: batch size, the context's option epoch and the scalar
        hyper-parameters (learning rates and step counters live on the device)."""
        fn = self.value_update_packed if value_only else self.update_packed
        if not self._graphable():
            return fn(pb)
        key = (pb.B, value_only, getattr(self.ctx, "epoch", 0), self.tau, self.alpha, self.discount, self.beta)
        entry = self._graphs.get(key)
        if entry is None:
            if len(self._graphs) >= 16:
                self._graphs.clear()
            self._graphs[key] = False
            return fn(pb)
        if entry is False:
            entry = torch.cuda.CUDAGraph()
            with torch.cuda.graph(entry):
                fn(pb)
            self._graphs[key] = entry
        entry.replay()

    # ------------------------------------------------------------------ checkpoint extension
    def training_state_dict(self) -> dict:
        """The agent's share of a bitwise-identical resume, on top of the reference-layout `state_dict()` that
        por_train.py:97 / sorl_train.py:74 save: both Adam groups (moments, step, current LR of the cosine
        recursion) and the device RNG state the encoder's DropPath draws from.  The data order (sampler generator
        state or device-sampler epoch, update and episode counters) is added by the training loop under "loop"
        (porl_b200/train/offline_loop.py)."""
        return {"model": self.state_dict(), "v_optimizer": self.v_group.state_dict(),
                "policy_optimizer": self.p_group.state_dict(),
                "cuda_rng_state": torch.cuda.get_rng_state(self.device)}

    def load_training_state_dict(self, sd: dict) -> None:
        self.load_state_dict(sd["model"])
        self.v_group.load_state_dict(sd["v_optimizer"])
        self.p_group.load_state_dict(sd["policy_optimizer"])
        if sd.get("cuda_rng_state") is not None:
            torch.cuda.set_rng_state(sd["cuda_rng_state"].cpu(), self.device)

    def losses_async(self) -> "PendingLosses":
        """Enqueue the device->host copy of (v_loss, policy_loss, status word) of the update just issued into a pinned
        slot and return a handle; `handle.result()` waits for THAT copy only, so a caller can keep several updates in
        flight and read each step's losses one or two steps late (what `update_async` is for)."""
        ring = self.__dict__.get("_loss_ring")
        if ring is None:     # eight slots: a handle stays valid until eight later updates have been issued
            ring = [(_C.pinned(2, torch.float32), _C.pinned(1, torch.int32), torch.cuda.Event()) for _ in range(8)]
            self.__dict__["_loss_ring"], self.__dict__["_loss_next"] = ring, 0
        k = self.__dict__["_loss_next"]
        host, status, event = ring[k % len(ring)]
        self.__dict__["_loss_next"] = k + 1
        host[0:1].copy_(self.v_group.tail[0:1], non_blocking=True)
        host[1:2].copy_(self.p_group.tail[0:1], non_blocking=True)
        status.copy_(self._status, non_blocking=True)
        event.record()
        return PendingLosses(self, host, status, event)

    def _raise_on_flags(self, flags: int) -> None:
        if flags:
            self._status.zero_()
            msgs = []
            if flags & _C.FLAG_NONFINITE_LOSS:
                msgs.append("a loss is NaN/Inf")
            if flags & _C.FLAG_NONPOSITIVE_NLL:
                msgs.append("non-positive per-sample NLL (the reference drops into pdb here, agent/por.py:104)")
            if flags & _C.FLAG_BAD_INDEX:
                msgs.append("gather index out of range")
            if flags & _C.FLAG_SCAN_RANGE:
                msgs.append("scan value outside [0, 4.0001) u (8, inf): state2costmap's scatter_ raises here "
                            "(util/costmap.py:29)")
            if flags & _C.FLAG_PEER_TIMEOUT:
                msgs.append("a data-parallel peer did not signal its gradient bucket in time (porl_adam_step_peer)")
            raise FloatingPointError("; ".join(msgs) or f"device status {flags}")

    def read_losses(self) -> tuple[float, float]:
        """Synchronising read of (v_loss, policy_loss) of the last update; raises on device flags."""
        self._host[0:1].copy_(self.v_group.tail[0:1], non_blocking=True)
        self._host[1:2].copy_(self.p_group.tail[0:1], non_blocking=True)
        self._host_status.copy_(self._status, non_blocking=True)
        _C.sync_stream()
        self._raise_on_flags(int(self._host_status[0]))
        return float(self._host[0]), float(self._host[1])


class PendingLosses:
    """Losses of one update on their way to the host (see ExpectileAgent.losses_async)."""

    def __init__(self, agent, host, status, event):
        self.agent, self.host, self.status, self.event = agent, host, status, event

    def result(self) -> tuple[float, float]:
        self.event.synchronize()
        self.agent._raise_on_flags(int(self.status[0]))
        return float(self.host[0]), float(self.host[1])
