"""FasterNet scan encoder, drop-in for agent/fasternet.py:264-451 as the agents use it
(`FasterNet(3, args.feature_dim)`, sorl_train.py:29-32): state (B, 362) -> state2costmap -> patch embedding ->
MLPBlock stages with partial 3x3 convolutions -> pooled head -> (B, num_classes) features.

The torch modules below only HOLD the parameters, with the reference's module tree, construction order and
initialisation (so state_dict() keys, their order, and the values drawn from a seeded global RNG are the
reference's); the arithmetic runs in libporl_b200.so (csrc/encoder.cu, porl_fasternet_forward).  Train-mode
semantics, as in the reference where nobody calls .eval(): batch-statistic BatchNorm with running-stat
updates, per-sample DropPath.  Forward only -- no optimizer owns these parameters in the reference
(agent/sorl.py:60,63; agent/por.py:61,64), so there is no backward pass to reproduce.
"""
from __future__ import annotations

import ctypes as C
from functools import partial

import torch
from torch import nn

STATE_SIZE = 362


def _trunc_normal_(m: nn.Module) -> None:
    """cls_init_weights (agent/fasternet.py:381-392): trunc-normal(std .02) conv / linear weights, zero biases."""
    if isinstance(m, (nn.Linear, nn.Conv1d, nn.Conv2d)):
        nn.init.trunc_normal_(m.weight, std=.02)
        if m.bias is not None:
            nn.init.constant_(m.bias, 0)
    elif isinstance(m, (nn.LayerNorm, nn.GroupNorm)):
        nn.init.constant_(m.bias, 0)
        nn.init.constant_(m.weight, 1.0)


class DropPath(nn.Module):
    """Holder of a block's drop probability (agent/fasternet.py:96-108)."""

    def __init__(self, drop_prob: float = 0., scale_by_keep: bool = True):
        super().__init__()
        self.drop_prob, self.scale_by_keep = drop_prob, scale_by_keep

    def extra_repr(self):
        return f'drop_prob={round(self.drop_prob, 3):0.3f}'


class Partial_conv3(nn.Module):
    def __init__(self, dim, n_div, forward='split_cat'):
        super().__init__()
        self.dim_conv3 = dim // n_div
        self.dim_untouched = dim - self.dim_conv3
        self.partial_conv3 = nn.Conv2d(self.dim_conv3, self.dim_conv3, 3, 1, 1, bias=False)


class MLPBlock(nn.Module):
    def __init__(self, dim, n_div, mlp_ratio, drop_path, act_layer, norm_layer, pconv_fw_type):
        super().__init__()
        self.dim, self.mlp_ratio, self.n_div = dim, mlp_ratio, n_div
        self.drop_path = DropPath(drop_path) if drop_path > 0. else nn.Identity()
        hidden = int(dim * mlp_ratio)
        self.mlp = nn.Sequential(nn.Conv2d(dim, hidden, 1, bias=False), norm_layer(hidden), act_layer(),
                                 nn.Conv2d(hidden, dim, 1, bias=False))
        self.spatial_mixing = Partial_conv3(dim, n_div, pconv_fw_type)


class BasicStage(nn.Module):
    def __init__(self, dim, depth, n_div, mlp_ratio, drop_path, norm_layer, act_layer, pconv_fw_type):
        super().__init__()
        self.blocks = nn.Sequential(*[MLPBlock(dim, n_div, mlp_ratio, drop_path[i], act_layer, norm_layer, pconv_fw_type)
                                      for i in range(depth)])


class PatchEmbed(nn.Module):
    def __init__(self, patch_size, patch_stride, in_chans, embed_dim, norm_layer):
        super().__init__()
        self.proj = nn.Conv2d(in_chans, embed_dim, kernel_size=patch_size, stride=patch_stride, bias=False)
        self.norm = norm_layer(embed_dim)


class PatchMerging(nn.Module):
    def __init__(self, patch_size2, patch_stride2, dim, norm_layer):
        super().__init__()
        self.reduction = nn.Conv2d(dim, 2 * dim, kernel_size=patch_size2, stride=patch_stride2, bias=False)
        self.norm = norm_layer(2 * dim)


class FasterNet(nn.Module):
    """Same constructor as agent/fasternet.py:266-290.  Only the configuration the agents instantiate is built
    as kernels: in_chans 3, embed_dim 96, depths (1, 2), mlp_ratio 2, n_div 4, patch 4/4 and 2/2, BN + ReLU,
    classification head; anything else raises NotImplementedError."""

    def __init__(self, in_chans=3, num_classes=1000, embed_dim=96, depths=(1, 2), mlp_ratio=2., n_div=4, patch_size=4,
                 patch_stride=4, patch_size2=2, patch_stride2=2, patch_norm=True, feature_dim=1280, drop_path_rate=0.1,
                 layer_scale_init_value=0, norm_layer='BN', act_layer='RELU', fork_feat=False, init_cfg=None,
                 pretrained=None, pconv_fw_type='split_cat', **kwargs):
        super().__init__()
        built = dict(in_chans=3, embed_dim=96, depths=(1, 2), mlp_ratio=2., n_div=4, patch_size=4, patch_stride=4,
                     patch_size2=2, patch_stride2=2, patch_norm=True, feature_dim=1280, layer_scale_init_value=0,
                     norm_layer='BN', act_layer='RELU', fork_feat=False, pconv_fw_type='split_cat')
        given = dict(in_chans=in_chans, embed_dim=embed_dim, depths=tuple(depths), mlp_ratio=mlp_ratio, n_div=n_div,
                     patch_size=patch_size, patch_stride=patch_stride, patch_size2=patch_size2,
                     patch_stride2=patch_stride2, patch_norm=patch_norm, feature_dim=feature_dim,
                     layer_scale_init_value=layer_scale_init_value, norm_layer=norm_layer, act_layer=act_layer,
                     fork_feat=fork_feat, pconv_fw_type=pconv_fw_type)
        diff = {k: v for k, v in given.items() if v != built[k]}
        if diff or num_classes <= 0:
            raise NotImplementedError(f"porl_b200 FasterNet kernels cover the agents' configuration only; got {diff}")
        norm, act = nn.BatchNorm2d, partial(nn.ReLU, inplace=True)
        self.num_classes, self.num_stages, self.embed_dim = num_classes, len(depths), embed_dim
        self.patch_norm, self.mlp_ratio, self.depths = patch_norm, mlp_ratio, depths
        self.num_features = int(embed_dim * 2 ** (self.num_stages - 1))
        self.patch_embed = PatchEmbed(patch_size, patch_stride, in_chans, embed_dim, norm)
        dpr = [x.item() for x in torch.linspace(0, drop_path_rate, sum(depths))]
        stages = []
        for i in range(self.num_stages):
            stages.append(BasicStage(int(embed_dim * 2 ** i), depths[i], n_div, mlp_ratio,
                                     dpr[sum(depths[:i]):sum(depths[:i + 1])], norm, act, pconv_fw_type))
            if i < self.num_stages - 1:
                stages.append(PatchMerging(patch_size2, patch_stride2, int(embed_dim * 2 ** i), norm))
        self.stages = nn.Sequential(*stages)
        self.fork_feat = fork_feat
        self.avgpool_pre_head = nn.Sequential(nn.AdaptiveAvgPool2d(1), nn.Conv2d(self.num_features, feature_dim, 1, bias=False),
                                              act())
        self.head = nn.Linear(feature_dim, num_classes)
        self.apply(_trunc_normal_)
        self.requires_grad_(False)            # frozen in effect in the reference (module docstring)
        self.__dict__["_cache"] = {}

    # ------------------------------------------------------------------ kernel-layout parameters
    def _blocks(self):
        return [self.stages[0].blocks[0], self.stages[2].blocks[0], self.stages[2].blocks[1]]

    def _bns(self):
        return [self.patch_embed.norm, self.stages[0].blocks[0].mlp[1], self.stages[1].norm,
                self.stages[2].blocks[0].mlp[1], self.stages[2].blocks[1].mlp[1]]

    def _params(self, keeps):
        """porl_fasternet_t over device tensors; weights are re-laid-out once per parameter version."""
        from .. import _C
        ws = [self.patch_embed.proj.weight, self.stages[1].reduction.weight, self.avgpool_pre_head[1].weight,
              self.head.weight, self.head.bias]
        for blk in self._blocks():
            ws += [blk.spatial_mixing.partial_conv3.weight, blk.mlp[0].weight, blk.mlp[3].weight]
        key = tuple((w.data_ptr(), w._version) for w in ws)
        c = self._cache
        if c.get("key") != key:
            f = lambda t: t.detach().to(torch.float32).contiguous()
            c["key"] = key
            c["pe_w"] = f(self.patch_embed.proj.weight.detach().reshape(96, 48).t())              # [(c,kh,kw)][o]
            c["merge_w"] = f(self.stages[1].reduction.weight.detach().permute(0, 2, 3, 1).reshape(192, 384))  # [o][(kh,kw,c)]
            c["head_conv_w"] = f(self.avgpool_pre_head[1].weight.detach().reshape(-1, 192))
            c["head_w"] = f(self.head.weight)
            hb = torch.zeros((self.num_classes + 3) // 4 * 4, dtype=torch.float32, device=self.head.bias.device)
            hb[:self.num_classes] = self.head.bias.detach()
            c["head_b"] = hb
            c["blocks"] = [(f(b.spatial_mixing.partial_conv3.weight.detach().permute(2, 3, 1, 0)),      # [kh][kw][ci][co]
                            f(b.mlp[0].weight.detach().flatten(1)), f(b.mlp[3].weight.detach().flatten(1)))
                           for b in self._blocks()]
        p = _C.FasterNetParams()
        ptr = lambda t: t.data_ptr()

        def bn(dst, m):
            for t in (m.weight, m.bias, m.running_mean, m.running_var):
                if t.dtype != torch.float32 or not t.is_contiguous():
                    raise TypeError("FasterNet BatchNorm tensors must be contiguous float32")
            dst.weight, dst.bias = ptr(m.weight), ptr(m.bias)
            dst.running_mean, dst.running_var = ptr(m.running_mean), ptr(m.running_var)

        bns = self._bns()
        p.pe_w = ptr(c["pe_w"])
        bn(p.pe_bn, bns[0])
        for i, (blk_bn, (pw, w1, w2)) in enumerate(zip((bns[1], bns[3], bns[4]), c["blocks"])):
            p.block[i].pconv_w, p.block[i].w1, p.block[i].w2 = ptr(pw), ptr(w1), ptr(w2)
            bn(p.block[i].bn, blk_bn)
            p.block[i].keep = ptr(keeps[i]) if keeps[i] is not None else None
        p.merge_w = ptr(c["merge_w"])
        bn(p.merge_bn, bns[2])
        p.head_conv_w, p.head_w, p.head_b = ptr(c["head_conv_w"]), ptr(c["head_w"]), ptr(c["head_b"])
        p.feature_dim = self.num_classes
        # data parallel: BatchNorm statistics cover the GLOBAL batch (sums all-reduced between the statistics
        # and the normalisation of every layer), so N ranks x B/N rows reproduce the single-process forward
        sync = self._cache.get("stats_sync")
        if sync is None:
            from .. import dist as pdist
            sync = (pdist.world_size(), pdist.allreduce_sum_) if pdist.is_dp() else (1, None)
        world, reduce_fn = sync
        p.world = world
        if world > 1:
            dev = self.head.weight.device
            buf = c.get("stats_buf")
            if buf is None or buf.device != dev:
                buf = c["stats_buf"] = torch.zeros(768, dtype=torch.float64, device=dev)

                def _cb(user, ptr_, n, stream, _buf=buf, _self=self):
                    try:                        # ptr_ is _buf's storage (p.stats_buf); torch needs the tensor
                        from .. import dist as pdist
                        fn = (_self._cache.get("stats_sync") or (0, None))[1] or pdist.allreduce_sum_
                        fn(_buf[:n])
                        return 0
                    except Exception as e:      # never unwind through the C frame
                        _self._cache["stats_error"] = e
                        return 1

                c["stats_cb"] = _C.ALLREDUCE_F64_FN(_cb)
            p.stats_buf = buf.data_ptr()
            p.stats_allreduce = C.cast(c["stats_cb"], C.c_void_p)
        return p

    def sync_batch_stats(self, world: int, reduce_fn) -> None:
        """Override how BatchNorm sums are combined across `world` ranks: reduce_fn(tensor_f64) must sum IN PLACE
        over ranks on the current stream.  Default (not called): torch.distributed's default group when it is
        initialised with more than one rank."""
        self._cache["stats_sync"] = (int(world), reduce_fn)
        self._cache.pop("stats_buf", None)

    def _keep_scales(self, x, B):
        """drop_path (agent/fasternet.py:76-93): one Bernoulli(keep)/keep factor per sample, drawn with the same
        torch calls (shape (B,1,1,1), device generator) the reference makes, in block order."""
        out = []
        for blk in self._blocks():
            dp = blk.drop_path
            if not isinstance(dp, DropPath) or dp.drop_prob == 0. or not self.training:
                out.append(None)
                continue
            keep = 1 - dp.drop_prob
            r = x.new_empty((B, 1, 1, 1)).bernoulli_(keep)
            if keep > 0.0 and dp.scale_by_keep:
                r.div_(keep)
            out.append(r.reshape(B))
        return out

    def replay_keep_scales(self, per_forward: list) -> None:
        """Parity hook: the next len(per_forward) forwards use these DropPath factors ([None, (B,), (B,)] each)
        instead of drawing them (the tests feed the oracle and this module the same masks)."""
        self._cache["keep_queue"] = list(per_forward)

    # ------------------------------------------------------------------ forward
    def forward_features(self, state: torch.Tensor, ld_state: int, B: int, out: torch.Tensor, keep_scales=None) -> None:
        """Kernel-level entry: `state` is a float32 CUDA buffer of B rows with stride ld_state (clamped in
        place), `out` a (B, >= num_classes) float32 buffer."""
        from .. import _C
        from .._C import check, lib
        if not self.training:
            raise NotImplementedError("eval-mode BatchNorm is not built: the reference never leaves train mode")
        ctx = _C.context(state.device)
        if keep_scales is None and self._cache.get("keep_queue"):
            keep_scales = self._cache["keep_queue"].pop(0)
        keeps = keep_scales if keep_scales is not None else self._keep_scales(state, B)
        keeps = [None if k is None else k.to(state.device, torch.float32).contiguous() for k in keeps]
        n = int(lib.porl_fasternet_workspace_floats(B))
        ws = self._cache.get("ws")
        if ws is None or ws.numel() < n or ws.device != state.device:
            ws = torch.empty(n, dtype=torch.float32, device=state.device)
            self._cache["ws"] = ws
        p = self._params(keeps)
        check(lib.porl_fasternet_forward(ctx.handle, C.byref(p), state.data_ptr(), ld_state, B, out.data_ptr(),
                                         out.stride(0), ws.data_ptr(), ws.numel(), _C.stream_ptr()))
        torch._foreach_add_([m.num_batches_tracked for m in self._bns()], 1)

    def forward_cls(self, x: torch.Tensor, keep_scales=None) -> torch.Tensor:
        """agent/fasternet.py:428-438.  x: (B, 362) float32 on the GPU; like the reference's state2costmap it is
        modified in place (values > 8 -> 0)."""
        if x.dim() != 2 or x.shape[1] != STATE_SIZE:
            raise ValueError(f"expected (B, {STATE_SIZE}) states, got {tuple(x.shape)}")
        B = int(x.shape[0])
        direct = x.dtype == torch.float32 and x.stride(1) == 1 and x.stride(0) >= STATE_SIZE
        buf = x if direct else x.to(torch.float32).contiguous()
        out = torch.empty(B, self.num_classes, dtype=torch.float32, device=x.device)
        if B:
            self.forward_features(buf, buf.stride(0), B, out, keep_scales)
            if not direct:
                x.copy_(buf)
        return out

    forward = forward_cls
