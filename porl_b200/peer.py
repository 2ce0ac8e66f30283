"""Peer-mapped gradient buckets for the data-parallel step.

One process per GPU.  Each rank allocates its gradient bucket and a small signal pad with
`porl_peer_alloc` (plain cudaMalloc: CUDA IPC cannot export a slice of torch's caching allocator), sends the
two IPC handles to the other ranks of the node through torch.distributed (plumbing), and maps theirs.  The
Adam kernel then sums the gradients straight out of the peers' HBM over NVLink (porl_adam_step_peer in
include/porl_b200.h): no NCCL launch and no extra pass over the bucket between backward and optimizer.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from . import _C
from ._C import check, lib

# one-shot reads cost (world-1) * bytes per rank over NVLink: beyond this NCCL's ring / NVLS wins
MAX_PEER_READ_BYTES = int(os.environ.get("PORL_PEER_MAX_BYTES", str(64 << 20)))


def shard_range(n_floats: int, world: int, rank: int) -> tuple[int, int]:
    """[begin, end) in floats of the shard rank `rank` owns in the two-shot step (mirrors adam_sharded_kernel:
    chunks of a multiple of 32 float4)."""
    n4 = n_floats // 4
    chunk = ((n4 + world - 1) // world + 31) // 32 * 32
    s0 = min(rank * chunk, n4)
    return 4 * s0, 4 * min(s0 + chunk, n4)


class PeerBuffer:
    """cudaMalloc'ed, zero-filled device memory that other processes can map; `tensor(dtype)` is a torch view."""

    def __init__(self, ctx: _C.Context, nbytes: int):
        self.ctx, self.nbytes = ctx, int(nbytes)
        ptr = C.c_void_p()
        check(lib.porl_peer_alloc(ctx.handle, self.nbytes, C.byref(ptr)))
        self.ptr = ptr.value

    def handle(self) -> bytes:
        buf = C.create_string_buffer(64)
        check(lib.porl_peer_export(self.ctx.handle, self.ptr, buf))
        return buf.raw

    def tensor(self, dtype=torch.float32) -> torch.Tensor:
        item = torch.empty(0, dtype=dtype).element_size()
        typestr = {torch.float32: "<f4", torch.int32: "<i4", torch.uint8: "|u1"}[dtype]
        holder = _ArrayInterface(self, (self.nbytes // item,), typestr)
        return torch.as_tensor(holder, device=torch.device("cuda", self.ctx.device_index))

    def __del__(self):
        try:
            lib.porl_peer_free(self.ctx.handle, self.ptr)
        except Exception:
            pass


class _ArrayInterface:
    def __init__(self, owner: PeerBuffer, shape, typestr):
        self.owner = owner               # keeps the allocation alive as long as the tensor
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (owner.ptr, False), "version": 3}


class PeerReduce:
    """This rank's view of one optimizer group's peer set: fills porl_peer_group_t."""

    def __init__(self, ctx: _C.Context, rank: int, buckets: list[int], pads: list[int], tail_floats: int,
                 keep=(), wait_ms: int = 0, params: list[int] | None = None, targets: list[int] | None = None):
        world = len(buckets)
        if not (1 <= world <= _C.MAX_PEERS):
            raise ValueError(f"peer reduce supports 1..{_C.MAX_PEERS} ranks per node, got {world}")
        dev = torch.device("cuda", ctx.device_index)
        self.epoch = torch.zeros(4, dtype=torch.int32, device=dev)
        self.tail_out = torch.zeros(max(tail_floats, 1), dtype=torch.float32, device=dev)
        self.keep = keep
        g = _C.PeerGroup()
        g.world, g.rank = world, rank
        for r in range(world):
            g.bucket[r], g.pad[r] = buckets[r], pads[r]
        g.epoch, g.tail_out = self.epoch.data_ptr(), self.tail_out.data_ptr()
        g.tail_floats, g.wait_ms = tail_floats, wait_ms
        # two-shot (sharded) mode: every rank's parameter (and target-net) buffer as mapped by this rank
        self.sharded = params is not None
        if params is not None:
            for r in range(world):
                g.params[r] = params[r]
                g.target[r] = targets[r] if targets is not None else None
        self.struct = g
        self.world, self.rank = world, rank


def exchange(ctx: _C.Context, bucket: PeerBuffer, pad: PeerBuffer, tail_floats: int, group=None,
             params: PeerBuffer | None = None, target: PeerBuffer | None = None) -> PeerReduce | None:
    """Collective over the default (or given) process group: every rank sends the handles of its bucket and pad (and,
    for the two-shot step, of its parameter and target-net buffers) and maps the others'.  Returns None (caller keeps
    the NCCL all-reduce) when some pair of devices has no P2P path or the ranks are not on one node."""
    import torch.distributed as dist

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    bufs = [bucket, pad] + ([params] if params is not None else []) + ([target] if target is not None else [])
    mine = (tuple(b.handle() for b in bufs), os.uname().nodename, int(torch.cuda.current_device()))
    everyone: list = [None] * world
    dist.all_gather_object(everyone, mine, group=group)
    ok = world <= _C.MAX_PEERS and all(e[1] == mine[1] and len(e[0]) == len(bufs) for e in everyone)
    mapped, opened = [[0] * world for _ in bufs], []
    if ok:
        for r, (handles, _, _) in enumerate(everyone):
            if r == rank:
                for k, b in enumerate(bufs):
                    mapped[k][r] = b.ptr
                continue
            for k, h in enumerate(handles):
                ptr = C.c_void_p()
                st = lib.porl_peer_open(ctx.handle, h, C.byref(ptr))
                if st != 0:
                    ok = False
                    break
                mapped[k][r] = ptr.value
                opened.append(ptr.value)
            if not ok:
                break
    buckets, pads = mapped[0], mapped[1]
    verdict = [None] * world
    dist.all_gather_object(verdict, bool(ok), group=group)
    if not all(verdict):
        for ptr in opened:
            lib.porl_peer_close(ctx.handle, ptr)
        return None
    p_ptrs = mapped[2] if params is not None else None
    t_ptrs = mapped[3] if (params is not None and target is not None) else None
    return PeerReduce(ctx, rank, buckets, pads, tail_floats, keep=(*bufs, _Opened(ctx, opened)), params=p_ptrs,
                      targets=t_ptrs)


class _Opened:
    def __init__(self, ctx, ptrs):
        self.ctx, self.ptrs = ctx, ptrs

    def __del__(self):
        for p in self.ptrs:
            try:
                lib.porl_peer_close(self.ctx.handle, p)
            except Exception:
                pass
