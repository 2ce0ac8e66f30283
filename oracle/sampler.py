"""CPU restatement of the device index sampler (porl_b200/csrc/sampler.cu).  TEST INFRASTRUCTURE ONLY.

There is no reference code for this kernel (the reference shuffles on the host: torch.randperm in
DataLoader(shuffle=True), por_train.py:61; np.random.choice(size, B, replace=False), replaybuffer.py:61); what the
kernel must reproduce is their CONTRACT -- B distinct rows, every row once per epoch, a fresh shuffle per key --
which tests/test_sampler.py checks as properties, and this restatement (numpy uint64 / uint32 arithmetic, same
constants) pins the kernel bit for bit."""
from __future__ import annotations

import numpy as np

ROUNDS = 8
M64 = (1 << 64) - 1


def splitmix64(x: int) -> int:
    x = (x + 0x9E3779B97F4A7C15) & M64
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & M64
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & M64
    return x ^ (x >> 31)


def _round(x: np.ndarray, k: int) -> np.ndarray:
    x = x ^ np.uint32(k)
    x = x * np.uint32(0x9E3779B1)
    x = x ^ (x >> np.uint32(15))
    x = x * np.uint32(0x85EBCA77)
    x = x ^ (x >> np.uint32(13))
    x = x * np.uint32(0xC2B2AE3D)
    return x ^ (x >> np.uint32(16))


def permutation(n: int, seed: int, stream_id: int, first: int, count: int) -> np.ndarray:
    bits = 2
    while (1 << bits) < n:
        bits += 1
    lo_bits = bits // 2
    hi_bits = bits - lo_bits
    s = splitmix64(seed ^ splitmix64(stream_id))
    keys = []
    for _ in range(ROUNDS):
        s = splitmix64(s)
        keys.append(s >> 32)
    mlo, mhi = np.uint32((1 << lo_bits) - 1), np.uint32((1 << hi_bits) - 1)
    v = ((first + np.arange(count, dtype=np.int64)) % n).astype(np.uint64)
    todo = np.ones(count, dtype=bool)
    with np.errstate(over="ignore"):
        while todo.any():
            x = v[todo]
            lo = (x & np.uint64(mlo)).astype(np.uint32)
            hi = ((x >> np.uint64(lo_bits)) & np.uint64(mhi)).astype(np.uint32)
            for r in range(ROUNDS):
                if r & 1:
                    hi = hi ^ (_round(lo, keys[r]) & mhi)
                else:
                    lo = lo ^ (_round(hi, keys[r]) & mlo)
            x = (hi.astype(np.uint64) << np.uint64(lo_bits)) | lo.astype(np.uint64)
            v[todo] = x
            todo[todo] = x >= np.uint64(n)
    return v.astype(np.int64)
