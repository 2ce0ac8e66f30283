"""Time the scan encoder forward (porl_fasternet_forward) per kernel category.  usage: python tools/encoder_probe.py [B]"""
import sys

import torch

from porl_b200 import _C
from porl_b200.agent.fasternet import FasterNet
from porl_b200.dataloader.dataset import synthetic_rows

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
QUICK = len(sys.argv) > 2 and sys.argv[2] in ("quick", "agree")      # one warm-up + one forward in mode 1 (for ncu)
dev = torch.device("cuda:0")
ctx = _C.context(dev)
torch.manual_seed(0)
net = FasterNet(3, 256).to(dev)
st = torch.from_numpy(synthetic_rows(B, 362, 1)[:, :362].copy()).to(dev)
out = torch.empty(B, 256, device=dev)
for mode in ((1,) if QUICK else (0, 1)):
    ctx.set_gemm_mode(mode)
    for _ in range(1 if QUICK else 3):
        net.forward_features(st, st.stride(0), B, out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 1 if QUICK else 10
    l0 = ctx.launches
    e0.record()
    for _ in range(n):
        net.forward_features(st, st.stride(0), B, out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    flops = 2 * 0.860e9 * B
    print(f"B={B} mode={mode}: {ms:.3f} ms/forward  {flops / ms * 1e-9:.1f} TFLOP/s (0.86 GMAC/sample)  "
          f"{(ctx.launches - l0) // n} launches/forward", flush=True)
    if QUICK:
        break
    ctx.profile(True)
    net.forward_features(st, st.stride(0), B, out)
    torch.cuda.synchronize()
    p = ctx.profile_read()
    ctx.profile(False)
    for k, v in p.items():
        if v["launches"]:
            unit = "TFLOP/s" if k == "contraction" else "GB/s"
            rate = v["work"] / v["ms"] * (1e-9 if k == "contraction" else 1e-6)
            print(f"    {k:12s} {v['launches']:3d} launches {v['ms']:8.3f} ms  {rate:8.1f} {unit}", flush=True)
ctx.set_gemm_mode(0)
if len(sys.argv) > 2 and sys.argv[2] == "agree":
    # large-B index check (offsets beyond 2^31 elements from B = 1942): a batch made of R copies of the same 256
    # states has the batch statistics of the 256-state batch, so every copy must reproduce its features
    ctx.set_gemm_mode(1)
    R = B // 256
    base = st[:256].contiguous()
    f256 = torch.empty(256, 256, device=dev)
    net.forward_features(base.clone(), 362, 256, f256, keep_scales=[None, None, None])
    big = base.repeat(R, 1).contiguous()
    fbig = torch.empty(R * 256, 256, device=dev)
    net.forward_features(big, 362, R * 256, fbig, keep_scales=[None, None, None])
    torch.cuda.synchronize()
    ctx.set_gemm_mode(0)
    d = max(float((fbig[i * 256:(i + 1) * 256].double() - f256.double()).norm() / f256.double().norm()) for i in range(R))
    print(f"B={R * 256} ({R} copies of 256 states) vs B=256: max relative L2 difference of a copy {d:.3e}; "
          f"peak memory {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB", flush=True)
