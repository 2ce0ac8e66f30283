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
if len(sys.argv) > 2 and sys.argv[2] == "agree":       # both contraction modes on the same batch (large-B index check)
    outs = []
    for mode in (0, 1):
        ctx.set_gemm_mode(mode)
        o = torch.empty(B, 256, device=dev)
        net.forward_features(st, st.stride(0), B, o, keep_scales=[None, None, None])
        torch.cuda.synchronize()
        outs.append(o)
    ctx.set_gemm_mode(0)
    d = float((outs[0].double() - outs[1].double()).norm() / outs[0].double().norm())
    print(f"B={B}: FFMA vs tcgen05 features relative L2 difference {d:.3e}", flush=True)
