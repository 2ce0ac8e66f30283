"""Per-tensor comparison of the MLP forward/backward in gemm mode 1 (tcgen05) against mode 0 (FFMA) and
an fp64 autograd reference.   python tools/mlp_probe.py 365,256,256,1 2 64"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle.nets import build_mlp
from oracle.compare import rel_l2
from porl_b200 import _C, nets

dims = [int(x) for x in sys.argv[1].split(",")]
n_nets, B = int(sys.argv[2]), int(sys.argv[3])
dev = torch.device("cuda:0")
ctx = _C.context(dev)
torch.manual_seed(0)
plan = nets.FlatMLP(ctx, dims, n_nets=n_nets)
flat = torch.zeros(plan.n_floats, device=dev)
mods = [nets.MLPParams(plan, flat, i, None) for i in range(n_nets)]
x = torch.randn(B, dims[0])
ldx = nets.pad_to(dims[0])
xd = torch.zeros(B, ldx); xd[:, :dims[0]] = x; xd = xd.to(dev)
up = torch.randn(n_nets, B, dims[-1])
refs = []
for m in mods:
    ref = build_mlp(dims).double()
    ref.load_state_dict({k: v.detach().cpu().double() for k, v in m.state_dict().items()})
    refs.append(ref)
res = {}
for mode in (0, 1):
    ctx.set_gemm_mode(mode)
    grads = torch.full((plan.n_floats,), 3.0, device=dev)
    gm = [nets.MLPParams(plan, flat, i, grads, init=False) for i in range(n_nets)]
    saved = torch.zeros(plan.saved_floats(B), device=dev)
    out = torch.zeros(n_nets, B, plan.out_ld, device=dev)
    call = nets.mlp_call(plan, flat, xd, ldx, saved, out)
    t0 = ctx.tc_launches
    nets.mlp_forward(ctx, [call], B)
    dout = torch.zeros(n_nets, B, plan.out_ld); dout[:, :, :dims[-1]] = up
    nets.mlp_backward(ctx, call, B, dout.to(dev), grads)
    torch.cuda.synchronize()
    print(f"mode {mode}: tc launches {ctx.tc_launches - t0}")
    res[mode] = (out.cpu().clone(), [{k: p.grad.cpu().clone() for k, p in m.named_parameters()} for m in gm])
ctx.set_gemm_mode(0)
for i, ref in enumerate(refs):
    y = ref(x.double())
    (y * up[i].double()).sum().backward()
    print(f"net {i} forward rel_l2 mode0 {rel_l2(res[0][0][i, :, :dims[-1]], y):.2e} mode1 {rel_l2(res[1][0][i, :, :dims[-1]], y):.2e}")
    for k, p in ref.named_parameters():
        g0, g1 = res[0][1][i][k], res[1][1][i][k]
        d = (g1.double() - p.grad).abs()
        print(f"  {k:12s} rel_l2 mode0 {rel_l2(g0, p.grad):.2e} mode1 {rel_l2(g1, p.grad):.2e}  max|d| {float(d.max()):.2e} at {tuple(int(v) for v in torch.nonzero(d == d.max())[0])} shape {tuple(p.shape)}")
