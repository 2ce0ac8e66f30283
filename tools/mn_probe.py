"""Debug probe of the MN-major operand path of the tcgen05 kernel: structured operands, prints what comes back."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from porl_b200 import _C
dev = torch.device("cuda:0")
ctx = _C.context(dev)
ctx.set_gemm_mode(1)
M, N, K = 128, 128, 64

def run(a_k, b_k, A, B):
    C = torch.full((M, N), 7.0, device=dev)
    before = ctx.tc_launches
    _C.check(_C.lib.porl_gemm(ctx.handle, a_k, b_k, M, N, K, A.data_ptr(), A.stride(0), B.data_ptr(), B.stride(0),
                              C.data_ptr(), N, None, 0, None, 0, N, None, None, 0, _C.stream_ptr()))
    torch.cuda.synchronize()
    return C.cpu(), ctx.tc_launches - before

for k0 in (0, 1, 5, 8, 31, 33):
    # B MN-major: Bm[k][n]; only row k0 non-zero = n + 1
    Bm = torch.zeros(K, N)
    Bm[k0] = torch.arange(1, N + 1).float()
    A = torch.ones(M, K)
    C, tc = run(1, 0, A.to(dev), Bm.to(dev))
    ref = A @ Bm
    print(f"b_mn k0={k0} tc={tc} ok={torch.equal(C, ref)} nz={int((C != 0).sum())} C[0,:8]={C[0,:8].tolist()} C[0,30:36]={C[0,30:36].tolist()} C[5,:4]={C[5,:4].tolist()} max={float(C.max())}")
for k0 in (0, 1, 8, 33):
    Am = torch.zeros(K, M)                      # A MN-major: Am[k][m]
    Am[k0] = torch.arange(1, M + 1).float()
    Bk = torch.ones(N, K)
    C, tc = run(0, 1, Am.to(dev), Bk.to(dev))
    ref = Am.t() @ Bk.t()
    print(f"a_mn k0={k0} tc={tc} ok={torch.equal(C, ref)} nz={int((C != 0).sum())} C[:8,0]={C[:8,0].tolist()} C[30:36,0]={C[30:36,0].tolist()} max={float(C.max())}")
g = torch.Generator().manual_seed(1)
Am, Bm = torch.randn(K, M, generator=g), torch.randn(K, N, generator=g)
C, tc = run(0, 0, Am.to(dev), Bm.to(dev))
ref = Am.t().double() @ Bm.double()
print("both mn random: tc", tc, "max err", float((C.double() - ref).abs().max()), "max ref", float(ref.abs().max()))
