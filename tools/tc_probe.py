"""Probe of the tcgen05 3xTF32 contraction on a B200: one form per process (run under `timeout`),
prints max abs error vs an fp64 matmul and whether the tensor-core kernel was the one launched."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from porl_b200 import _C

def main():
    a_k, b_k, M, N, K = (int(x) for x in sys.argv[1:6])
    mode = int(sys.argv[6]) if len(sys.argv) > 6 else 1
    dev = torch.device("cuda:0")
    ctx = _C.context(dev)
    ctx.set_gemm_mode(mode)
    g = torch.Generator().manual_seed(1)
    pad = lambda x: (x + 3) // 4 * 4
    A = torch.randn((M, pad(K)) if a_k else (K, pad(M)), generator=g)
    B = torch.randn((N, pad(K)) if b_k else (K, pad(N)), generator=g)
    Al = A[:, :K] if a_k else A[:, :M].t()
    Bl = B[:, :K].t() if b_k else B[:, :N]
    ref = Al.double() @ Bl.double()
    ldc = pad(N)
    Cd = torch.full((M, ldc), 7.0, device=dev)
    Ad, Bd = A.to(dev), B.to(dev)
    t0 = ctx.tc_launches
    _C.check(_C.lib.porl_gemm(ctx.handle, a_k, b_k, M, N, K, Ad.data_ptr(), Ad.stride(0), Bd.data_ptr(), Bd.stride(0),
                              Cd.data_ptr(), ldc, None, 0, None, 0, N, None, None, 0, _C.stream_ptr()))
    torch.cuda.synchronize()
    got = Cd.cpu()[:, :N].double()
    err = (got - ref).abs()
    print(f"form=({a_k},{b_k}) M={M} N={N} K={K} mode={mode} tc_launches={ctx.tc_launches - t0} "
          f"max_err={float(err.max()):.3e} ref_max={float(ref.abs().max()):.3e} "
          f"bad_rows={int((err.max(1).values > 1e-3).sum())} bad_cols={int((err.max(0).values > 1e-3).sum())}")
    if float(err.max()) > 1e-3:
        # locate the error pattern
        r = err.max(1).values
        c = err.max(0).values
        print(" first bad rows:", torch.nonzero(r > 1e-3).flatten()[:16].tolist())
        print(" first bad cols:", torch.nonzero(c > 1e-3).flatten()[:16].tolist())
        print(" got[0,:8]", got[0, :8].tolist())
        print(" ref[0,:8]", ref[0, :8].tolist())

main()
