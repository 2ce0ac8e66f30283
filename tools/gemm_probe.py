"""Accuracy + timing probe of the contraction paths on a B200 (run under `timeout`).

  python tools/gemm_probe.py acc      error statistics vs fp64 for the shapes of the SORL step
  python tools/gemm_probe.py time     per-shape device time (CUDA events inside the library)
"""
import os, sys, math
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from porl_b200 import _C

dev = torch.device("cuda:0")
ctx = _C.context(dev)
pad = lambda x: (x + 3) // 4 * 4
SHAPES = [(1, 1, 4096, 512, 362), (1, 1, 4096, 512, 512), (1, 1, 512, 512, 4096), (1, 1, 512, 362, 4096),
          (1, 1, 4096, 1024, 1024), (1, 1, 1024, 1024, 4096), (1, 1, 256, 512, 512), (1, 1, 512, 512, 256),
          (1, 1, 16384, 512, 512)]
ENC_SHAPES = [(1, 1, 147456, 192, 96), (1, 1, 147456, 96, 192), (1, 1, 36864, 192, 384), (1, 1, 36864, 384, 192)]


def operands(a_k, b_k, M, N, K, seed=1, relu_like=True):
    g = torch.Generator().manual_seed(seed)
    A = torch.randn((M, pad(K)) if a_k else (K, pad(M)), generator=g)
    B = torch.randn((N, pad(K)) if b_k else (K, pad(N)), generator=g) * 0.05
    if relu_like:
        A = A.clamp_min(0)          # activations after ReLU: one-signed sums expose truncation bias
    return A, B


def run(a_k, b_k, M, N, K, Ad, Bd, Cd):
    _C.check(_C.lib.porl_gemm(ctx.handle, a_k, b_k, M, N, K, Ad.data_ptr(), Ad.stride(0), Bd.data_ptr(), Bd.stride(0),
                              Cd.data_ptr(), Cd.stride(0), None, 0, None, 0, N, None, None, 0, _C.stream_ptr()))


def acc():
    for (a_k, b_k, M, N, K) in SHAPES[:4] + [(1, 1, 100, 70, 133), (1, 1, 513, 129, 47)]:
        A, B = operands(a_k, b_k, M, N, K)
        Al = A[:, :K] if a_k else A[:, :M].t()
        Bl = B[:, :K].t() if b_k else B[:, :N]
        ref = Al.double() @ Bl.double()
        Ad, Bd = A.to(dev), B.to(dev)
        scale = float(ref.abs().mean())
        rows = []
        for mode, chunk, atm in ((0, 1, 0), (1, 1, 0), (1, 1, 1), (1, 1, 2), (1, 2, 1), (1, 4, 1), (1, 64, 0)):
            ctx.set_gemm_mode(mode)
            ctx.set_tc_chunk(chunk)
            ctx.set_tc_variant(atm)
            Cd = torch.zeros(M, pad(N), device=dev)
            run(a_k, b_k, M, N, K, Ad, Bd, Cd)
            torch.cuda.synchronize()
            e = Cd.cpu()[:, :N].double() - ref
            rows.append((f"mode{mode}_chunk{chunk}_atm{atm}", e))
        ctx.set_tc_chunk(1)
        ctx.set_tc_variant(0)
        ctx.set_gemm_mode(0)
        torch.backends.cuda.matmul.allow_tf32 = False
        e = (Al.to(dev) @ Bl.to(dev)).cpu().double() - ref
        rows.append(("cublas_fp32", e))
        e = (Al @ Bl).double() - ref
        rows.append(("cpu_fp32", e))
        for name, e in rows:
            print(f"acc form=({a_k},{b_k}) M={M} N={N} K={K} {name:12s} rms_rel={float(e.pow(2).mean().sqrt())/scale:.3e} "
                  f"max_rel={float(e.abs().max())/scale:.3e} bias_rel={float(e.mean())/scale:+.3e}", flush=True)


def time_shapes():
    for (a_k, b_k, M, N, K) in (ENC_SHAPES if len(sys.argv) > 2 and sys.argv[2] == "enc" else SHAPES):
        A, B = operands(a_k, b_k, M, N, K)
        Ad, Bd = A.to(dev), B.to(dev)
        Cd = torch.zeros(M, pad(N), device=dev)
        stamps = torch.zeros(16 * 148, dtype=torch.int64, device=dev)
        for mode, chunk, atm in ((0, 1, 0), (1, 1, 0), (1, 1, 1), (1, 1, 2)):
            ctx.set_gemm_mode(mode)
            ctx.set_tc_chunk(chunk)
            ctx.set_tc_variant(atm)
            ctx.set_tc_debug(stamps.data_ptr() if mode == 1 else None)
            for _ in range(3):
                run(a_k, b_k, M, N, K, Ad, Bd, Cd)
            torch.cuda.synchronize()
            ctx.profile(True)
            for _ in range(20):
                run(a_k, b_k, M, N, K, Ad, Bd, Cd)
            torch.cuda.synchronize()
            p = ctx.profile_read()
            ctx.profile(False)
            c = p["contraction"]
            us = c["ms"] / c["launches"] * 1e3
            extra = ""
            if mode == 1:
                t = stamps.cpu().view(148, 16)
                t = t[t[:, 15] > 0]
                if len(t):
                    d = t[:, 15] - t[:, 0]
                    extra = (f" ctas={len(t)} cta_us min/mean/max={float(d.min())/1e3:.1f}/"
                             f"{float(d.double().mean())/1e3:.1f}/{float(d.max())/1e3:.1f} "
                             f"span={float(t[:,15].max()-t[:,0].min())/1e3:.1f}\n     cta0 stamps (us from start): "
                             + " ".join(f"[{i}]{(int(t[0, i]) - int(t[0, 0]))/1e3:.1f}" for i in (1, 2, 3, 6, 7, 8, 9, 10, 15))
                             + "\n     cta0 wait share of CTA clocks: "
                             + " ".join(f"{n}={int(t[0, i]) / max(int(t[0, 5]), 1):.2f}" for n, i in
                                        (("producer/empty", 11), ("split/raw_full", 12), ("mma/split_done", 13),
                                         ("mma/tmem_empty", 14), ("acc/tmem_full", 4))) + f" clocks={int(t[0, 5])}")
                stamps.zero_()
            print(f"time form=({a_k},{b_k}) M={M} N={N} K={K} mode={mode} chunk={chunk} atm={atm} {us:8.1f} us/launch "
                  f"{2.0*M*N*K/us*1e-6:7.1f} TFLOP/s{extra}", flush=True)
        ctx.set_tc_debug(None)
        ctx.set_tc_variant(0)
        ctx.set_tc_chunk(1)
        ctx.set_gemm_mode(0)


def epi():
    """python tools/gemm_probe.py epi: cost of the epilogue features on the step's shapes (tensor-core mode)."""
    for (M, N, K) in ((4096, 512, 512), (16384, 512, 512), (16384, 512, 384), (8192, 512, 512)):
        A, B = operands(1, 1, M, N, K)
        Ad, Bd = A.to(dev), B.to(dev)
        Cd = torch.randn(M, pad(N), device=dev)
        Ct = torch.zeros(N, pad(M), device=dev)
        bias = torch.randn(pad(N), device=dev)
        stamps = torch.zeros(16 * 148, dtype=torch.int64, device=dev)
        ctx.set_gemm_mode(1)
        for name, kw in (("plain", {}), ("bias+relu", dict(bias=1, epi=1)), ("bias+relu+Ct", dict(bias=1, epi=1, ct=1)),
                         ("mask(in place)+Ct", dict(mask=1, ct=1)), ("mask(in place)", dict(mask=1)), ("Ct", dict(ct=1))):
            def go():
                _C.check(_C.lib.porl_gemm(ctx.handle, 1, 1, M, N, K, Ad.data_ptr(), Ad.stride(0), Bd.data_ptr(), Bd.stride(0),
                                          Cd.data_ptr(), Cd.stride(0), bias.data_ptr() if kw.get("bias") else None,
                                          kw.get("epi", 0), Cd.data_ptr() if kw.get("mask") else None, Cd.stride(0), N, None,
                                          Ct.data_ptr() if kw.get("ct") else None, Ct.stride(0), _C.stream_ptr()))
            ctx.set_tc_debug(stamps.data_ptr())
            for _ in range(3):
                go()
            torch.cuda.synchronize()
            ctx.profile(True)
            for _ in range(20):
                Cd.normal_()
                go()
            torch.cuda.synchronize()
            c = ctx.profile_read()["contraction"]
            ctx.profile(False)
            us = c["ms"] / c["launches"] * 1e3
            t = stamps.cpu().view(148, 16)
            t = t[t[:, 15] > 0]
            d = t[:, 15] - t[:, 0]
            print(f"epi M={M} N={N} K={K} {name:20s} {us:7.1f} us/launch {2.0*M*N*K/us*1e-6:6.1f} TFLOP/s  cta_us mean/max "
                  f"{float(d.double().mean())/1e3:.1f}/{float(d.max())/1e3:.1f}  cta0: first tile drained "
                  f"{(int(t[0,8])-int(t[0,0]))/1e3:.1f} epilogue done {(int(t[0,9])-int(t[0,0]))/1e3:.1f} all {(int(t[0,10])-int(t[0,0]))/1e3:.1f}"
                  f"  waits mma/tmem_empty={int(t[0,14])/max(int(t[0,5]),1):.2f} mma/split_done={int(t[0,13])/max(int(t[0,5]),1):.2f} "
                  f"acc/tmem_full={int(t[0,4])/max(int(t[0,5]),1):.2f} prod/empty={int(t[0,11])/max(int(t[0,5]),1):.2f}", flush=True)
            stamps.zero_()
        ctx.set_tc_debug(None)
        ctx.set_gemm_mode(0)


def forms():
    """python tools/gemm_probe.py forms: the four operand forms (K-major / MN-major A and B) on the weight-gradient and
    forward shapes, with the per-role wait shares."""
    for (M, N, K) in ((512, 512, 4096), (4096, 512, 512), (8192, 8192, 1024)):
        g = torch.Generator().manual_seed(1)
        Ak, Am = torch.randn(M, K, generator=g).to(dev), torch.randn(K, M, generator=g).to(dev)
        Bk, Bm = torch.randn(N, K, generator=g).to(dev), torch.randn(K, N, generator=g).to(dev)
        Cd = torch.zeros(M, N, device=dev)
        stamps = torch.zeros(16 * 148, dtype=torch.int64, device=dev)
        ctx.set_gemm_mode(1)
        for a_k, b_k in ((1, 1), (1, 0), (0, 1), (0, 0)):
            A, B = (Ak if a_k else Am), (Bk if b_k else Bm)
            def go():
                _C.check(_C.lib.porl_gemm(ctx.handle, a_k, b_k, M, N, K, A.data_ptr(), A.stride(0), B.data_ptr(), B.stride(0),
                                          Cd.data_ptr(), N, None, 0, None, 0, N, None, None, 0, _C.stream_ptr()))
            ctx.set_tc_debug(stamps.data_ptr())
            for _ in range(3):
                go()
            torch.cuda.synchronize()
            ctx.profile(True)
            for _ in range(20):
                go()
            torch.cuda.synchronize()
            c = ctx.profile_read()["contraction"]
            ctx.profile(False)
            us = c["ms"] / c["launches"] * 1e3
            t = stamps.cpu().view(148, 16)
            t = t[t[:, 15] > 0]
            d = t[:, 15] - t[:, 0]
            w = lambda i: int(t[0, i]) / max(int(t[0, 5]), 1)
            print(f"forms M={M} N={N} K={K} a_kmajor={a_k} b_kmajor={b_k} {us:7.1f} us {2.0*M*N*K/us*1e-6:6.1f} TFLOP/s cta mean/max "
                  f"{float(d.double().mean())/1e3:.1f}/{float(d.max())/1e3:.1f} drained {(int(t[0,8])-int(t[0,0]))/1e3:.1f} waits: prod/empty={w(11):.2f} "
                  f"split/raw_full={w(12):.2f} mma/split_done={w(13):.2f} mma/tmem_empty={w(14):.2f} acc/tmem_full={w(4):.2f}", flush=True)
            stamps.zero_()
        ctx.set_tc_debug(None)
        ctx.set_gemm_mode(0)


def one():
    """python tools/gemm_probe.py one M N K [atm]: three launches of one shape in tensor-core mode (for ncu)."""
    M, N, K = (int(x) for x in sys.argv[2:5])
    A, B = operands(1, 1, M, N, K)
    Ad, Bd = A.to(dev), B.to(dev)
    Cd = torch.zeros(M, pad(N), device=dev)
    ctx.set_gemm_mode(1)
    ctx.set_tc_variant(int(sys.argv[5]) if len(sys.argv) > 5 else 0)
    for _ in range(3):
        run(1, 1, M, N, K, Ad, Bd, Cd)
    torch.cuda.synchronize()
    ctx.set_gemm_mode(0)


if __name__ == "__main__" and len(sys.argv) > 1:
    {"acc": acc, "time": time_shapes, "one": one, "epi": epi, "forms": forms}[sys.argv[1]]()
