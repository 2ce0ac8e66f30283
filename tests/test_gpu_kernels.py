"""Kernel-level parity (run on the B200 with -m gpu): every C-ABI entry point against the CPU oracle
(oracle.formulas / eager torch fp64) on seeded inputs.  Integer/byte work is bit-exact; fp32 maths is
checked at the tolerance written next to each assert."""
import ctypes as C
import math

import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import formulas, synth  # noqa: E402


@pytest.fixture(scope="module")
def env():
    from porl_b200 import _C
    dev = torch.device("cuda:0")
    return _C, _C.context(dev), dev


def rel_l2(a, b):
    a, b = a.double().cpu().flatten(), b.double().cpu().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


# ------------------------------------------------------------------------------------------ K1
@pytest.mark.parametrize("S,A,B", [(362, 2, 256), (365, 2, 100), (22, 2, 1), (8, 3, 37)])
def test_gather_rows_and_transitions_bit_exact(env, S, A, B):
    from porl_b200.dataloader.dataset import TransitionDataset
    _C, ctx, dev = env
    rows = synth.laser_rows(500, S, 7, A)
    ds = TransitionDataset(rows, S, A, dev)
    g = torch.Generator().manual_seed(3)
    idx = torch.randint(0, 500, (B,), generator=g)
    out = ds.gather(idx)
    ref = torch.from_numpy(rows)[idx]
    assert torch.equal(out.cpu(), ref)
    assert torch.equal(ds[5].cpu(), torch.from_numpy(rows[5]))
    pb = ds.gather_packed(idx)
    assert torch.equal(pb.observations.cpu(), ref[:, :S])
    assert torch.equal(pb.next_observations.cpu(), ref[:, S + 1:2 * S + 1])
    assert torch.equal(pb.rewards.cpu(), ref[:, S])
    assert torch.equal(pb.dones.cpu(), ref[:, 2 * S + 1])
    assert torch.equal(pb.action_values.cpu(), ref[:, 2 * S + 2:])
    assert float(pb.obs[:, S:].abs().sum()) == 0.0 and float(pb.next_obs[:, S:].abs().sum()) == 0.0


def test_gather_empty_and_bad_index(env):
    from porl_b200.dataloader.dataset import TransitionDataset
    _C, ctx, dev = env
    rows = synth.laser_rows(10, 22, 1)
    ds = TransitionDataset(rows, 22, 2, dev)
    assert ds.gather(torch.zeros(0, dtype=torch.int64)).shape == (0, 48)
    out = ds.gather(torch.tensor([3, 99, -1]))          # out of range: clamped + flagged, never a fault
    torch.cuda.synchronize()
    status = ctx.status_tensor()
    assert int(status[0]) & _C.FLAG_BAD_INDEX
    status.zero_()
    assert torch.equal(out[0].cpu(), torch.from_numpy(rows[3]))


def test_replay_buffer_matches_golden(env, golden):
    """Same pushes + same np.random stream as the reference run behind tests/golden/replay_buffer.npz."""
    from porl_b200 import ReplayBuffer
    _C, ctx, dev = env
    g = golden("replay_buffer")
    seed = int(g["seed"])
    rb = ReplayBuffer(50, (7,), dev)
    rng = np.random.RandomState(seed)
    for i in range(70):
        rb.push(rng.rand(7).astype(np.float32), int(rng.randint(5)), float(rng.randn()),
                rng.rand(7).astype(np.float32), bool(rng.rand() < 0.2))
    assert len(rb) == int(g["size"]) and rb.position == int(g["position"])
    np.random.seed(seed)
    s, a, r, s2, d = rb.sample(16)
    assert a.dtype == torch.int64 and s.dtype == torch.float32 and s.device.type == "cuda"
    for got, key in ((s, "states"), (a, "actions"), (r, "rewards"), (s2, "next_states"), (d, "dones")):
        np.testing.assert_array_equal(got.cpu().numpy(), g[key])


def test_replay_buffer_push_batch_equals_push(env):
    from porl_b200 import ReplayBuffer
    _C, ctx, dev = env
    rng = np.random.RandomState(0)
    S, n = 9, 25
    st, nx = rng.rand(n, S).astype(np.float32), rng.rand(n, S).astype(np.float32)
    ac, rw, dn = rng.randint(0, 4, n), rng.randn(n).astype(np.float32), rng.rand(n) < 0.3
    a, b = ReplayBuffer(16, (S,), dev), ReplayBuffer(16, (S,), dev)
    for i in range(n):
        a.push(st[i], ac[i], rw[i], nx[i], dn[i])
    b.push_batch(st[:10], ac[:10], rw[:10], nx[:10], dn[:10])
    b.push_batch(st[10:], ac[10:], rw[10:], nx[10:], dn[10:])
    assert a.size == b.size and a.position == b.position
    for x, y in ((a.states, b.states), (a.next_states, b.next_states), (a.actions, b.actions),
                 (a.rewards, b.rewards), (a.dones, b.dones)):
        assert torch.equal(x, y)


# ------------------------------------------------------------------------------------------ contraction
def _gemm(env, a_k, b_k, M, N, K, epilogue=0, bias=False, relu=False, colsum=False, n_store=None, seed=0,
          transposed=False):
    _C, ctx, dev = env
    g = torch.Generator().manual_seed(seed)
    pad = lambda x: (x + 3) // 4 * 4
    A = torch.randn((M, pad(K)) if a_k else (K, pad(M)), generator=g)
    Bm = torch.randn((N, pad(K)) if b_k else (K, pad(N)), generator=g)
    Al = A[:, :K] if a_k else A[:, :M].t()
    Bl = Bm[:, :K].t() if b_k else Bm[:, :N]
    ref = Al.double() @ Bl.double()
    bias_t = torch.randn(N, generator=g) if bias else None
    if bias:
        ref = ref + bias_t.double()
    if epilogue == 1:
        ref = ref.clamp_min(0)
    elif epilogue == 2:
        ref = ref.tanh()
    n_store = n_store or N
    ldc = pad(n_store)
    relu_src = torch.randn(M, ldc, generator=g) if relu else None
    if relu:
        ref = ref * (relu_src[:, :N] > 0)
    Cd = torch.full((M, ldc), 7.0, device=dev)
    Ad, Bd = A.to(dev), Bm.to(dev)
    bd = bias_t.to(dev) if bias else None
    rd = relu_src.to(dev) if relu else None
    cs = torch.zeros(M, device=dev) if colsum else None
    p = lambda t: t.data_ptr() if t is not None else None
    ldct = pad(M)
    Ct = torch.full((N, ldct), 7.0, device=dev) if transposed else None
    _C.check(_C.lib.porl_gemm(ctx.handle, int(a_k), int(b_k), M, N, K, p(Ad), Ad.stride(0), p(Bd), Bd.stride(0),
                              p(Cd), ldc, p(bd), epilogue, p(rd), ldc, n_store, p(cs), p(Ct), ldct, _C.stream_ptr()))
    torch.cuda.synchronize()
    got = Cd.cpu()
    if transposed:
        assert torch.equal(Ct.cpu()[:, :M].t(), got[:, :N])                # bitwise the same values, transposed
        if ldct > M:
            assert bool((Ct.cpu()[:, M:] == 7.0).all())
    scale = float(ref.abs().max())
    err = float((got[:, :N].double() - ref).abs().max())
    assert err <= 2e-6 * max(scale, 1.0) * math.sqrt(K), (err, scale)      # fp32 accumulation over K terms
    if n_store > N:
        assert float(got[:, N:n_store].abs().max()) == 0.0
    if ldc > n_store:
        assert bool((got[:, n_store:] == 7.0).all())                       # never written past n_store
    if colsum:
        np.testing.assert_allclose(cs.cpu().numpy(), Al.double().sum(1).numpy(), rtol=0, atol=3e-5 * math.sqrt(K))
    return got


@pytest.mark.parametrize("M,N,K,kw", [(4096, 512, 362, {}), (512, 512, 4096, {}), (1000, 200, 96, dict(bias=True, epilogue=1)),
                                      (2048, 384, 512, dict(relu=True, transposed=True)), (300, 129, 640, dict(n_store=132))])
def test_gemm_tc_variants_bitwise_identical(env, M, N, K, kw):
    """The three tcgen05 kernel variants (operands in shared memory / A in tensor memory / CTA pairs with
    cta_group::2) issue the same MMAs on the same values in the same order: identical bits, incl. split-K.  The
    A-in-tensor-memory variant also serves MN-major operands (the weight- and input-gradient forms): checked against the
    shared-memory variant in all four operand forms."""
    ctx = env[1]
    ctx.set_gemm_mode(1)
    try:
        outs = []
        for variant in (0, 1, 2):
            ctx.set_tc_variant(variant)
            before = ctx.tc_launches
            outs.append(_gemm(env, 1, 1, M, N, K, seed=5, **kw))
            assert ctx.tc_launches == before + 1
        assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
        # split-K: the last-arriver fix-up (PORL_TC_LAST_ARRIVER: one slice sums all partials) against the distributed one
        os.environ["PORL_TC_LAST_ARRIVER"] = "1"
        try:
            for variant in (0, 1):
                ctx.set_tc_variant(variant)
                assert torch.equal(_gemm(env, 1, 1, M, N, K, seed=5, **kw), outs[0]), ("last arriver", variant)
        finally:
            del os.environ["PORL_TC_LAST_ARRIVER"]
        if not kw.get("transposed"):
            for a_k, b_k in ((1, 0), (0, 0), (0, 1)):
                forms = []
                for variant in (0, 1):
                    ctx.set_tc_variant(variant)
                    forms.append(_gemm(env, a_k, b_k, M, N, K, seed=6, **kw))
                assert torch.equal(forms[0], forms[1]), (a_k, b_k)
    finally:
        ctx.set_tc_variant(int(os.environ.get("PORL_TC_A_IN_TMEM", "1")))      # the library's default
        ctx.set_gemm_mode(0)


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("a_k,b_k", [(1, 1), (1, 0), (0, 0), (0, 1)])
@pytest.mark.parametrize("M,N,K", [(256, 512, 362), (100, 70, 33), (1, 5, 64), (513, 129, 17), (4096, 512, 512),
                                   (512, 362, 4096), (300, 129, 640)])
def test_gemm_forms(env, mode, a_k, b_k, M, N, K):
    env[1].set_gemm_mode(mode)
    try:
        before = env[1].tc_launches
        _gemm(env, a_k, b_k, M, N, K, seed=M + N + K)
        # the tcgen05 kernel must be the one that ran for tensor-core-sized problems (no silent FFMA fallback)
        # (every operand form: K-major and MN-major shared-memory descriptors)
        eligible = mode == 1 and min(M, N, K) >= 32 and M * N * K >= 2 ** 20 and N % 4 == 0
        assert (env[1].tc_launches - before) == (1 if eligible else 0)
    finally:
        env[1].set_gemm_mode(0)


@pytest.mark.parametrize("mode", [0, 1])
def test_gemm_epilogues_and_split_k(env, mode):
    env[1].set_gemm_mode(mode)
    try:
        _gemm(env, 1, 1, 300, 200, 365, epilogue=1, bias=True, n_store=224)      # forward hidden layer
        _gemm(env, 1, 1, 64, 2, 512, epilogue=2, bias=True, n_store=4)           # tanh head
        _gemm(env, 1, 0, 300, 200, 5, relu=True, n_store=224)                    # ReLU backward mask
        _gemm(env, 0, 0, 512, 362, 4096, colsum=True, n_store=384)               # weight grad, split-K + bias grad
        _gemm(env, 0, 0, 1, 512, 4096, colsum=True)                              # scalar-head weight grad
        _gemm(env, 0, 0, 64, 64, 100000, colsum=True)                            # deep split-K
        _gemm(env, 1, 1, 512, 362, 4096, n_store=384)                            # weight grad, K-major form, split-K
        _gemm(env, 1, 1, 64, 64, 100000)                                         # deep split-K, K-major
        _gemm(env, 1, 1, 200, 100, 365, relu=True, n_store=128)                  # ReLU backward mask, K-major
        if mode == 1:
            _gemm(env, 1, 1, 300, 200, 365, epilogue=1, bias=True, n_store=224, transposed=True)
            _gemm(env, 1, 1, 4096, 512, 512, relu=True, transposed=True)
            _gemm(env, 1, 1, 37, 33, 1000, n_store=36, transposed=True)
    finally:
        env[1].set_gemm_mode(0)


def test_gemm_split_k_is_deterministic(env):
    _C, ctx, dev = env
    g = torch.Generator().manual_seed(5)
    A = torch.randn(8192, 128, generator=g).to(dev)
    Bm = torch.randn(8192, 256, generator=g).to(dev)
    outs = []
    for _ in range(3):
        Cd = torch.zeros(128, 256, device=dev)
        _C.check(_C.lib.porl_gemm(ctx.handle, 0, 0, 128, 256, 8192, A.data_ptr(), 128, Bm.data_ptr(), 256,
                                  Cd.data_ptr(), 256, None, 0, None, 0, 256, None, None, 0, _C.stream_ptr()))
        outs.append(Cd.cpu())
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])


def test_gemm_rejects_bad_arguments(env):
    _C, ctx, dev = env
    A = torch.zeros(8, 6, device=dev)
    rc = _C.lib.porl_gemm(ctx.handle, 1, 1, 8, 8, 6, A.data_ptr(), 6, A.data_ptr(), 6, A.data_ptr(), 8, None, 0,
                          None, 0, 8, None, None, 0, _C.stream_ptr())
    assert rc == -1 and b"leading dimensions" in _C.lib.porl_last_error()


# ------------------------------------------------------------------------------------------ K4 / K5 / K6
@pytest.mark.parametrize("B,tau", [(256, 0.9), (37, 0.5), (4096, 0.7)])
def test_expectile_loss(env, B, tau):
    _C, ctx, dev = env
    g = torch.Generator().manual_seed(B)
    v1, v2, n1, n2, r = (torch.randn(B, generator=g) for _ in range(5))
    done = (torch.rand(B, generator=g) < 0.1).float()
    t64, loss64, d1, d2 = formulas.expectile_twin(v1.double(), v2.double(), n1.double(), n2.double(), r.double(),
                                                  done.double(), 0.99, tau)
    ld = 4
    def spread(x):
        o = torch.zeros(B, ld)
        o[:, 0] = x
        return o
    vv = torch.stack([spread(v1), spread(v2)]).to(dev)
    nn_ = torch.stack([spread(n1), spread(n2)]).to(dev)
    dv = torch.zeros(2, B, ld, device=dev)
    tv = torch.zeros(B, device=dev)
    loss = torch.zeros(8, device=dev)
    rd, dd = r.to(dev), done.to(dev)
    _C.check(_C.lib.porl_expectile_twin_loss(ctx.handle, vv[0].data_ptr(), vv[1].data_ptr(), nn_[0].data_ptr(),
                                             nn_[1].data_ptr(), ld, rd.data_ptr(), dd.data_ptr(), 0.99, tau, B, B,
                                             tv.data_ptr(), dv[0].data_ptr(), dv[1].data_ptr(), loss.data_ptr(),
                                             _C.stream_ptr()))
    torch.cuda.synchronize()
    assert abs(float(loss[0]) - float(loss64)) <= 1e-6 * abs(float(loss64))          # loss: 1e-6 relative
    np.testing.assert_allclose(tv.cpu().numpy(), t64.numpy(), rtol=1e-6, atol=1e-6)
    assert rel_l2(dv[0, :, 0], d1) <= 1e-6 and rel_l2(dv[1, :, 0], d2) <= 1e-6


@pytest.mark.parametrize("B,D,mode,tanh", [(256, 2, "sorl", True), (64, 365, "por", False), (1000, 22, "por", False),
                                           (4096, 2, "sorl", True)])
def test_awr_gauss_nll(env, B, D, mode, tanh):
    _C, ctx, dev = env
    g = torch.Generator().manual_seed(B + D)
    pre = torch.randn(B, D, generator=g) * 0.5
    mean = pre.tanh() if tanh else pre
    log_std = torch.randn(D, generator=g) * 0.3
    log_std[0] = 2.5                                    # outside the clamp: gradient must be zero
    x = torch.randn(B, D, generator=g)
    tv, v1, v2 = (torch.randn(B, generator=g) for _ in range(3))
    alpha = 3.0
    m64 = pre.double().requires_grad_(True)
    ls64 = log_std.double().requires_grad_(True)
    mu = m64.tanh() if tanh else m64
    loss64, _, _, _, nll = formulas.awr_gauss_nll(mu, ls64, x.double(), tv.double(), torch.minimum(v1, v2).double(),
                                                  alpha, mode)
    loss64.backward()
    ldm, ldx = (D + 3) // 4 * 4, (D + 31) // 32 * 32
    md = torch.zeros(B, ldm); md[:, :D] = mean
    xd = torch.zeros(B, ldx); xd[:, :D] = x
    md, xd = md.to(dev), xd.to(dev)
    vv = torch.zeros(2, B, 4); vv[0, :, 0] = v1; vv[1, :, 0] = v2
    vv = vv.to(dev)
    dz = torch.full((B, ldm), 9.0, device=dev)
    dls = torch.zeros(D, device=dev)
    loss = torch.zeros(8, device=dev)
    lsd, tvd = log_std.to(dev), tv.to(dev)
    _C.check(_C.lib.porl_awr_gauss_nll(ctx.handle, md.data_ptr(), ldm, lsd.data_ptr(), xd.data_ptr(), ldx,
                                       tvd.data_ptr(), vv[0].data_ptr(), vv[1].data_ptr(), 4, alpha,
                                       _C.AWR_POR if mode == "por" else _C.AWR_SORL, int(tanh), B, D, B,
                                       dz.data_ptr(), dls.data_ptr(), loss.data_ptr(), _C.stream_ptr()))
    torch.cuda.synchronize()
    assert abs(float(loss[0]) - float(loss64)) <= 2e-6 * abs(float(loss64))
    assert rel_l2(dz[:, :D], m64.grad) <= 2e-6
    assert rel_l2(dls, ls64.grad) <= 5e-6 and float(dls[0]) == 0.0
    if ldm > D:
        assert float(dz[:, D:].abs().max()) == 0.0


@pytest.mark.parametrize("B,A", [(64, 5), (4096, 4), (33, 2)])
def test_cql_head(env, B, A):
    _C, ctx, dev = env
    g = torch.Generator().manual_seed(B + A)
    q, qn = torch.randn(B, A, generator=g), torch.randn(B, A, generator=g)
    acts = torch.randint(0, A, (B,), generator=g)
    r = torch.randn(B, generator=g)
    d = (torch.rand(B, generator=g) < 0.2).float()
    loss64, td64, pen64, dq64, _ = formulas.cql_head(q.double(), qn.double(), acts, r.double(), d.double(), 0.99, 1.0)
    ld = (A + 3) // 4 * 4
    qd = torch.zeros(B, ld); qd[:, :A] = q
    nd = torch.zeros(B, ld); nd[:, :A] = qn
    qd, nd, ad, rd, dd = qd.to(dev), nd.to(dev), acts.to(dev), r.to(dev), d.to(dev)
    dq = torch.full((B, ld), 9.0, device=dev)
    out = torch.zeros(8, device=dev)
    _C.check(_C.lib.porl_cql_head(ctx.handle, qd.data_ptr(), nd.data_ptr(), ld, ad.data_ptr(), rd.data_ptr(),
                                  dd.data_ptr(), 0.99, 1.0, B, A, B, dq.data_ptr(), out.data_ptr(), _C.stream_ptr()))
    torch.cuda.synchronize()
    for got, ref in zip(out[:3].tolist(), (loss64, td64, pen64)):
        assert abs(got - float(ref)) <= 2e-6 * abs(float(ref))
    assert rel_l2(dq[:, :A], dq64) <= 2e-6
    if ld > A:
        assert float(dq[:, A:].abs().max()) == 0.0


# ------------------------------------------------------------------------------------------ K7
@pytest.mark.parametrize("n,t_max,ema,hard", [(1000, 0, 0.005, 0), (4099, 4, 0.0, 0), (257, 0, 0.0, 3)])
def test_adam_matches_torch(env, n, t_max, ema, hard):
    """12 steps against torch.optim.Adam (+CosineAnnealingLR, + Polyak / hard copy) on CPU."""
    from porl_b200.nets import ParamGroup
    _C, ctx, dev = env
    g = torch.Generator().manual_seed(n)
    p0 = torch.randn(n, generator=g)
    tgt0 = torch.randn(n, generator=g)
    grp = ParamGroup(ctx, n, dev, lr=1e-3, t_max=t_max)
    grp.params[:n].copy_(p0)
    tgt = torch.zeros(grp.n, device=dev); tgt[:n].copy_(tgt0)
    pref = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([pref], lr=1e-3)
    sch = torch.optim.lr_scheduler.CosineAnnealingLR(opt, t_max) if t_max else None
    tref = tgt0.clone()
    for step in range(1, 13):
        grad = torch.randn(n, generator=g) * (0.1 if step % 2 else 10.0)
        grp.grads[:n].copy_(grad)
        grp.step(target=tgt if (ema or hard) else None, ema_beta=ema, hard_every=hard)
        pref.grad = grad.clone()
        opt.step()
        if sch:
            sch.step()
        if ema:
            tref.mul_(1 - ema).add_(pref.detach(), alpha=ema)
        if hard and step % hard == 0:
            tref.copy_(pref.detach())
        # per step: parameters within 2 fp32 ulp-ish of torch (elementwise 1e-6 relative + 1e-9)
        np.testing.assert_allclose(grp.params[:n].cpu().numpy(), pref.detach().numpy(), rtol=2e-6, atol=2e-9)
    clock = grp.read_clock()
    assert clock.step == 12
    assert abs(clock.lr - opt.param_groups[0]["lr"]) <= 1e-15 + 1e-12 * opt.param_groups[0]["lr"]
    if ema or hard:
        np.testing.assert_allclose(tgt[:n].cpu().numpy(), tref.numpy(), rtol=2e-6, atol=2e-9)
    assert float(grp.params[n:].abs().sum()) == 0.0      # padding stays zero


@pytest.mark.parametrize("world,n", [(2, 4096), (3, 1028), (8, 2048), (1, 512)])
def test_adam_peer_sum_equals_presummed_adam(env, world, n):
    """porl_adam_step_peer: `world` replicas on `world` streams of ONE device (their buckets stand in for
    peer-mapped memory; grids small enough to be co-resident).  Each replica must end up bitwise equal to
    porl_adam_step on the gradients summed in rank order, the tails must add up, and six consecutive epochs
    exercise both handshakes."""
    from porl_b200.nets import ParamGroup, GRAD_TAIL
    from porl_b200.peer import PeerReduce
    _C, ctx, dev = env
    g = torch.Generator().manual_seed(world * 100 + n)
    p0 = torch.randn(n, generator=g).to(dev)
    ref = ParamGroup(ctx, n, dev, lr=1e-3, t_max=5)
    ref.params.copy_(p0)
    ref_t = torch.zeros(n, device=dev)
    reps = [ParamGroup(ctx, n, dev, lr=1e-3, t_max=5) for _ in range(world)]
    tgts = [torch.zeros(n, device=dev) for _ in range(world)]
    pads = [torch.zeros(_C.PEER_PAD_WORDS, dtype=torch.int32, device=dev) for _ in range(world)]
    for r, grp in enumerate(reps):
        grp.params.copy_(p0)
        grp.peers = PeerReduce(ctx, r, [q.bucket.data_ptr() for q in reps], [q.data_ptr() for q in pads],
                               GRAD_TAIL, wait_ms=5000)
    streams = [torch.cuda.Stream(dev) for _ in range(world)]
    for epoch in range(6):
        grads = [torch.randn(n + GRAD_TAIL, generator=g).to(dev) for _ in range(world)]
        total = grads[0].clone()
        for r in range(1, world):
            total += grads[r]
        for r, grp in enumerate(reps):
            grp.bucket.copy_(grads[r])
        ref.bucket.copy_(total)
        torch.cuda.synchronize()
        for r in reversed(range(world)):                      # launch order != rank order
            with torch.cuda.stream(streams[r]):
                reps[r].step(target=tgts[r], ema_beta=0.005)
        ref.step(target=ref_t, ema_beta=0.005)
        torch.cuda.synchronize()
        assert int(ctx.status_tensor()[0]) & _C.FLAG_PEER_TIMEOUT == 0
        for r, grp in enumerate(reps):
            assert torch.equal(grp.params, ref.params), (epoch, r)
            assert torch.equal(tgts[r], ref_t)
            assert torch.equal(grp.tail, total[n:])
            assert torch.equal(grp.bucket, grads[r])          # the bucket itself is only read
    assert all(grp.read_clock().step == 6 for grp in reps)


@pytest.mark.parametrize("world,n,ema,hard", [(2, 4096, 0.005, 0), (3, 1028, 0.0, 0), (8, 20480, 0.005, 0), (4, 512, 0.0, 2),
                                              (8, 132, 0.005, 0)])
def test_adam_two_shot_sharded_step_equals_presummed_adam(env, world, n, ema, hard):
    """Two-shot data-parallel step (porl_peer_group_t.params set): `world` ranks emulated on ONE device as a single
    cooperative launch (blockIdx.y = rank; porl_adam_step_peer_emulated), each with its own parameter / target /
    moment / bucket buffers.  Every rank must end up with parameters and target net BITWISE equal to porl_adam_step on
    the gradients summed in rank order, moments equal on the shard it owns and untouched elsewhere, tails summed, over
    six epochs (both handshakes, LR recursion, hard target copy every 2 steps)."""
    import ctypes as C
    from porl_b200.nets import ParamGroup, GRAD_TAIL
    from porl_b200.peer import PeerReduce, shard_range
    _C, ctx, dev = env
    g = torch.Generator().manual_seed(world * 1000 + n)
    p0 = torch.randn(n, generator=g).to(dev)
    use_t = ema > 0 or hard > 0
    ref = ParamGroup(ctx, n, dev, lr=1e-3, t_max=5, with_target=True)
    ref.params.copy_(p0)
    reps = [ParamGroup(ctx, n, dev, lr=1e-3, t_max=5, with_target=True) for _ in range(world)]
    pads = [torch.zeros(_C.PEER_PAD_WORDS, dtype=torch.int32, device=dev) for _ in range(world)]
    for r, grp in enumerate(reps):
        grp.params.copy_(p0)
        grp.peers = PeerReduce(ctx, r, [q.bucket.data_ptr() for q in reps], [q.data_ptr() for q in pads], GRAD_TAIL,
                               wait_ms=5000, params=[q.params.data_ptr() for q in reps],
                               targets=[q.target.data_ptr() for q in reps])
    ranks = (_C.PeerRank * world)()
    for r, grp in enumerate(reps):
        ranks[r].p, ranks[r].m, ranks[r].v = grp.params.data_ptr(), grp.m.data_ptr(), grp.v.data_ptr()
        ranks[r].target = grp.target.data_ptr() if use_t else None
        ranks[r].state = grp.state.data_ptr()
        ranks[r].peers = C.pointer(grp.peers.struct)
    ctx.status_tensor().zero_()
    for epoch in range(6):
        grads = [torch.randn(n + GRAD_TAIL, generator=g).to(dev) for _ in range(world)]
        total = grads[0].clone()
        for r in range(1, world):
            total += grads[r]
        for r, grp in enumerate(reps):
            grp.bucket.copy_(grads[r])
        ref.bucket.copy_(total)
        torch.cuda.synchronize()
        _C.check(_C.lib.porl_adam_step_peer_emulated(ctx.handle, world, ranks, n, 0.9, 0.999, 1e-8, ema, hard,
                                                     _C.stream_ptr()))
        ref.step(target=ref.target if use_t else None, ema_beta=ema, hard_every=hard)
        torch.cuda.synchronize()
        assert int(ctx.status_tensor()[0]) & _C.FLAG_PEER_TIMEOUT == 0
        for r, grp in enumerate(reps):
            assert torch.equal(grp.params, ref.params), (epoch, r)
            assert torch.equal(grp.target, ref.target), (epoch, r)
            assert torch.equal(grp.peers.tail_out, total[n:])
            assert torch.equal(grp.bucket, grads[r])          # the bucket itself is only read
            s0, s1 = shard_range(n, world, r)
            assert torch.equal(grp.m[s0:s1], ref.m[s0:s1]) and torch.equal(grp.v[s0:s1], ref.v[s0:s1])
            assert float(grp.m[:s0].abs().sum()) == 0.0 and float(grp.m[s1:].abs().sum()) == 0.0   # other shards untouched
    assert all(grp.read_clock().step == 6 for grp in reps)
    clocks = [grp.read_clock().lr for grp in reps]
    assert all(abs(c - ref.read_clock().lr) == 0.0 for c in clocks)
    # the shards tile the group exactly
    cover = sorted(shard_range(n, world, r) for r in range(world))
    assert cover[0][0] == 0 and cover[-1][1] == n and all(a[1] == b[0] for a, b in zip(cover, cover[1:]))


def test_adam_two_shot_timeout_skips_the_step(env):
    """Two-shot step, a peer that never signals: flag raised, nothing written (neither here nor into the peer)."""
    from porl_b200.nets import ParamGroup, GRAD_TAIL
    from porl_b200.peer import PeerReduce
    _C, ctx, dev = env
    n = 512
    grp, other = (ParamGroup(ctx, n, dev, lr=1e-3, with_target=True) for _ in range(2))
    pads = [torch.zeros(_C.PEER_PAD_WORDS, dtype=torch.int32, device=dev) for _ in range(2)]
    grp.peers = PeerReduce(ctx, 0, [grp.bucket.data_ptr(), other.bucket.data_ptr()], [q.data_ptr() for q in pads], GRAD_TAIL,
                           wait_ms=50, params=[grp.params.data_ptr(), other.params.data_ptr()],
                           targets=[grp.target.data_ptr(), other.target.data_ptr()])
    status = ctx.status_tensor()
    status.zero_()
    grp.params.normal_()
    grp.bucket.normal_()
    before = grp.params.clone()
    grp.step(target=grp.target, ema_beta=0.5)
    torch.cuda.synchronize()
    assert int(status[0]) & _C.FLAG_PEER_TIMEOUT
    assert torch.equal(grp.params, before) and float(other.params.abs().sum()) == 0.0
    assert float(grp.m.abs().sum()) == 0.0 and float(grp.target.abs().sum()) == 0.0 and grp.read_clock().step == 0
    status.zero_()


def test_adam_peer_timeout_sets_flag_and_returns(env):
    """A peer that never signals: the kernel gives up after wait_ms, raises PORL_FLAG_PEER_TIMEOUT and finishes
    WITHOUT applying the update (the other rank's bucket is not this step's gradient)."""
    from porl_b200.nets import ParamGroup, GRAD_TAIL
    from porl_b200.peer import PeerReduce
    _C, ctx, dev = env
    n = 256
    grp, other = ParamGroup(ctx, n, dev, lr=1e-3), ParamGroup(ctx, n, dev, lr=1e-3)
    pads = [torch.zeros(_C.PEER_PAD_WORDS, dtype=torch.int32, device=dev) for _ in range(2)]
    grp.peers = PeerReduce(ctx, 0, [grp.bucket.data_ptr(), other.bucket.data_ptr()], [q.data_ptr() for q in pads],
                           GRAD_TAIL, wait_ms=50)
    status = ctx.status_tensor()
    status.zero_()
    grp.params.normal_()
    grp.bucket.normal_()
    before = grp.params.clone()
    tgt = torch.ones(grp.n, device=dev)
    grp.step(target=tgt, ema_beta=0.5)
    torch.cuda.synchronize()
    assert int(status[0]) & _C.FLAG_PEER_TIMEOUT
    # the lost step leaves parameters, moments, target net and the optimizer clock untouched
    assert torch.equal(grp.params, before) and float(grp.m.abs().sum()) == 0.0 and float(grp.v.abs().sum()) == 0.0
    assert torch.equal(tgt, torch.ones_like(tgt)) and grp.read_clock().step == 0
    status.zero_()


def test_peer_buffer_alloc_view_export(env):
    from porl_b200.peer import PeerBuffer
    _C, ctx, dev = env
    buf = PeerBuffer(ctx, 4096)
    t = buf.tensor(torch.float32)
    assert t.shape == (1024,) and t.data_ptr() == buf.ptr and float(t.abs().sum()) == 0.0
    t.fill_(3.0)
    assert float(buf.tensor(torch.float32)[5]) == 3.0
    assert len(buf.handle()) == 64


# ------------------------------------------------------------------------------------------ K2 / K3
@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("dims,n_nets,ln,tanh,B", [
    ([362, 512, 512, 1], 2, False, False, 256),
    ([22, 32, 32, 1], 2, True, False, 16),
    ([365, 64, 64, 365], 1, False, False, 50),
    ([362, 64, 128, 64, 5], 1, False, False, 64),
    ([22, 32, 32, 2], 1, False, True, 33),
    ([40, 48, 2], 1, True, True, 7),
])
def test_mlp_forward_backward_matches_autograd(env, mode, dims, n_nets, ln, tanh, B):
    from oracle.nets import build_mlp
    from porl_b200 import nets
    _C, ctx, dev = env
    ctx.set_gemm_mode(mode)
    try:
        torch.manual_seed(sum(dims) + B)
        plan = nets.FlatMLP(ctx, dims, n_nets=n_nets, layer_norm=ln, tanh_out=tanh)
        flat = torch.zeros(plan.n_floats, device=dev)
        grads = torch.zeros(plan.n_floats, device=dev)
        mods = [nets.MLPParams(plan, flat, i, grads) for i in range(n_nets)]
        refs = []
        for m in mods:
            ref = build_mlp(dims, layer_norm=ln, tanh_out=tanh).double()
            sd = {k: v.detach().cpu().double() for k, v in m.state_dict().items()}
            if ln:   # non-trivial LayerNorm affine
                for k in sd:
                    if sd[k].ndim == 1 and k.split(".")[0] in ("1", "4"):
                        sd[k] = sd[k] + 0.1 * torch.randn_like(sd[k])
                m.load_state_dict({k: v.float() for k, v in sd.items()})
            ref.load_state_dict(sd)
            refs.append(ref)
        x = torch.randn(B, dims[0])
        ldx = nets.pad_to(dims[0])
        xd = torch.zeros(B, ldx); xd[:, :dims[0]] = x
        xd = xd.to(dev)
        saved = torch.zeros(plan.saved_floats(B), device=dev)
        out = torch.zeros(n_nets, B, plan.out_ld, device=dev)
        call = nets.mlp_call(plan, flat, xd, ldx, saved, out)
        nets.mlp_forward(ctx, [call], B)
        dout = torch.zeros(n_nets, B, plan.out_ld)
        up = torch.randn(n_nets, B, dims[-1])
        for i, ref in enumerate(refs):
            y = ref(x.double())
            assert rel_l2(out[i, :, :dims[-1]], y) <= 3e-6, "forward"
            if tanh:   # the loss kernels hand the gradient w.r.t. the PRE-activation to the backward
                dout[i, :, :dims[-1]] = up[i] * (1 - y.detach().float() ** 2)
            else:
                dout[i, :, :dims[-1]] = up[i]
            (y * up[i].double()).sum().backward()
        doutd = dout.to(dev)
        nets.mlp_backward(ctx, call, B, doutd, grads)
        torch.cuda.synchronize()
        for m, ref in zip(mods, refs):
            rp = dict(ref.named_parameters())
            for k, p in m.named_parameters():
                # per-tensor gradient rel-L2 <= 1e-5; a bias gradient is a cancelling fp32 sum over the batch, so
                # its error is bounded against the summed magnitudes instead (3e-7 * sum_b |dout|)
                d = float((p.grad.detach().cpu().double() - rp[k].grad).abs().max())
                assert rel_l2(p.grad, rp[k].grad) <= 1e-5 or d <= 3e-7 * float(up.abs().sum()), k
        # padding of the gradient bucket is written as exact zeros
        w0 = plan.layout(0, 0, _C.PARAM_WEIGHT)
        g0 = grads[w0[0]:w0[0] + dims[1] * w0[1]].view(dims[1], w0[1])
        assert float(g0[:, dims[0]:].abs().sum()) == 0.0
    finally:
        ctx.set_gemm_mode(0)
