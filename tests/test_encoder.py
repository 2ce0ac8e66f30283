"""Scan encoder (state2costmap + FasterNet.forward_cls): oracle vs the reference's outputs (CPU), and the CUDA
path vs the oracle and the same fixture (GPU).  Fixture: tests/golden/encoder_small.npz, written by
tests/golden/make_golden.py from the unmodified reference (FasterNet(3, 256), deterministic weights,
6 states of which 5 carry the goal-bound cases of the reference's CostmapTestCase, injected DropPath factors).

Tolerances (relative L2 of the (B, 256) features).  The fixture is the reference's fp32 CPU forward on one
thread; it sits 1.3e-5 from an fp64 evaluation of the same network (worst sample 3.8e-5; 1.6e-6 with 8 threads:
the distance is dominated by torch's sequential fp32 BatchNorm statistics over 34,560 values per channel).  So:
vs the fixture 5e-5 (GOLD_TOL), vs the fp64 oracle 5e-6 (FEAT_TOL) -- the second is the meaningful bound and the
CUDA path is held to it.  The costmap and the in-place clamp are bit-exact.
"""
import numpy as np
import pytest
import torch

from oracle import encoder, refimport

FEAT_TOL = 5e-6
GOLD_TOL = 5e-5


def rel(a, b):
    a, b = torch.as_tensor(a).double(), torch.as_tensor(b).double()
    return float((a - b).norm() / b.norm())


def dense_from_nonzero(nz, B):
    cm = torch.zeros(B, 3, 360, 256)
    cm[nz[:, 0], nz[:, 1], nz[:, 2], nz[:, 3]] = 1.0
    return cm


# ------------------------------------------------------------------------------------------ CPU: oracle pinning
def test_oracle_costmap_matches_reference_fixture(golden):
    g = golden("encoder_small")
    seed, B = (int(x) for x in g["meta"])
    st, _ = encoder.encoder_inputs(seed, B)
    cm = encoder.state2costmap(st)
    assert torch.equal(cm, dense_from_nonzero(torch.from_numpy(g["costmap_nonzero"]).long(), B))
    assert np.array_equal(st.numpy(), g["clamped_state"])              # in-place clamp, costmap.py:17
    assert cm.stride() == (276480, 1, 768, 3)                            # permuted NHWC view, costmap.py:64


def test_oracle_fasternet_matches_reference_fixture(golden):
    g = golden("encoder_small")
    seed, B = (int(x) for x in g["meta"])
    st, keeps = encoder.encoder_inputs(seed, B)
    sd = encoder.det_encoder_state_dict(seed + 1)
    assert list(sd.keys()) != [] and sorted(sd.keys()) == sorted(g["keys"].tolist())
    sd64 = {k: (v.double() if v.is_floating_point() else v.clone()) for k, v in sd.items()}
    f32 = encoder.fasternet_forward(sd, st.clone(), keeps)
    f64 = encoder.fasternet_forward(sd64, st.clone(), keeps)
    assert rel(f32, g["features"]) < GOLD_TOL and rel(g["features"], f64) < GOLD_TOL and rel(f32, f64) < FEAT_TOL
    for k in g:
        if k.startswith("stats/"):
            name = k[len("stats/"):]
            if "tracked" in name:
                assert int(sd[name]) == int(g[k]) == 1
            else:
                scale = float(np.abs(g[k]).max())
                assert float((sd64[name] - torch.from_numpy(g[k]).double()).abs().max()) <= 2e-5 * scale + 1e-5, name   # the fixture's own fp32 noise


@pytest.mark.skipif(not refimport.available(), reason="reference not mounted (build container only)")
def test_oracle_matches_live_reference_random_goal_cells():
    """Bit-exact costmaps for 512 random states, against the imported reference."""
    rng = np.random.default_rng(5)
    st = rng.uniform(0.0, 4.0, (512, 362)).astype(np.float32)
    st[rng.random(st.shape) < 0.05] = 10.0
    st[:, -2:] = rng.uniform(-6.0, 6.0, (512, 2)).astype(np.float32)
    a, b = torch.from_numpy(st.copy()), torch.from_numpy(st.copy())
    assert torch.equal(refimport.costmap_fn()(a), encoder.state2costmap(b)) and torch.equal(a, b)


def test_fasternet_holder_matches_reference_layout_and_init(golden):
    """The parameter-holding module tree: reference key order, shapes, and (seeded) initial values."""
    from porl_b200.agent.fasternet import FasterNet
    g = golden("encoder_small")
    seed = int(g["meta"][0])
    torch.manual_seed(seed)
    sd = FasterNet(3, 256).state_dict()
    assert list(sd.keys()) == g["keys"].tolist()
    for k, v in sd.items():
        if v.is_floating_point():
            got = np.array([v.double().sum().item(), v.double().abs().sum().item()])
            np.testing.assert_allclose(got, g["init/" + k], rtol=1e-12, atol=1e-12, err_msg=k)
    with pytest.raises(NotImplementedError):
        FasterNet(3, 256, depths=(1, 2, 8, 2))


# ------------------------------------------------------------------------------------------ GPU: CUDA path
@pytest.mark.gpu
def test_cuda_costmap_bit_exact(golden):
    from porl_b200.util.costmap import state2costmap
    g = golden("encoder_small")
    seed, B = (int(x) for x in g["meta"])
    st, _ = encoder.encoder_inputs(seed, B)
    d = st.cuda()
    cm = state2costmap(d)
    assert cm.shape == (B, 3, 360, 256) and cm.stride() == (276480, 1, 768, 3)
    assert torch.equal(cm.cpu(), dense_from_nonzero(torch.from_numpy(g["costmap_nonzero"]).long(), B))
    assert np.array_equal(d.cpu().numpy(), g["clamped_state"])
    # larger random batch against the oracle, through a strided column view of a wider row batch
    rng = np.random.default_rng(6)
    rows = rng.uniform(0.0, 4.0, (300, 730)).astype(np.float32)
    rows[rng.random(rows.shape) < 0.05] = 10.0
    rows[:, 360:362] = rng.uniform(-6.0, 6.0, (300, 2)).astype(np.float32)
    dev_rows = torch.from_numpy(rows).cuda()
    view = dev_rows[:, :362]
    want_state = torch.from_numpy(rows[:, :362].copy())
    want = encoder.state2costmap(want_state)
    got = state2costmap(view)
    assert torch.equal(got.cpu(), want) and torch.equal(view.cpu(), want_state)
    assert torch.equal(dev_rows[:, 362:].cpu(), torch.from_numpy(rows[:, 362:]))   # nothing else touched


@pytest.mark.gpu
def test_cuda_costmap_flags_out_of_domain_scan():
    from porl_b200.util.costmap import state2costmap
    st = torch.full((2, 362), 1.0, device="cuda")
    st[1, 7] = 5.0                                   # in (4.0001, 8]: the reference's scatter_ raises
    with pytest.raises(RuntimeError):
        state2costmap(st)


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [0, 1])
def test_cuda_fasternet_matches_fixture_and_oracle(golden, mode):
    from porl_b200 import _C
    from porl_b200.agent.fasternet import FasterNet
    g = golden("encoder_small")
    seed, B = (int(x) for x in g["meta"])
    st, keeps = encoder.encoder_inputs(seed, B)
    sd = encoder.det_encoder_state_dict(seed + 1)
    net = FasterNet(3, 256)
    net.load_state_dict(encoder.order_like(sd, net.state_dict().keys()))
    net = net.cuda()
    ctx = _C.context(torch.device("cuda"))
    ctx.set_gemm_mode(mode)
    try:
        d = st.cuda()
        feats = net(d, keep_scales=keeps)
        torch.cuda.synchronize()
    finally:
        ctx.set_gemm_mode(0)
    sd64 = {k: (v.double() if v.is_floating_point() else v.clone()) for k, v in sd.items()}
    f64 = encoder.fasternet_forward(sd64, st.clone(), keeps)
    assert feats.shape == (B, 256)
    assert rel(feats.cpu(), g["features"]) < GOLD_TOL, rel(feats.cpu(), g["features"])
    assert rel(feats.cpu(), f64) < FEAT_TOL, rel(feats.cpu(), f64)
    assert np.array_equal(d.cpu().numpy(), g["clamped_state"])
    got = net.state_dict()
    for k in g:
        if k.startswith("stats/"):
            name = k[len("stats/"):]
            if "tracked" in name:
                assert int(got[name]) == 1
            else:
                scale = float(np.abs(g[k]).max())
                err = float((got[name].cpu().double() - sd64[name]).abs().max())
                assert err <= 2e-6 * scale + 1e-7, (name, err)


@pytest.mark.gpu
def test_cuda_fasternet_larger_batch_vs_oracle_and_determinism():
    """B = 48 (tcgen05 contractions on every layer incl. the head), random DropPath from the device generator
    replayed into the oracle; two runs from the same state are bitwise identical."""
    from porl_b200 import _C
    from porl_b200.agent.fasternet import FasterNet
    from oracle import synth
    B = 48
    st = torch.from_numpy(synth.laser_rows(B, 362, 77)[:, :362].copy())
    sd = encoder.det_encoder_state_dict(9)
    ctx = _C.context(torch.device("cuda"))
    ctx.set_gemm_mode(1)
    try:
        outs = []
        for _ in range(2):
            net = FasterNet(3, 256)
            net.load_state_dict(encoder.order_like({k: v.clone() for k, v in sd.items()}, net.state_dict().keys()))
            net = net.cuda()
            torch.manual_seed(3)
            keeps = net._keep_scales(st.cuda(), B)
            assert keeps[0] is None and keeps[1].shape == (B,)
            outs.append(net(st.cuda(), keep_scales=keeps))
        torch.cuda.synchronize()
    finally:
        ctx.set_gemm_mode(0)
    assert torch.equal(outs[0], outs[1])
    sd64 = {k: (v.double() if v.is_floating_point() else v.clone()) for k, v in sd.items()}
    f64 = encoder.fasternet_forward(sd64, st.clone(), [None if k is None else k.cpu() for k in keeps])
    assert rel(outs[0].cpu(), f64) < FEAT_TOL, rel(outs[0].cpu(), f64)


@pytest.mark.gpu
def test_cuda_fasternet_fused_batchnorm_matches_separate_passes(monkeypatch):
    """BatchNorm scale / shift / ReLU applied on the consuming contraction's operand (the default in tensor-core mode)
    against the same encoder with the apply as a pass of its own (PORL_ENC_FUSE_BN=0): the same arithmetic per element in
    the same order, so features and running statistics must be BITWISE equal; both agree with the fp64 oracle."""
    from porl_b200 import _C
    from porl_b200.agent.fasternet import FasterNet
    from oracle import synth
    B = 40                                           # 40 * 5760 rows: whole 128-row tiles plus a ragged last one per layer
    st = torch.from_numpy(synth.laser_rows(B, 362, 78)[:, :362].copy())
    sd = encoder.det_encoder_state_dict(10)
    ctx = _C.context(torch.device("cuda"))
    ctx.set_gemm_mode(1)
    try:
        outs, stats, launches = [], [], []
        for fuse in ("1", "0"):
            monkeypatch.setenv("PORL_ENC_FUSE_BN", fuse)     # read by every porl_fasternet_forward call
            net = FasterNet(3, 256)
            net.load_state_dict(encoder.order_like({k: v.clone() for k, v in sd.items()}, net.state_dict().keys()))
            net = net.cuda()
            torch.manual_seed(3)
            keeps = net._keep_scales(st.cuda(), B)
            l0 = ctx.launches
            outs.append(net(st.cuda(), keep_scales=keeps))
            torch.cuda.synchronize()
            launches.append(ctx.launches - l0)
            stats.append({k: v.clone() for k, v in net.state_dict().items() if "running" in k})
    finally:
        ctx.set_gemm_mode(0)
    assert launches[0] == launches[1] - 3, launches          # three bn_apply passes gone (one per block)
    assert torch.equal(outs[0], outs[1]), rel(outs[0].cpu(), outs[1].cpu())
    for k in stats[0]:
        assert torch.equal(stats[0][k], stats[1][k]), k
    sd64 = {k: (v.double() if v.is_floating_point() else v.clone()) for k, v in sd.items()}
    f64 = encoder.fasternet_forward(sd64, st.clone(), [None if k is None else k.cpu() for k in keeps])
    assert rel(outs[0].cpu(), f64) < FEAT_TOL, rel(outs[0].cpu(), f64)


# ------------------------------------------------------------------------------------------ GPU: agents with a backbone
class OracleBackbone:
    """The oracle's encoder as the callable the oracle agents expect (mutates its input, consumes one set of
    DropPath factors per call, updates the running statistics held in `sd`)."""

    def __init__(self, sd, keeps_per_call):
        self.sd, self.queue = sd, list(keeps_per_call)

    def __call__(self, x):
        return encoder.fasternet_forward(self.sd, x, self.queue.pop(0))


@pytest.mark.gpu
@pytest.mark.parametrize("kind,mode", [("sorl", 0), ("sorl", 1), ("por", 1)])
def test_agent_with_backbone_vs_oracle(kind, mode):
    """agent/sorl.py:81-83 / agent/por.py:75-77: both observation sets go through the train-mode encoder, the
    heads read its 256 features; POR regresses the raw next state AFTER the encoder's in-place clamp."""
    from porl_b200 import POR, SORL
    from porl_b200.agent.fasternet import FasterNet
    from oracle import synth
    from oracle.buffer import slice_row
    from oracle.compare import assert_params_close
    from oracle.steps import OraclePOR, OracleSORL, dims
    S, H, A, B, steps = 362, 64, 2, 40, 2
    args = dims(S, H, 2, False, 256, A)
    enc_sd = encoder.det_encoder_state_dict(21)
    g = torch.Generator().manual_seed(8)
    probs = encoder.drop_probs()
    keeps = [[None] + [torch.bernoulli(torch.full((B,), 1 - p), generator=g) / (1 - p) for p in probs[1:]]
             for _ in range(2 * steps)]
    ocls, cls = (OraclePOR, POR) if kind == "por" else (OracleSORL, SORL)
    hyper = dict(max_steps=1000, tau=0.8, alpha=10.0 if kind == "por" else 3.0)
    oracle = ocls(args, backbone=OracleBackbone({k: v.clone() for k, v in enc_sd.items()}, keeps), **hyper)
    head_sd = synth.det_state_dict(oracle, 5)
    oracle.load_state_dict(head_sd)
    net = FasterNet(3, 256)
    agent = cls(args, backbone=net, device=torch.device("cuda:0"), **hyper)
    full = {("backbone." + k): v for k, v in encoder.order_like(enc_sd, net.state_dict().keys()).items()}
    full.update(head_sd)
    assert list(agent.state_dict().keys())[:len(enc_sd)] == ["backbone." + k for k in net.state_dict().keys()]
    agent.load_state_dict(full)
    agent.backbone.replay_keep_scales(keeps)
    agent.ctx.set_gemm_mode(mode)
    try:
        rows = torch.from_numpy(synth.laser_rows(B * steps, S, 19, A))
        for i in range(steps):
            chunk = rows[i * B:(i + 1) * B]
            bo, bd = slice_row(chunk.clone(), S, A), slice_row(chunk.clone().cuda(), S, A)
            if kind == "por":
                want = oracle.por_residual_update(bo["observations"], bo["next_observations"], bo["rewards"], bo["terminals"])
                got = agent.por_residual_update(bd["observations"], bd["next_observations"], bd["rewards"], bd["terminals"])
            else:
                want = oracle.update(bo["observations"], bo["actions"], bo["rewards"], bo["next_observations"], bo["terminals"])
                got = agent.update(bd["observations"], bd["actions"], bd["rewards"], bd["next_observations"], bd["terminals"])
            np.testing.assert_allclose(got, want, rtol=2e-5, atol=0)
        sd, osd = agent.state_dict(), oracle.state_dict()
        for k, v in osd.items():
            assert_params_close(sd[k], v, lr=1e-4, n_steps=steps, name=k)
        for k, v in oracle.backbone.sd.items():
            if "running" in k:
                scale = float(v.abs().max())
                assert float((sd["backbone." + k].cpu() - v).abs().max()) <= 2e-6 * scale + 1e-6, k
            elif "tracked" in k:
                assert int(sd["backbone." + k]) == 2 * steps
    finally:
        agent.ctx.set_gemm_mode(0)


@pytest.mark.gpu
def test_cuda_fasternet_data_parallel_batch_stats():
    """Two 'ranks' (threads, one stream each, same GPU) run half of the batch each; the BatchNorm sums are
    combined through porl_fasternet_t.stats_allreduce, so every rank reproduces the single-process features of
    its rows and the single-process running statistics (what torch SyncBatchNorm would do for the reference)."""
    import threading
    from porl_b200 import _C
    from porl_b200.agent.fasternet import FasterNet
    from oracle import synth
    B, W = 16, 2
    st = torch.from_numpy(synth.laser_rows(B, 362, 91)[:, :362].copy())
    sd = encoder.det_encoder_state_dict(13)
    g = torch.Generator().manual_seed(4)
    keeps = [None] + [torch.bernoulli(torch.full((B,), 1 - p), generator=g) / (1 - p) for p in encoder.drop_probs()[1:]]

    def make():
        net = FasterNet(3, 256)
        net.load_state_dict(encoder.order_like({k: v.clone() for k, v in sd.items()}, net.state_dict().keys()))
        return net.cuda()

    ctx = _C.context(torch.device("cuda"))
    ctx.set_gemm_mode(0)                 # per-rank row counts differ from the full batch: keep the FFMA contraction
    whole = make()
    want = whole(st.cuda(), keep_scales=keeps)
    torch.cuda.synchronize()

    nets = [make() for _ in range(W)]
    slots, barrier, outs, errs = [None] * W, threading.Barrier(W), [None] * W, []
    # The two "ranks" share one library context here (its reduction tickets and scratch are per context; real ranks
    # are separate processes), so their GPU segments must not overlap: rank 1 runs each segment (the kernels
    # between two reductions) only after rank 0 has finished the same segment.
    cond, progress = threading.Condition(), {"rank0_calls": 0, "rank0_done": False}

    def wait_rank0(calls):
        with cond:
            cond.wait_for(lambda: progress["rank0_calls"] >= calls or progress["rank0_done"], timeout=60)

    def reducer(rank):
        state = {"k": 0}

        def fn(t):                                   # sum over "ranks" in rank order, in place
            state["k"] += 1
            torch.cuda.current_stream().synchronize()
            if rank == 0:
                with cond:
                    progress["rank0_calls"] = state["k"]
                    cond.notify_all()
            slots[rank] = t.clone()
            barrier.wait()
            total = slots[0] + slots[1]
            barrier.wait()
            t.copy_(total)
            torch.cuda.current_stream().synchronize()
            if rank == 1:
                wait_rank0(state["k"] + 1)
        return fn

    def work(rank):
        try:
            with torch.cuda.stream(torch.cuda.Stream()):
                nets[rank].sync_batch_stats(W, reducer(rank))
                rows = slice(rank * B // W, (rank + 1) * B // W)
                x = st[rows].cuda()
                ks = [None if k is None else k[rows].cuda() for k in keeps]
                torch.cuda.current_stream().synchronize()
                if rank == 1:
                    wait_rank0(1)
                outs[rank] = nets[rank](x, keep_scales=ks)
                torch.cuda.current_stream().synchronize()
                if rank == 0:
                    with cond:
                        progress["rank0_done"] = True
                        cond.notify_all()
        except Exception as e:                       # pragma: no cover
            errs.append(e)
            barrier.abort()

    threads = [threading.Thread(target=work, args=(r,)) for r in range(W)]
    [t.start() for t in threads]
    [t.join(120) for t in threads]
    assert not errs, errs
    got = torch.cat(outs)
    assert rel(got.cpu(), want.cpu()) < 2e-6, rel(got.cpu(), want.cpu())
    a, b = whole.state_dict(), nets[1].state_dict()
    for k in a:
        if "running" in k:
            assert float((a[k] - b[k]).abs().max()) <= 1e-6 * float(a[k].abs().max()) + 1e-7, k
