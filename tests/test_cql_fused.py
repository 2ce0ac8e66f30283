"""The single-kernel CQL update (porl_cql_fused_grad, csrc/cql_fused.cu) against the CPU oracle and against the
layer-by-layer CUDA path (B200, -m gpu).

Tolerances (fp32): losses 1e-5 relative; per-tensor gradient rel-L2 <= 1e-5 (against the fp64 oracle 2e-6 is
typical); the gather is bit-exact, so pre-gathered rows and in-kernel gathered rows give bitwise equal updates.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import synth  # noqa: E402
from oracle.buffer import slice_row  # noqa: E402
from oracle.steps import OracleCQL  # noqa: E402

DEV = "cuda:0"


@pytest.fixture(params=[(0, 8), (1, 16), (1, 8)], ids=["ffma", "mma3xtf32_16warps", "mma3xtf32"])
def mode(request, monkeypatch):
    """Contraction mode of the context: 0 = fp32 FFMA tiles, 1 = mma.sync 3xTF32 tiles inside the fused kernel (as a
    256-thread CTA with 16 x 16 warp tiles, the default, or a 512-thread CTA with 16 x 8 warp tiles)."""
    from porl_b200 import _C
    ctx = _C.context(torch.device(DEV))
    ctx.set_gemm_mode(request.param[0])
    monkeypatch.setenv("PORL_CQL_WARPS", str(request.param[1]))
    yield request.param[0]
    ctx.set_gemm_mode(0)


def make(S, A, n_rows, seed, tmp="/tmp/porl_b200_logs"):
    from porl_b200 import CQLTrainer, QNetwork
    from porl_b200.buffer.replaybuffer import ReplayBuffer
    t = CQLTrainer(state_size=S, action_size=A, gamma=0.99, epsilon=1.0, epsilon_min=0.05, epsilon_decay=0.99,
                   update_target_freq=10, device=torch.device(DEV), network=QNetwork, log_dir=tmp)
    o = OracleCQL(S, A)
    sd = synth.det_state_dict(o.q_network, seed)
    o.q_network.load_state_dict(sd)
    t.q_network.load_state_dict(sd)
    sd_t = synth.det_state_dict(o.q_network, seed + 50)          # a target net that differs from the online one
    o.target_network.load_state_dict(sd_t)
    t.target_network.load_state_dict(sd_t)
    rows = synth.laser_rows(n_rows, S, seed + 1)
    acts = synth.discrete_actions(n_rows, A, seed + 2)
    b = slice_row(rows, S)
    t.replay_buffer = ReplayBuffer(n_rows, (S,), torch.device(DEV))
    t.replay_buffer.push_batch(b["observations"], acts, b["rewards"], b["next_observations"], b["terminals"])
    return t, o, b, acts


def oracle_grads(o, b, acts, idx, dtype=torch.float32):
    """Loss and gradients of cql_trainer.py:88-112 on rows idx (autograd on the oracle's network)."""
    import copy
    q, tgt = copy.deepcopy(o.q_network).to(dtype), copy.deepcopy(o.target_network).to(dtype)
    st = torch.from_numpy(b["observations"][idx]).to(dtype)
    nx = torch.from_numpy(b["next_observations"][idx]).to(dtype)
    r = torch.from_numpy(b["rewards"][idx]).to(dtype)
    d = torch.from_numpy(b["terminals"][idx]).to(dtype)
    a = torch.from_numpy(acts[idx]).long()
    with torch.no_grad():
        targets = r + 0.99 * tgt(nx).max(dim=1).values * (1 - d)
    qv = q(st)
    qa = qv.gather(1, a[:, None]).squeeze(1)
    td = ((qa - targets) ** 2).mean()
    pen = (torch.logsumexp(qv, dim=1) - np.log(qv.shape[1]) - qa).mean()
    loss = td + 1.0 * pen
    loss.backward()
    return float(loss.detach()), float(td.detach()), float(pen.detach()), {k: p.grad for k, p in q.named_parameters()}


def relu_margin(net, obs):
    """Per row: the smallest |pre-activation| over every ReLU of `net` (fp64).  A unit whose pre-activation lies inside
    fp32 rounding noise of zero may take either side of the ReLU in two correct fp32 implementations and then moves
    the gradient by O(1/B) -- tests/test_gpu_agents.py::relu_margin explains the same filter for the SORL/POR tests."""
    import copy
    with torch.no_grad():
        x = torch.from_numpy(obs).double()
        m = torch.full((x.shape[0],), float("inf"), dtype=torch.float64)
        for layer in copy.deepcopy(net.model).double():
            if isinstance(layer, torch.nn.ReLU):
                m = torch.minimum(m, x.abs().min(dim=1).values)
            x = layer(x)
    return m.numpy()


def grads_of(t):
    """Gradient bucket of the trainer as {state_dict key: tensor}."""
    out = {}
    plan = t.q_network.plan
    names = list(t.q_network.state_dict().keys())
    for li in range(plan.n_layers):
        out[names[2 * li]] = plan.view(t.q_group.grads, 0, li, 0).clone()
        out[names[2 * li + 1]] = plan.view(t.q_group.grads, 0, li, 1).clone()
    return out


def rel_l2(a, b):
    a, b = a.detach().double().cpu().flatten(), b.detach().double().cpu().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


@pytest.mark.parametrize("S,A,B,n_rows", [(362, 5, 4096, 6000), (362, 5, 77, 500), (8, 4, 64, 300),
                                           (22, 5, 1, 40), (100, 8, 148 * 32 * 2 + 5, 9600)])
def test_fused_gradients_match_oracle(S, A, B, n_rows, mode):
    t, o, b, acts = make(S, A, n_rows, 11)
    assert t.fused
    rng = np.random.default_rng(5)
    ok = np.flatnonzero(relu_margin(o.q_network, b["observations"]) > 2e-5)
    assert len(ok) >= min(B, n_rows // 2)
    idx = rng.choice(ok, B, replace=B > len(ok))
    loss, td, pen, want = oracle_grads(o, b, acts, idx)
    loss64, td64, pen64, want64 = oracle_grads(o, b, acts, idx, torch.float64)
    p0 = t.q_group.params.clone()
    t.learn_indices(torch.as_tensor(idx))
    got_l = t.read_losses()
    # the penalty is a small difference of O(1) terms (logsumexp - log A - q[a]): absolute floor of a few fp32 ulps of those
    np.testing.assert_allclose(got_l, (loss64, td64, pen64), rtol=1e-5, atol=3e-7)
    got = grads_of(t)
    for k in want:
        ref_err = rel_l2(want[k], want64[k])
        err = rel_l2(got[k], want64[k])
        assert err <= max(1e-5, 3 * ref_err), (k, err, ref_err)
    # pad columns / pad entries of the bucket are zero (the Adam kernel runs over the padded extent)
    plan = t.q_network.plan
    mask = torch.ones_like(t.q_group.grads, dtype=torch.bool)
    for li in range(plan.n_layers):
        for kind in (0, 1):
            off, ld = plan.layout(0, li, kind)
            rows_, cols = (plan.dims[li + 1], plan.dims[li]) if kind == 0 else (1, plan.dims[li + 1])
            for r in range(rows_):
                mask[off + r * ld: off + r * ld + cols] = False
    assert torch.count_nonzero(t.q_group.grads[mask]) == 0
    # the Adam step that followed used these gradients: first step moves every touched weight by ~lr
    assert float((t.q_group.params - p0).abs().max()) <= 5e-4 * 1.001


def test_fused_equals_layerwise_path_and_is_deterministic(monkeypatch, mode):
    S, A, B, n_rows = 362, 5, 1000, 3000
    t, o, b, acts = make(S, A, n_rows, 3)
    u, _, _, _ = make(S, A, n_rows, 3)
    idx = torch.as_tensor(np.random.default_rng(1).choice(n_rows, B, replace=False))
    t.learn_indices(idx)
    l_f = t.read_losses()
    g_f = t.q_group.grads.clone()
    # same rows, already gathered (idx == NULL branch): bitwise the same gradient
    pb = u.replay_buffer.gather_packed(idx)
    u.learn_packed(pb)
    assert u.read_losses() == l_f
    assert torch.equal(u.q_group.grads, g_f)
    assert torch.equal(u.q_group.params, t.q_group.params)
    # run-to-run bitwise determinism (partials are summed in CTA order)
    v, _, _, _ = make(S, A, n_rows, 3)
    v.learn_indices(idx)
    v.read_losses()
    assert torch.equal(v.q_group.grads, g_f)
    # layer-by-layer kernels on the same rows
    monkeypatch.setenv("PORL_CQL_FUSED", "0")
    w, _, _, _ = make(S, A, n_rows, 3)
    assert not w.fused
    w.learn_indices(idx)
    l_w = w.read_losses()
    np.testing.assert_allclose(l_f, l_w, rtol=1e-5)
    assert rel_l2(g_f, w.q_group.grads) <= 1e-3      # unfiltered rows: a ReLU flip moves the gradient by O(1/B)


def test_fused_flags_bad_rows_and_actions(mode):
    """Out-of-range actions / rows are clamped and reported through the status word, as in porl_cql_head / K1."""
    from porl_b200 import _C
    t, o, b, acts = make(22, 5, 64, 2)
    status = t.ctx.status_tensor()
    status.zero_()
    t.replay_buffer.actions[3] = 17
    t.learn_indices(torch.arange(64))
    t.read_losses()
    assert int(status[0]) & _C.FLAG_BAD_ACTION
    status.zero_()
    t.replay_buffer.actions[3] = 1
    idx = torch.arange(64)
    idx[5] = 10 ** 6
    t.learn_indices(idx)
    t.read_losses()
    assert int(status[0]) & _C.FLAG_BAD_INDEX and not int(status[0]) & _C.FLAG_BAD_ACTION
    status.zero_()


def test_fused_data_parallel_shards_sum_to_the_global_gradient(monkeypatch, mode):
    """two ranks take halves of one global minibatch and scale by 1/B_global; the sum of their buckets
    (what the peer-memory / NCCL gradient sum produces) equals the single-process gradient on all rows up to
    summation order (<= 1e-6 rel-L2), and the summed loss tails equal the global losses."""
    from porl_b200 import dist as pdist
    S, A, B, n_rows = 362, 5, 512, 2000
    t, o, b, acts = make(S, A, n_rows, 21)
    idx = torch.as_tensor(np.random.default_rng(2).choice(n_rows, B, replace=False)).to(DEV)
    rb = t.replay_buffer
    args = (rb._states, rb._next_states, rb.ld, rb.actions, rb.rewards, rb.dones)
    t._fused_grad(*args, idx, rb.capacity, B)
    torch.cuda.synchronize()
    g_all, l_all = t.q_group.grads.clone(), t.q_group.tail[0:3].clone()
    monkeypatch.setattr(pdist, "world_size", lambda: 2)
    parts, tails = [], []
    for r in range(2):
        t._fused_grad(*args, idx[r * B // 2:(r + 1) * B // 2].contiguous(), rb.capacity, B // 2)
        torch.cuda.synchronize()
        parts.append(t.q_group.grads.clone())
        tails.append(t.q_group._local_tail[0:3].clone())
    assert rel_l2(parts[0] + parts[1], g_all) <= 1e-6
    np.testing.assert_allclose((tails[0] + tails[1]).cpu().numpy(), l_all.cpu().numpy(), rtol=2e-6, atol=1e-8)


def test_adam_folded_into_the_reduction_is_bitwise_equal_to_the_separate_optimizer_launch(monkeypatch, mode):
    """porl_cql_fused_step (Adam + hard target copy + optimizer clock inside the gradient reduction: two launches per
    update) against porl_cql_fused_grad + porl_adam_step (three launches), six updates with a hard copy every third:
    parameters, target net, moments, optimizer clock and losses must be bitwise equal."""
    S, A, n_rows, B = 362, 5, 4000, 512
    runs = []
    for fold in ("1", "0"):
        monkeypatch.setenv("PORL_CQL_FOLD_ADAM", fold)
        t, _, _, _ = make(S, A, n_rows, seed=9)
        assert t.fused and t._two_launch() == (fold == "1")
        g = torch.Generator().manual_seed(31)
        launches0 = t.ctx.launches
        losses = []
        for _ in range(6):
            t.learn_indices(torch.randint(0, n_rows, (B,), generator=g), hard_every=3)
            losses.append(t.read_losses())
        clock = t.q_group.read_clock()
        runs.append((t.ctx.launches - launches0, losses, (int(clock.step), float(clock.lr)),
                     {k: v.clone() for k, v in t.q_network.state_dict().items()},
                     {k: v.clone() for k, v in t.target_network.state_dict().items()}, t.q_group.m.clone(),
                     t.q_group.v.clone(), t.q_group.grads.clone()))
    (n1, l1, c1, p1, t1, m1, v1, g1), (n0, l0, c0, p0, t0, m0, v0, g0) = runs
    assert n1 == 12 and n0 == 18, (n1, n0)               # 2 vs 3 launches per update
    assert l1 == l0 and c1 == c0 and c1[0] == 6
    for k in p1:
        assert torch.equal(p1[k], p0[k]), k
        assert torch.equal(t1[k], t0[k]), k
    assert torch.equal(m1, m0) and torch.equal(v1, v0) and torch.equal(g1, g0)
    any_moved = any(not torch.equal(p1[k], t1[k]) for k in p1)
    assert not any_moved                                  # update 6 ended with a hard copy: target == online
