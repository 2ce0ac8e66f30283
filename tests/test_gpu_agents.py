"""Step-level parity of the drop-in agents (B200, -m gpu) against
  (1) the committed outputs of the UNMODIFIED reference (tests/golden/*.npz), and
  (2) the CPU oracle run side by side on the same seeded inputs (teacher-forced and free-running).

Tolerances (fp32; BASELINE.json north_star: <= 1e-5 relative per step): scalar losses 1e-5 relative;
per-tensor gradient rel-L2 <= 1e-5; parameters after n Adam steps per oracle/compare.py
(per-tensor rel-L2 <= 5e-5, <= 0.01 % of elements outside rtol 5e-5 / atol 2e-7, every element within
2*lr*n -- Adam turns rounding noise on near-zero gradients into +-lr-sized steps).
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import synth  # noqa: E402
from oracle.compare import assert_params_close  # noqa: E402
from oracle.buffer import slice_row  # noqa: E402
from oracle.steps import OracleCQL, OraclePOR, OracleSORL, dims  # noqa: E402

DEV = "cuda:0"


def rel_l2(a, b):
    a, b = a.detach().double().cpu().flatten(), b.detach().double().cpu().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def make_agent(kind, S, H, A, max_steps, tau, alpha, seed, gemm_mode=0, layer_norm=False):
    from porl_b200 import POR, SORL
    args = dims(S, H, 2, layer_norm, 256, A)
    cls = POR if kind == "por" else SORL
    agent = cls(args, max_steps=max_steps, tau=tau, alpha=alpha, device=torch.device(DEV))
    agent.ctx.set_gemm_mode(gemm_mode)
    ocls = OraclePOR if kind == "por" else OracleSORL
    oracle = ocls(args, max_steps=max_steps, tau=tau, alpha=alpha)
    sd = synth.det_state_dict(oracle, seed)
    oracle.load_state_dict(sd)
    agent.load_state_dict(sd)
    return agent, oracle


def relu_margin(oracle, kind, obs):
    """Per sample: the smallest |pre-activation| over every ReLU the update back-propagates through
    (online value nets and policy net on `obs`, oracle weights, fp64).  A sample with a pre-activation inside
    fp32 rounding noise of zero can take either side of the ReLU in two correct fp32 implementations; its
    gradient contribution then differs by O(1/B), which says nothing about either implementation."""
    import copy
    value = oracle.vf if kind == "por" else oracle.v_net
    policy = oracle.goal_policy if kind == "por" else oracle.policy
    margin = torch.full((obs.shape[0],), float("inf"), dtype=torch.float64)
    for seq in (value.v1, value.v2, policy.net):
        x = obs.double()
        for layer in copy.deepcopy(seq).double():
            if isinstance(layer, torch.nn.ReLU):
                margin = torch.minimum(margin, x.abs().min(dim=1).values)
            x = layer(x)
    return margin


def pick_rows(oracle, kind, pool, S, B, min_margin=2e-5):
    """First B rows of `pool` whose ReLU margin under the oracle's current weights exceeds min_margin."""
    with torch.no_grad():
        keep = torch.nonzero(relu_margin(oracle, kind, pool[:, :S]) > min_margin).flatten()
    assert keep.numel() >= B, "row pool too small"
    return pool[keep[:B]]


def step(agent, kind, b):
    if kind == "por":
        return agent.por_residual_update(b["observations"], b["next_observations"], b["rewards"], b["terminals"])
    return agent.update(b["observations"], b["actions"], b["rewards"], b["next_observations"], b["terminals"])


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("name,kind", [("por_small", "por"), ("sorl_small", "sorl"), ("sorl_small_tau5", "sorl"),
                                       ("por_real", "por"), ("sorl_real", "sorl")])
def test_agent_matches_reference_golden(golden, name, kind, mode):
    g = golden(name)
    S, H, B, steps, max_steps, seed, A = (int(x) for x in g["meta"])
    tau, alpha = (float(x) for x in g["hyper"])
    agent, _ = make_agent(kind, S, H, A, max_steps, tau, alpha, seed, gemm_mode=mode)
    try:
        rows = torch.from_numpy(synth.laser_rows(B * steps, S, seed + 1, A)).to(DEV)
        losses = []
        for i in range(steps):
            losses.append(step(agent, kind, slice_row(rows[i * B:(i + 1) * B], S, A)))
        np.testing.assert_allclose(np.array(losses), g["losses"], rtol=1e-5, atol=0)
        sd = agent.state_dict()
        full = [k[len("final/"):] for k in g if k.startswith("final/")]
        if full:
            assert sorted(full) == sorted(sd.keys())
            for k in full:
                assert tuple(sd[k].shape) == g["final/" + k].shape and sd[k].dtype == torch.float32
                assert_params_close(sd[k], g["final/" + k], lr=1e-4, n_steps=steps, name=k)
        else:
            for k, v in sd.items():
                v = v.double().flatten().cpu()
                ref = g["sum/" + k]
                got = np.array([v.sum().item(), v.abs().sum().item(), *v[:6].tolist()])
                # per-tensor checksums (sum, abs-sum) and the first six entries of the reference run
                assert abs(got[1] - ref[1]) <= 5e-5 * ref[1] + 1e-6, k
                assert abs(got[0] - ref[0]) <= 5e-5 * ref[1] + 1e-6, k
                np.testing.assert_allclose(got[2:], ref[2:], rtol=5e-5, atol=2 * 1e-4 * steps, err_msg=k)
        sched = agent.goal_lr_schedule if kind == "por" else agent.lr_schedule
        np.testing.assert_allclose(sched.get_last_lr(), g["last_lr"], rtol=1e-12)
    finally:
        agent.ctx.set_gemm_mode(0)


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("kind,S,H,A,B,ln", [("sorl", 362, 512, 2, 256, False), ("por", 365, 256, 2, 64, False),
                                             ("sorl", 30, 64, 2, 48, True)])
def test_agent_teacher_forced_vs_oracle(kind, S, H, A, B, ln, mode):
    """Each step starts from the ORACLE's state (teacher forcing): compare losses, gradients and
    the parameters after one update."""
    agent, oracle = make_agent(kind, S, H, A, 1000, 0.8, 3.0 if kind == "sorl" else 10.0, 5, gemm_mode=mode,
                               layer_norm=ln)
    try:
        rows = torch.from_numpy(synth.laser_rows(B * 6, S, 17, A))
        for i in range(3):
            agent.load_state_dict(oracle.state_dict())
            chunk = pick_rows(oracle, kind, rows[i * 2 * B:(i + 1) * 2 * B], S, B)
            want = step(oracle, kind, slice_row(chunk.clone(), S, A))
            got = step(agent, kind, slice_row(chunk.to(DEV), S, A))
            np.testing.assert_allclose(got, want, rtol=1e-5)
            ref_grads = {k: p.grad for k, p in oracle.named_parameters() if p.grad is not None}
            for k, p in agent.named_parameters():
                if k in ref_grads:
                    assert rel_l2(p.grad, ref_grads[k]) <= 1e-5, (i, k)
            if i == 0:   # optimizer moments are zero on both sides only at the first step
                osd = oracle.state_dict()
                for k, v in agent.state_dict().items():
                    assert_params_close(v, osd[k], lr=1e-4, n_steps=1, name=k)
    finally:
        agent.ctx.set_gemm_mode(0)


def test_state_dict_interchange_and_api_surface():
    agent, oracle = make_agent("por", 22, 32, 2, 10, 0.9, 10.0, 3)
    assert list(agent.state_dict().keys()) == list(oracle.state_dict().keys())
    for (k, a), (_, o) in zip(agent.state_dict().items(), oracle.state_dict().items()):
        assert tuple(a.shape) == tuple(o.shape) and a.dtype == o.dtype
        assert torch.equal(a.cpu(), o)
    oracle.load_state_dict({k: v.cpu() for k, v in agent.state_dict().items()})     # and back
    for attr in ("goal_policy", "vf", "v_target", "v_optimizer", "goal_policy_optimizer", "goal_lr_schedule",
                 "tau", "alpha", "discount", "beta", "step", "pretrain_step"):
        assert hasattr(agent, attr), attr
    assert not any(p.requires_grad for p in agent.v_target.parameters())
    sorl, so = make_agent("sorl", 22, 32, 2, 10, 0.9, 10.0, 3)
    assert list(sorl.state_dict().keys()) == list(so.state_dict().keys())
    x = torch.from_numpy(synth.laser_rows(5, 22, 1))[:, :22]
    np.testing.assert_allclose(sorl.select_action(x.to(DEV)), so.select_action(x), rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(sorl.v_net(x.to(DEV)).cpu().numpy(), so.v_net(x).detach().numpy(), rtol=1e-5, atol=1e-6)
    b = slice_row(torch.from_numpy(synth.laser_rows(8, 22, 2)), 22, 2)
    want = so.vf_update(b["observations"], b["actions"], b["rewards"], b["next_observations"], b["terminals"])
    got = sorl.vf_update(b["observations"], b["actions"], b["rewards"], b["next_observations"], b["terminals"])
    assert abs(got - want) <= 1e-5 * abs(want)


def test_update_does_not_modify_inputs_and_accepts_strided_views():
    agent, _ = make_agent("sorl", 22, 32, 2, 10, 0.9, 10.0, 3)
    rows = torch.from_numpy(synth.laser_rows(16, 22, 4)).to(DEV)
    keep = rows.clone()
    b = slice_row(rows, 22, 2)                     # column views with row stride 48
    step(agent, "sorl", b)
    assert torch.equal(rows, keep)


def test_free_running_drift_vs_oracle():
    """200 free-running updates: report drift; bound it loosely (the reference's own fp32-vs-fp64
    drift over the same horizon is the yardstick, printed beside it)."""
    kind, S, H, A, B, n = "sorl", 64, 64, 2, 64, 200
    agent, oracle = make_agent(kind, S, H, A, 1000, 0.7, 3.0, 9)
    o64 = OracleSORL(dims(S, H, 2, False, 256, A), max_steps=1000, tau=0.7, alpha=3.0)
    o64.load_state_dict(oracle.state_dict())
    o64.double()
    rows = torch.from_numpy(synth.laser_rows(B * n, S, 23, A))
    for i in range(n):
        chunk = rows[i * B:(i + 1) * B]
        la = step(agent, kind, slice_row(chunk.to(DEV), S, A))
        lo = step(oracle, kind, slice_row(chunk.clone(), S, A))
        l64 = step(o64, kind, slice_row(chunk.double(), S, A))
    flat = lambda sd: torch.cat([v.double().flatten().cpu() for v in sd.values()])
    ref64 = flat(o64.state_dict())
    d_gpu = float((flat(agent.state_dict()) - ref64).norm() / ref64.norm())
    d_ref = float((flat(oracle.state_dict()) - ref64).norm() / ref64.norm())
    print(f"drift after {n} updates vs fp64: gpu {d_gpu:.3e}, reference-fp32 {d_ref:.3e}; "
          f"last losses gpu {la} ref {lo} fp64 {l64}")
    assert d_gpu <= max(10 * d_ref, 1e-4)
    np.testing.assert_allclose(la, l64, rtol=1e-3)


@pytest.mark.parametrize("mode", [0, 1, "fused", "fused_mma"])
@pytest.mark.parametrize("name", ["cql_small", "cql_real"])
def test_cql_matches_reference_golden(golden, name, mode, monkeypatch):
    """mode 0 / 1: layer-by-layer kernels with FFMA / tcgen05 contractions; "fused" / "fused_mma": the single-kernel update with FFMA / mma.sync 3xTF32 tiles."""
    from porl_b200 import CQLTrainer, QNetwork
    monkeypatch.setenv("PORL_CQL_FUSED", "1" if str(mode).startswith("fused") else "0")
    mode = {"fused": 0, "fused_mma": 1}.get(mode, mode)
    g = golden(name)
    S, A, B, n_rows, iters, freq, seed = (int(x) for x in g["meta"])
    t = CQLTrainer(state_size=S, action_size=A, gamma=0.99, epsilon=1.0, epsilon_min=0.05, epsilon_decay=0.99,
                   update_target_freq=freq, device=torch.device(DEV), network=QNetwork, log_dir="/tmp/porl_b200_logs")
    t.ctx.set_gemm_mode(mode)
    try:
        t.batch_size = B
        ref_q = OracleCQL(S, A).q_network
        sd = synth.det_state_dict(ref_q, seed)
        t.q_network.load_state_dict(sd)
        t.target_network.load_state_dict(sd)
        rows = synth.laser_rows(n_rows, S, seed + 1)
        acts = synth.discrete_actions(n_rows, A, seed + 2)
        b = slice_row(rows, S)
        for i in range(n_rows):
            t.replay_buffer.push(b["observations"][i], acts[i], b["rewards"][i], b["next_observations"][i],
                                 bool(b["terminals"][i]))
        np.random.seed(seed + 3)
        losses = t.train_offline(num_iterations=iters, verbose=False)
        # the loss is td + penalty with the penalty a small difference of O(1) terms: absolute floor of a few fp32 ulps
        np.testing.assert_allclose(losses, g["losses"], rtol=1e-5, atol=2e-7)
        for k, v in t.q_network.state_dict().items():
            assert_params_close(v, g["q/" + k], lr=5e-4, n_steps=iters, name=k)
        for k, v in t.target_network.state_dict().items():
            assert_params_close(v, g["target/" + k], lr=5e-4, n_steps=iters, name=k)
        # API surface of cql_trainer.py
        st = torch.from_numpy(rows[:9, :S]).to(DEV)
        pen = t.compute_cql_penalty(st, torch.from_numpy(acts[:9]).to(DEV))
        oq = OracleCQL(S, A)
        oq.q_network.load_state_dict({k: v.cpu() for k, v in t.q_network.state_dict().items()})
        want = oq.compute_cql_penalty(torch.from_numpy(rows[:9, :S]), torch.from_numpy(acts[:9]))
        assert abs(float(pen) - float(want)) <= 1e-5 * abs(float(want))
        assert t.get_action(rows[0, :S]) == oq.get_action(rows[0, :S])
    finally:
        t.ctx.set_gemm_mode(0)


def test_cql_device_side_hard_copy_matches_host_loop():
    """The graph-friendly path (hard target copy inside the Adam kernel) equals train_offline's host loop."""
    from porl_b200 import CQLTrainer, QNetwork
    mk = lambda: CQLTrainer(state_size=22, action_size=5, gamma=0.99, epsilon=1.0, epsilon_min=0.05,
                            epsilon_decay=0.99, update_target_freq=3, device=torch.device(DEV), network=QNetwork,
                            log_dir="/tmp/porl_b200_logs")
    a, b = mk(), mk()
    b.q_network.load_state_dict(a.q_network.state_dict())
    b.target_network.load_state_dict(a.target_network.state_dict())
    rows = synth.laser_rows(300, 22, 8)
    acts = synth.discrete_actions(300, 5, 9)
    s = slice_row(rows, 22)
    for t in (a, b):
        t.replay_buffer.push_batch(s["observations"], acts, s["rewards"], s["next_observations"], s["terminals"])
        t.batch_size = 32
    np.random.seed(1)
    idx = [np.random.choice(300, 32, replace=False) for _ in range(7)]
    np.random.seed(1)
    a.train_offline(num_iterations=7, verbose=False)
    for i in range(7):
        pb = b.replay_buffer.gather_packed(torch.as_tensor(idx[i]))
        b.learn_packed(pb, hard_every=3)
    torch.cuda.synchronize()
    assert torch.equal(a.q_group.params, b.q_group.params)
    assert torch.equal(a.target_network.flat, b.target_network.flat)


@pytest.mark.parametrize("kind", ["sorl", "por"])
def test_training_checkpoint_resumes_bitwise(kind, tmp_path):
    """weights + both Adam groups + LR clocks in one checkpoint; a resumed run continues bit for bit
    (the reference saves weights only, por_train.py:97, so its resume restarts the optimizers and the cosine LR)."""
    S, H, A, B = 30, 64, 2, 32
    rows = torch.from_numpy(synth.laser_rows(B * 6, S, 23, A)).to(DEV)
    agent, _ = make_agent(kind, S, H, A, 50, 0.8, 3.0, 9)
    for i in range(3):
        step(agent, kind, slice_row(rows[i * B:(i + 1) * B], S, A))
    path = tmp_path / "train.pt"
    torch.save(agent.training_state_dict(), path)
    tail = [step(agent, kind, slice_row(rows[i * B:(i + 1) * B], S, A)) for i in range(3, 6)]
    fresh, _ = make_agent(kind, S, H, A, 50, 0.8, 3.0, 1234)             # different weights before the load
    fresh.load_training_state_dict(torch.load(path, map_location=DEV, weights_only=False))
    tail2 = [step(fresh, kind, slice_row(rows[i * B:(i + 1) * B], S, A)) for i in range(3, 6)]
    assert tail == tail2
    a, b = agent.state_dict(), fresh.state_dict()
    assert all(torch.equal(a[k], b[k]) for k in a)
    sched = (lambda ag: ag.goal_lr_schedule if kind == "por" else ag.lr_schedule)
    assert sched(agent).get_last_lr() == sched(fresh).get_last_lr() and sched(fresh).last_epoch == 6
    act = agent.select_action(rows[:5, :S])
    assert act.shape == (5, S if kind == "por" else A) and np.isfinite(act).all()


@pytest.mark.parametrize("kind", ["sorl", "por"])
def test_update_graph_replay_is_bitwise_equal_to_eager_launches(kind, monkeypatch):
    """`agent.update()` replays its ~25 launches as one CUDA graph from the third call at a batch size on
    (`_run_update`); same kernels, same arguments, so losses and parameters equal the eager path bit for bit, also across
    a change of batch size, of a hyper-parameter baked into the launches, and of a context option."""
    S, H, A, B = 30, 64, 2, 32
    rows = torch.from_numpy(synth.laser_rows(B * 12, S, 31, A)).to(DEV)
    graphed, _ = make_agent(kind, S, H, A, 50, 0.8, 3.0, 9)
    eager, _ = make_agent(kind, S, H, A, 50, 0.8, 3.0, 9)
    for i in range(10):
        lo, hi = (i * B, (i + 1) * B) if i != 6 else (i * B, i * B + 17)        # one ragged batch in between
        b = slice_row(rows[lo:hi], S, A)
        if i == 4:
            graphed.tau = eager.tau = 0.6
        if i == 8:
            graphed.ctx.set_pdl(False)
        monkeypatch.setenv("PORL_UPDATE_GRAPH", "1")
        lg = step(graphed, kind, b)
        monkeypatch.setenv("PORL_UPDATE_GRAPH", "0")
        le = step(eager, kind, b)
        assert lg == le, (i, lg, le)
    graphed.ctx.set_pdl(True)
    assert any(isinstance(g, torch.cuda.CUDAGraph) for g in graphed._graphs.values())
    assert not eager._graphs
    for (k, a), (_, b) in zip(graphed.state_dict().items(), eager.state_dict().items()):
        assert torch.equal(a, b), k


@pytest.mark.parametrize("kind,S,H,B", [("sorl", 362, 512, 256), ("por", 365, 256, 512), ("sorl", 362, 512, 4096)])
def test_multi_phase_launches_are_bitwise_equal_to_one_launch_per_contraction(kind, S, H, B):
    """gemm_tc.cu TcPhase: the hidden layers of a forward and {dW_l, dx_l, dW_l-1} of a backward as phases of one
    persistent launch (grid barriers in between) must give the bits of one launch per contraction: same tiles, same
    split-K cuts, same summation order -- over three free-running updates (any race across a barrier would show)."""
    A = 2
    rows = torch.from_numpy(synth.laser_rows(B * 3, S, 41, A)).to(DEV)
    outs = []
    for phases in (True, False):
        agent, _ = make_agent(kind, S, H, A, 1000, 0.8, 3.0 if kind == "sorl" else 10.0, 5, gemm_mode=1)
        agent.ctx.set_tc_phases(phases)
        try:
            before = agent.ctx.launches
            losses = [step(agent, kind, slice_row(rows[i * B:(i + 1) * B], S, A)) for i in range(3)]
            n_launch = agent.ctx.launches - before
            outs.append((losses, {k: v.clone() for k, v in agent.state_dict().items()}, n_launch))
        finally:
            agent.ctx.set_tc_phases(False)
            agent.ctx.set_gemm_mode(0)
    assert outs[0][0] == outs[1][0]
    for k, v in outs[0][1].items():
        assert torch.equal(v, outs[1][1][k]), k
    assert outs[0][2] < outs[1][2]          # fewer launches with phases


@pytest.mark.parametrize("S,H,B,mode", [(362, 512, 256, 1), (362, 512, 4096, 1), (30, 64, 48, 0)])
def test_fused_policy_head_step_matches_three_launches(S, H, B, mode, monkeypatch):
    """porl_policy_head_step (value heads + advantage-weighted Gaussian NLL + policy head backward in one kernel) against
    head forward, porl_awr_gauss_nll and head backward as three launches (PORL_FUSED_POLICY_HEAD=0): the same formulas,
    the batch reductions in a different order -- losses to 2e-6, parameters after three free-running SORL updates to fp32
    summation-order noise, and two launches fewer per update."""
    A = 2
    rows = torch.from_numpy(synth.laser_rows(B * 3, S, 43, A)).to(DEV)
    outs = []
    for fused in ("1", "0"):
        monkeypatch.setenv("PORL_FUSED_POLICY_HEAD", fused)
        monkeypatch.setenv("PORL_UPDATE_GRAPH", "0")
        agent, _ = make_agent("sorl", S, H, A, 1000, 0.8, 3.0, 5, gemm_mode=mode)
        try:
            before = agent.ctx.launches
            losses = [step(agent, "sorl", slice_row(rows[i * B:(i + 1) * B], S, A)) for i in range(3)]
            outs.append((losses, {k: v.clone() for k, v in agent.state_dict().items()}, agent.ctx.launches - before))
        finally:
            agent.ctx.set_gemm_mode(0)
    assert outs[0][2] == outs[1][2] - 6, (outs[0][2], outs[1][2])          # 2 launches fewer per update
    for lf, ls in zip(outs[0][0], outs[1][0]):
        np.testing.assert_allclose(lf, ls, rtol=2e-6)
    for k, v in outs[0][1].items():
        w = outs[1][1][k]
        if v.is_floating_point():
            assert float((v - w).abs().max()) <= 2e-6 * float(w.abs().max()) + 2e-7, k      # 3 Adam steps of 1e-4
        else:
            assert torch.equal(v, w), k
