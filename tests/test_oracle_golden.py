"""Pin the oracle (oracle/) to outputs of the reference itself (tests/golden/*.npz, made by
tests/golden/make_golden.py from the unmodified /root/reference).  CPU only."""
import numpy as np
import pytest
import torch

from oracle import formulas, synth
from oracle.buffer import RingBuffer, slice_row
from oracle.steps import OracleCQL, OraclePOR, OracleSORL, dims

torch.set_num_threads(1)


def _run_agent(kind, g):
    S, H, B, steps, max_steps, seed, A = (int(x) for x in g["meta"])
    tau, alpha = (float(x) for x in g["hyper"])
    cls = OraclePOR if kind == "por" else OracleSORL
    agent = cls(dims(S, H, 2, False, 256, A), max_steps=max_steps, tau=tau, alpha=alpha)
    agent.load_state_dict(synth.det_state_dict(agent, seed))
    rows = torch.from_numpy(synth.laser_rows(B * steps, S, seed + 1, A))
    losses = []
    for i in range(steps):
        b = slice_row(rows[i * B:(i + 1) * B], S, A)
        if kind == "por":
            losses.append(agent.por_residual_update(b["observations"], b["next_observations"], b["rewards"], b["terminals"]))
        else:
            losses.append(agent.update(b["observations"], b["actions"], b["rewards"], b["next_observations"], b["terminals"]))
    return agent, np.array(losses)


@pytest.mark.parametrize("name,kind", [("por_small", "por"), ("sorl_small", "sorl"), ("sorl_small_tau5", "sorl")])
def test_agent_small_matches_reference(golden, name, kind):
    g = golden(name)
    agent, losses = _run_agent(kind, g)
    np.testing.assert_allclose(losses, g["losses"], rtol=1e-6, atol=0)
    sd = agent.state_dict()
    keys = [k[len("final/"):] for k in g if k.startswith("final/")]
    assert sorted(keys) == sorted(sd.keys())
    for k in keys:
        assert tuple(sd[k].shape) == g["final/" + k].shape and sd[k].dtype == torch.float32
        np.testing.assert_allclose(sd[k].numpy(), g["final/" + k], rtol=1e-5, atol=1e-7, err_msg=k)


@pytest.mark.parametrize("name,kind", [("por_real", "por"), ("sorl_real", "sorl")])
def test_agent_real_dims_matches_reference(golden, name, kind):
    g = golden(name)
    agent, losses = _run_agent(kind, g)
    np.testing.assert_allclose(losses, g["losses"], rtol=2e-6, atol=0)
    for k, v in agent.state_dict().items():
        v = v.double().flatten()
        got = np.array([v.sum().item(), v.abs().sum().item(), *v[:6].tolist()])
        ref = g["sum/" + k]
        np.testing.assert_allclose(got[1:], ref[1:], rtol=1e-5, atol=1e-7, err_msg=k)
        assert abs(got[0] - ref[0]) <= 1e-5 * ref[1] + 1e-7, k


@pytest.mark.parametrize("name", ["cql_small", "cql_real"])
def test_cql_matches_reference(golden, name):
    g = golden(name)
    S, A, B, n_rows, iters, freq, seed = (int(x) for x in g["meta"])
    rb = RingBuffer(100000, (S,))
    rows = synth.laser_rows(n_rows, S, seed + 1)
    acts = synth.discrete_actions(n_rows, A, seed + 2)
    b = slice_row(rows, S)
    for i in range(n_rows):
        rb.push(b["observations"][i], acts[i], b["rewards"][i], b["next_observations"][i], bool(b["terminals"][i]))
    t = OracleCQL(S, A, update_target_freq=freq, replay_buffer=rb, batch_size=B)
    t.q_network.load_state_dict(synth.det_state_dict(t.q_network, seed))
    t.target_network.load_state_dict(t.q_network.state_dict())
    np.random.seed(seed + 3)
    state = np.random.get_state()
    idx = np.stack([rb.draw_indices(B) for _ in range(iters)])
    np.testing.assert_array_equal(idx, g["indices"])  # same sampler stream as the reference
    np.random.set_state(state)
    losses = t.train_offline(iters)
    np.testing.assert_allclose(losses, g["losses"], rtol=2e-6)
    for k, v in t.q_network.state_dict().items():
        np.testing.assert_allclose(v.numpy(), g["q/" + k], rtol=1e-5, atol=1e-7, err_msg=k)
    for k, v in t.target_network.state_dict().items():
        np.testing.assert_allclose(v.numpy(), g["target/" + k], rtol=1e-5, atol=1e-7, err_msg=k)


def test_replay_buffer_matches_reference(golden):
    g = golden("replay_buffer")
    seed = int(g["seed"])
    rb = RingBuffer(50, (7,))
    rng = np.random.RandomState(seed)
    for i in range(70):
        rb.push(rng.rand(7).astype(np.float32), int(rng.randint(5)), float(rng.randn()), rng.rand(7).astype(np.float32), bool(rng.rand() < 0.2))
    assert len(rb) == int(g["size"]) and rb.position == int(g["position"])
    np.random.seed(seed)
    s, a, r, s2, d = rb.sample(16)
    assert a.dtype == torch.int64 and s.dtype == torch.float32
    for got, key in ((s, "states"), (a, "actions"), (r, "rewards"), (s2, "next_states"), (d, "dones")):
        np.testing.assert_array_equal(got.numpy(), g[key])  # bit-exact gather


def test_cosine_lr_matches_torch(golden):
    g = golden("cosine_lr")
    lrs = formulas.cosine_lr_sequence(float(g["base"]), int(g["t_max"]), len(g["lrs"]))
    np.testing.assert_allclose(lrs, g["lrs"], rtol=1e-12, atol=1e-20)
