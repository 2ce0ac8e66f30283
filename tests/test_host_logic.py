"""CPU-only checks of the boundary and the host logic (no compute without a GPU):
the C-ABI library loads and exports every symbol include/porl_b200.h declares, the product path
fails loudly without an sm_100 device, parameter containers expose the reference's state_dict
layout over the flat buffers, the update issues the expected kernel sequence, samplers reproduce
the reference's index streams, and the DP decomposition (shard + scale by 1/B_global + sum
all-reduce) equals the single-process update (gloo, world size 2, oracle as per-rank compute)."""
import os
import re

import numpy as np
import pytest
import torch

from oracle import synth
from oracle.buffer import epoch_batches, slice_row
from oracle.steps import OracleCQL, OraclePOR, OracleSORL, dims

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from porl_b200 import _C
    header = open(os.path.join(ROOT, "include", "porl_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(porl_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    assert declared == set(_C.EXPORTS)
    for name in declared:
        assert hasattr(_C.lib, name), name
    assert _C.lib.porl_version() == 100


def test_no_cpu_fallback():
    from porl_b200 import _C
    with pytest.raises(_C.PorlError, match="no CPU fallback"):
        _C.context("cpu")
    if not torch.cuda.is_available():
        with pytest.raises(_C.PorlError, match="no CUDA device"):
            _C.Context(0)


def test_mlp_geometry_is_padded_and_aligned():
    import ctypes
    from porl_b200 import _C, nets

    class Ctx:
        handle = ctypes.c_void_p(1)

    plan = nets.FlatMLP(Ctx(), [362, 512, 512, 1], n_nets=2, layer_norm=True)
    assert plan.out_ld == 4 and plan.in_ld == 384
    per_net = 512 * 384 + 512 + 2 * 512 + 512 * 512 + 512 + 2 * 512 + 1 * 512 + 4
    assert plan.n_floats == 2 * per_net
    flat = torch.arange(plan.n_floats, dtype=torch.float32)
    seen = torch.zeros(plan.n_floats, dtype=torch.bool)
    for net in range(2):
        for layer in range(3):
            kinds = [_C.PARAM_WEIGHT, _C.PARAM_BIAS] + ([_C.PARAM_LN_WEIGHT, _C.PARAM_LN_BIAS] if layer < 2 else [])
            for kind in kinds:
                off, ld = plan.layout(net, layer, kind)
                assert off % 4 == 0                                  # 16-byte aligned tensors
                v = plan.view(flat, net, layer, kind)
                idx = v.long().flatten()
                assert not seen[idx].any()                           # tensors do not overlap
                seen[idx] = True
    B = 256
    # per net, per hidden layer: h (B x 512), relu bits (B x 16 words), LayerNorm xhat (B x 512) and rstd (B); the
    # tcgen05 kernel reads MN-major operands in place, so no transposed copies are kept
    acts = 2 * 2 * (B * 512 + B * 16 + B * 512 + B)
    assert plan.saved_floats(B) == acts + 4
    with pytest.raises(_C.PorlError):
        plan.layout(0, 2, _C.PARAM_LN_WEIGHT)                        # no LayerNorm after the head


@pytest.mark.parametrize("kind", ["por", "sorl"])
def test_update_goes_through_the_whole_step_entry_point(dry_porl, kind):
    """update() = pack + ONE call of porl_por_step / porl_sorl_step (include/porl_b200.h) whose descriptor carries the
    agent's buffers and hyper-parameters; attribute changes (sweeps set tau / alpha) reach the next call."""
    import ctypes
    from porl_b200 import POR, SORL, _C
    S, H, A, B = 22, 32, 2, 16
    args = dims(S, H, 2, False, 256, A)
    agent = (POR if kind == "por" else SORL)(args, max_steps=7, tau=0.9, alpha=10.0, device=torch.device("cpu"))
    b = slice_row(torch.from_numpy(synth.laser_rows(B, S, 1, A)), S, A)
    agent.tau = 0.7
    dry_porl.clear()
    if kind == "por":
        agent.por_residual_update(b["observations"], b["next_observations"], b["rewards"], b["terminals"])
    else:
        agent.update(b["observations"], b["actions"], b["rewards"], b["next_observations"], b["terminals"])
    names = [c[0] for c in dry_porl]
    assert names == ["porl_pack_rows"] * (4 if kind == "por" else 5) + [f"porl_{kind}_step"]
    a = dry_porl[-1][1]
    d = ctypes.cast(a[1], ctypes.POINTER(_C.AgentStep)).contents
    assert a[2] == B and d.world == 1 and d.tau == pytest.approx(0.7) and d.alpha == pytest.approx(10.0)
    assert d.awr_mode == (0 if kind == "por" else 1) and d.tanh_out == (0 if kind == "por" else 1)
    assert d.policy_out == (S if kind == "por" else A)
    assert d.value.params == agent.v_group.params.data_ptr() and d.value.grads == agent.v_group.grads.data_ptr()
    assert d.policy.loss == agent.p_group.loss_ptr and d.value.n == agent.v_group.n and not d.value.peers
    tgt = agent.v_target if kind == "por" else agent.v_tgt
    assert d.target_params == tgt.flat.data_ptr() and d.ema_beta == pytest.approx(0.005)
    assert d.log_std_offset == (agent.goal_policy if kind == "por" else agent.policy).log_std_offset
    ws_floats = a[11] if kind == "por" else a[11]
    assert ws_floats == _C.lib.porl_step_workspace_floats(ctypes.byref(d), B) > 0
    pb = agent._packed[B]
    assert a[3] == pb.obs.data_ptr() and a[4] == pb.next_obs.data_ptr() and a[5] == pb.ld_obs
    assert a[6] == (pb.next_obs.data_ptr() if kind == "por" else pb.actions.data_ptr())


@pytest.mark.parametrize("kind", ["por", "sorl"])
def test_agent_state_dict_layout_and_kernel_sequence(dry_porl, kind, monkeypatch):
    from porl_b200 import POR, SORL
    monkeypatch.setenv("PORL_WHOLE_STEP", "0")        # the host-sequenced path (kept for the NCCL all-reduce fallback)
    S, H, A, B = 22, 32, 2, 16
    args = dims(S, H, 2, False, 256, A)
    cls, ocls = (POR, OraclePOR) if kind == "por" else (SORL, OracleSORL)
    agent = cls(args, max_steps=7, tau=0.9, alpha=10.0, device=torch.device("cpu"))
    oracle = ocls(args, max_steps=7, tau=0.9, alpha=10.0)
    assert list(agent.state_dict().keys()) == list(oracle.state_dict().keys())
    sd = synth.det_state_dict(oracle, 3)
    agent.load_state_dict(sd)
    for k, v in agent.state_dict().items():
        assert torch.equal(v, sd[k]), k
    # the named parameters are views of the flat buffers the kernels read
    value = agent.vf if kind == "por" else agent.v_net
    w = value.v2[2].weight
    off, ld = value.plan.layout(1, 1, 0)
    assert w.data_ptr() == agent.v_group.params.data_ptr() + 4 * off and w.stride() == (ld, 1)
    assert float(agent.v_group.params.abs().sum()) == pytest.approx(
        float(sum(v.abs().sum() for k, v in sd.items() if k.startswith(("vf.", "v_net.")))), rel=1e-5)
    # the target network is a private copy (deepcopy(vf).requires_grad_(False))
    tgt = agent.v_target if kind == "por" else agent.v_tgt
    assert tgt.flat.data_ptr() != agent.v_group.params.data_ptr()
    assert not any(p.requires_grad for p in tgt.parameters())
    assert all(p.requires_grad for p in value.parameters())
    # gradients are exposed on the parameters as views of the bucket
    assert w.grad.data_ptr() == agent.v_group.grads.data_ptr() + 4 * off
    rows = torch.from_numpy(synth.laser_rows(B, S, 1, A))
    b = slice_row(rows, S, A)
    dry_porl.clear()
    if kind == "por":
        agent.por_residual_update(b["observations"], b["next_observations"], b["rewards"], b["terminals"])
        n_pack = 4
    else:
        agent.update(b["observations"], b["actions"], b["rewards"], b["next_observations"], b["terminals"])
        n_pack = 5
    names = [c[0] for c in dry_porl]
    assert names == ["porl_pack_rows"] * n_pack + [
        "porl_mlp_forward", "porl_expectile_twin_loss", "porl_mlp_backward", "porl_adam_step",
        "porl_mlp_forward", "porl_awr_gauss_nll", "porl_mlp_backward", "porl_adam_step"]
    adam_v, adam_p = [c[1] for c in dry_porl if c[0] == "porl_adam_step"]
    assert adam_v[10] == tgt.flat.data_ptr() and adam_v[11] == pytest.approx(0.005)     # Polyak on the value group
    assert adam_p[10] is None and adam_p[12] == 0
    awr = [c[1] for c in dry_porl if c[0] == "porl_awr_gauss_nll"][0]
    assert awr[11] == (0 if kind == "por" else 1) and awr[12] == (0 if kind == "por" else 1)
    assert awr[14] == (S if kind == "por" else A)


def test_cql_trainer_host_plumbing(dry_porl, tmp_path, monkeypatch):
    from porl_b200 import CQLTrainer, QNetwork
    t = CQLTrainer(state_size=22, action_size=5, gamma=0.99, epsilon=1.0, epsilon_min=0.05, epsilon_decay=0.99,
                   update_target_freq=10, device=torch.device("cpu"), network=QNetwork, log_dir=str(tmp_path))
    o = OracleCQL(22, 5)
    assert list(t.q_network.state_dict().keys()) == list(o.q_network.state_dict().keys())
    for (k, a), (_, b) in zip(t.q_network.state_dict().items(), o.q_network.state_dict().items()):
        assert a.shape == b.shape
    for a, b in zip(t.q_network.state_dict().values(), t.target_network.state_dict().values()):
        assert torch.equal(a, b)                       # target starts as a copy (dqn_trainer.py:69)
    assert t.batch_size == 64 and t.replay_buffer.capacity == 100000 and t.alpha == 1
    assert t.q_group.base_lr == 0.0005
    rows = synth.laser_rows(100, 22, 1)
    s = slice_row(rows, 22)
    t.replay_buffer.push_batch(s["observations"], synth.discrete_actions(100, 5, 2), s["rewards"],
                               s["next_observations"], s["terminals"])
    assert len(t.replay_buffer) == 100
    dry_porl.clear()
    assert t.fused                                      # 22 -> 64 -> 128 -> 64 -> 5 is served by the single-kernel update
    t.learn()
    assert [c[0] for c in dry_porl] == ["porl_cql_fused_step"]       # single process: Adam rides in the gradient reduction
    fused = dry_porl[0][1]
    assert fused[6] == 32 and fused[11] == 100000 and fused[14] == 64 and fused[23] == 0     # ld, rows, B, hard_every
    assert t.training_step == 1
    monkeypatch.setenv("PORL_CQL_FOLD_ADAM", "0")        # the separate optimizer launch (what data parallelism uses)
    dry_porl.clear()
    t.learn()
    assert [c[0] for c in dry_porl] == ["porl_cql_fused_grad", "porl_adam_step"]
    fused = dry_porl[0][1]
    assert fused[6] == 32 and fused[11] == 100000 and fused[14] == 64 and fused[15] == 64   # ld, rows, B, B_norm
    assert t.training_step == 2
    monkeypatch.setenv("PORL_CQL_FUSED", "0")            # layer-by-layer path, kept for other network shapes
    dry_porl.clear()
    t.learn()
    assert [c[0] for c in dry_porl] == ["porl_gather_soa", "porl_mlp_forward", "porl_cql_head",
                                       "porl_mlp_backward", "porl_adam_step"]
    assert t.training_step == 3


def test_epoch_sampler_reproduces_dataloader_stream():
    from porl_b200.dataloader.dataset import EpochSampler
    n, bs = 103, 16
    g1, g2 = torch.Generator().manual_seed(5), torch.Generator().manual_seed(5)
    ds = torch.utils.data.TensorDataset(torch.arange(n))
    ref = [b[0] for b in torch.utils.data.DataLoader(ds, batch_size=bs, shuffle=True, generator=g1)]
    # DataLoader draws one extra int64 seed from its generator before the sampler's randperm
    _ = torch.empty((), dtype=torch.int64).random_(generator=g2)
    got = list(EpochSampler(n, bs, device="cpu", generator=g2))
    assert len(got) == len(ref) == 7 and got[-1].numel() == 7          # ragged last batch kept
    for a, b in zip(got, ref):
        assert torch.equal(a, b)
    g3 = torch.Generator().manual_seed(5)
    _ = torch.empty((), dtype=torch.int64).random_(generator=g3)
    assert all(torch.equal(a, b) for a, b in zip(epoch_batches(n, bs, g3), ref))


def test_synthetic_rows_shape_and_domain():
    from porl_b200.dataloader.dataset import synthetic_rows
    rows = synthetic_rows(1000, 362, 1)
    assert rows.shape == (1000, 728) and rows.dtype == np.float32
    scan = rows[:, :360]
    assert ((scan >= 0.13) & (scan <= 3.5) | (scan == 10.0)).all()       # state2costmap's valid domain
    assert set(np.unique(rows[:, 725])) <= {0.0, 1.0}
    assert np.abs(rows[:, 726:]).max() <= 1.0
    assert synthetic_rows(10, 365, 1).shape == (10, 734)


def test_shard_indices():
    from porl_b200 import dist as pdist
    idx = torch.arange(12)
    assert torch.equal(pdist.shard_indices(idx, 1, 4), torch.tensor([3, 4, 5]))
    assert torch.equal(torch.cat([pdist.shard_indices(idx, r, 3) for r in range(3)]), idx)
    with pytest.raises(ValueError):
        pdist.shard_indices(idx, 0, 5)


def _dp_worker(rank, world, port, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    import torch.distributed as dist
    from porl_b200 import dist as pdist
    torch.set_num_threads(1)
    pdist.init_from_env("gloo")
    S, H, A, B = 22, 32, 2, 32
    args = dims(S, H, 2, False, 256, A)
    agent = OracleSORL(args, max_steps=10, tau=0.8, alpha=3.0)
    agent.load_state_dict(synth.det_state_dict(agent, 4))
    rows = torch.from_numpy(synth.laser_rows(B, S, 6, A))
    gidx = torch.arange(B)
    b = slice_row(rows[pdist.shard_indices(gidx)], S, A)
    # per-rank value-loss gradient with per-sample terms scaled by 1/B_global (what the loss kernels do)
    with torch.no_grad():
        tv = b["rewards"] + (1 - b["terminals"]) * agent.discount * agent.v_tgt(b["next_observations"])
    v1, v2 = agent.v_net.both(b["observations"])
    loss = sum((torch.abs(agent.tau - ((tv - v) < 0).float()) * (tv - v) ** 2).sum() for v in (v1, v2)) / (2 * B)
    loss.backward()
    params = list(agent.v_net.parameters())
    bucket = torch.cat([p.grad.flatten() for p in params] + [loss.detach().reshape(1)])
    pdist.allreduce_sum_(bucket)
    if rank == 0:
        torch.save(bucket, os.path.join(tmp, "bucket.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_dp_bucket_allreduce_equals_single_process(tmp_path):
    import torch.multiprocessing as mp
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_dp_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    bucket = torch.load(os.path.join(tmp_path, "bucket.pt"))
    S, H, A, B = 22, 32, 2, 32
    agent = OracleSORL(dims(S, H, 2, False, 256, A), max_steps=10, tau=0.8, alpha=3.0)
    agent.load_state_dict(synth.det_state_dict(agent, 4))
    b = slice_row(torch.from_numpy(synth.laser_rows(B, S, 6, A)), S, A)
    with torch.no_grad():
        tv = b["rewards"] + (1 - b["terminals"]) * agent.discount * agent.v_tgt(b["next_observations"])
    v1, v2 = agent.v_net.both(b["observations"])
    from oracle.steps import expectile_l2
    loss = (expectile_l2(tv - v1, agent.tau) + expectile_l2(tv - v2, agent.tau)) / 2
    loss.backward()
    ref = torch.cat([p.grad.flatten() for p in agent.v_net.parameters()] + [loss.detach().reshape(1)])
    # DP(2) vs single process on the same global batch: only the summation order differs
    assert float((bucket - ref).norm() / ref.norm()) <= 1e-6
    assert abs(float(bucket[-1]) - float(ref[-1])) <= 1e-6 * abs(float(ref[-1]))
