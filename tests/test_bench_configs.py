"""Step-level parity AT THE CONFIGURATIONS bench.py TIMES (B200, -m gpu): SORL S=362/H=512 at batch 4096 and 256,
POR S=365/H=1024 at batch 256 and 4096 (por_train.py:127,139), SORL + FasterNet at batch 256, in both contraction
modes, against the CPU oracle (oracle/steps.py, pinned to the unmodified reference by tests/test_oracle_golden.py).

Per step, teacher-forced from the oracle's state: losses <= 1e-5 relative, per-tensor gradient rel-L2 <= 1e-5,
parameters per oracle/compare.py.  Two row sets per configuration:
  * rows whose fp64 ReLU margin exceeds 2e-5 (no ReLU of the update can land on different sides in two correct fp32
    implementations): the 1e-5 gradient bar is asserted as is;
  * the UNFILTERED rows (what the bench feeds): the number of rows inside the margin is reported, losses and
    post-step parameters keep their bars, and the gradient distance to the fp64 oracle is bounded against the
    reference's own fp32-vs-fp64 distance on the same rows, the yardstick two correct fp32
    implementations are apart by.  A unit that lands on the other side of a ReLU is a DISCRETE event (it switches one
    sample's contribution to one weight row on or off), each implementation draws its own few flips from the rows
    inside the margin, and a sample with a large TD error weighs more: measured on B200, POR S=365/H=1024/B=4096 (359
    of 4096 rows inside the margin) puts the fp32 reference 2.5e-6 and the FFMA path 1.1e-5 from fp64 on vf.v1.0.weight
    while every tensor agrees to <= 9e-7 on the filtered rows.  The bound is therefore 10x the reference's distance
    (not 3x: the ratio of two small-count draws), floored at the 1e-5 bar.
Plus a 1,000-update free-running drift run at the SORL dims with an asserted bound (tools/drift_10k.py is the
10,000-update version whose table is committed under profiles/), and the real multi-GPU checks when >= 2 GPUs exist.
"""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import synth  # noqa: E402
from oracle.buffer import slice_row  # noqa: E402
from oracle.compare import assert_params_close, rel_l2  # noqa: E402
from oracle.steps import OraclePOR, OracleSORL, dims  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEV = "cuda:0"
MARGIN = 2e-5


def _agents(kind, S, H, A, seed, mode):
    from porl_b200 import POR, SORL
    args = dims(S, H, 2, False, 256, A)
    hyper = dict(max_steps=1_000_000, tau=0.9, alpha=10.0)          # bench.py's hyper-parameters
    agent = (POR if kind == "por" else SORL)(args, device=torch.device(DEV), **hyper)
    agent.ctx.set_gemm_mode(mode)
    ocls = OraclePOR if kind == "por" else OracleSORL
    o32, o64 = ocls(args, **hyper), ocls(args, **hyper)
    sd = synth.det_state_dict(o32, seed)
    o32.load_state_dict(sd)
    o64.load_state_dict(sd)
    o64.double()
    agent.load_state_dict(sd)
    return agent, o32, o64


def _step(agent, kind, b):
    if kind == "por":
        return agent.por_residual_update(b["observations"], b["next_observations"], b["rewards"], b["terminals"])
    return agent.update(b["observations"], b["actions"], b["rewards"], b["next_observations"], b["terminals"])


def _margin(oracle64, kind, obs):
    """Smallest |pre-activation| over every ReLU the update back-propagates through, per sample (fp64 weights)."""
    value = oracle64.vf if kind == "por" else oracle64.v_net
    policy = oracle64.goal_policy if kind == "por" else oracle64.policy
    m = torch.full((obs.shape[0],), float("inf"), dtype=torch.float64)
    with torch.no_grad():
        for seq in (value.v1, value.v2, policy.net):
            x = obs.double()
            for layer in seq:
                if isinstance(layer, torch.nn.ReLU):
                    m = torch.minimum(m, x.abs().min(dim=1).values)
                x = layer(x)
    return m


def _grads(module):
    return {k: p.grad.detach().clone() for k, p in module.named_parameters() if p.grad is not None}


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("kind,S,H,B", [("sorl", 362, 512, 4096), ("sorl", 362, 512, 256), ("por", 365, 1024, 256),
                                        ("por", 365, 1024, 4096)])
def test_benchmarked_config_teacher_forced(kind, S, H, B, mode):
    A = 2
    agent, o32, o64 = _agents(kind, S, H, A, 11, mode)
    try:
        pool = torch.from_numpy(synth.laser_rows(5 * B, S, 31 + B, A))
        start = o32.state_dict()
        start = {k: v.clone() for k, v in start.items()}
        m = _margin(o64, kind, pool[:, :S])
        safe = pool[torch.nonzero(m > MARGIN).flatten()[:B]]
        assert safe.shape[0] == B, f"row pool too small: {int((m > MARGIN).sum())} of {pool.shape[0]} rows outside the margin"
        raw = pool[:B]
        inside = int((m[:B] <= MARGIN).sum())
        for label, rows in (("filtered", safe), ("unfiltered", raw)):
            for o in (o32, o64):
                o.load_state_dict(start)
            o64.double()
            agent.load_state_dict(start)
            want = _step(o32, kind, slice_row(rows.clone(), S, A))
            w64 = _step(o64, kind, slice_row(rows.double(), S, A))
            got = _step(agent, kind, slice_row(rows.to(DEV), S, A))
            np.testing.assert_allclose(got, want, rtol=1e-5, err_msg=f"{label} losses")
            g32, g64, worst, worst_ref = _grads(o32), _grads(o64), 0.0, 0.0
            for k, p in agent.named_parameters():
                if k not in g32:
                    continue
                d_gpu32 = rel_l2(p.grad, g32[k])
                d_gpu64, d_ref = rel_l2(p.grad, g64[k]), rel_l2(g32[k], g64[k])
                worst, worst_ref = max(worst, d_gpu64), max(worst_ref, d_ref)
                if label == "filtered":
                    assert d_gpu32 <= 1e-5, (label, k, d_gpu32)
                else:   # rows inside the ReLU margin may flip: bound by the reference's own fp32-vs-fp64 distance
                    assert d_gpu64 <= max(10.0 * d_ref, 1e-5), (label, k, d_gpu64, d_ref, inside)
            osd = o32.state_dict()                      # moments were zero on both sides: one Adam step
            for k, v in agent.state_dict().items():
                assert_params_close(v, osd[k], lr=1e-4, n_steps=1, name=f"{label}:{k}")
            print(f"[{kind} S={S} H={H} B={B} mode={mode} {label}] rows inside the ReLU margin: "
                  f"{0 if label == 'filtered' else inside}; worst gradient rel-L2 vs fp64 oracle: gpu {worst:.2e}, "
                  f"reference fp32 {worst_ref:.2e}; losses gpu {got} fp32 {want} fp64 {w64}")
    finally:
        agent.ctx.set_gemm_mode(0)


def test_sorl_fasternet_step_at_bench_batch():
    """bench.py's `also_encoder` configuration (the literal sorl_train.py:29-40 agent): one SORL update with the
    FasterNet scan encoder at batch 256, H = 512, tcgen05 contractions, against the oracle agent with the fp32 oracle
    encoder (recorded DropPath factors on both sides)."""
    from oracle import encoder
    from porl_b200 import SORL
    from porl_b200.agent.fasternet import FasterNet
    S, H, A, B = 362, 512, 2, 256
    args = dims(S, H, 2, False, 256, A)
    enc_sd = encoder.det_encoder_state_dict(21)
    g = torch.Generator().manual_seed(8)
    keeps = [[None] + [torch.bernoulli(torch.full((B,), 1 - p), generator=g) / (1 - p) for p in encoder.drop_probs()[1:]]
             for _ in range(2)]

    class Backbone:
        def __init__(self, sd, queue):
            self.sd, self.queue = sd, list(queue)

        def __call__(self, x):
            return encoder.fasternet_forward(self.sd, x, self.queue.pop(0))

    hyper = dict(max_steps=1_000_000, tau=0.9, alpha=10.0)
    oracle = OracleSORL(args, backbone=Backbone({k: v.clone() for k, v in enc_sd.items()}, keeps), **hyper)
    head_sd = synth.det_state_dict(oracle, 5)
    oracle.load_state_dict(head_sd)
    net = FasterNet(3, 256)
    agent = SORL(args, backbone=net, device=torch.device(DEV), **hyper)
    full = {("backbone." + k): v for k, v in encoder.order_like(enc_sd, net.state_dict().keys()).items()}
    full.update(head_sd)
    agent.load_state_dict(full)
    agent.backbone.replay_keep_scales(keeps)
    agent.ctx.set_gemm_mode(1)
    try:
        rows = torch.from_numpy(synth.laser_rows(B, S, 19, A))
        bo, bd = slice_row(rows.clone(), S, A), slice_row(rows.clone().to(DEV), S, A)
        want = oracle.update(bo["observations"], bo["actions"], bo["rewards"], bo["next_observations"], bo["terminals"])
        got = agent.update(bd["observations"], bd["actions"], bd["rewards"], bd["next_observations"], bd["terminals"])
        np.testing.assert_allclose(got, want, rtol=2e-5, atol=0)
        sd, osd = agent.state_dict(), oracle.state_dict()
        for k, v in osd.items():
            assert_params_close(sd[k], v, lr=1e-4, n_steps=1, name=k)
    finally:
        agent.ctx.set_gemm_mode(0)


def test_free_running_drift_1k_updates_at_sorl_dims():
    """1,000 free-running SORL updates at the benchmarked dims (S=362, H=512; batch 64 keeps the two CPU oracles at a
    few seconds), tcgen05 contractions: parameter distance to the fp64 oracle, next to the reference-fp32 oracle's own
    distance to it.  Asserted: the CUDA path drifts no more than 3x what the reference's fp32 arithmetic drifts."""
    kind, S, H, A, B, n = "sorl", 362, 512, 2, 64, 1000
    agent, o32, o64 = _agents(kind, S, H, A, 9, 1)
    try:
        rows = torch.from_numpy(synth.laser_rows(B * 250, S, 23, A))
        flat = lambda sd: torch.cat([v.double().flatten().cpu() for v in sd.values()])
        for i in range(n):
            chunk = rows[(i % 250) * B:((i % 250) + 1) * B]
            la = _step(agent, kind, slice_row(chunk.to(DEV), S, A))
            lo = _step(o32, kind, slice_row(chunk.clone(), S, A))
            l64 = _step(o64, kind, slice_row(chunk.double(), S, A))
        ref64 = flat(o64.state_dict())
        d_gpu = float((flat(agent.state_dict()) - ref64).norm() / ref64.norm())
        d_ref = float((flat(o32.state_dict()) - ref64).norm() / ref64.norm())
        print(f"drift after {n} free-running updates vs the fp64 oracle: gpu {d_gpu:.3e}, reference fp32 {d_ref:.3e}; "
              f"last losses gpu {la} fp32 {lo} fp64 {l64}")
        assert d_gpu <= 3.0 * d_ref + 1e-6, (d_gpu, d_ref)
        # free-running losses: the fp32 reference itself is several per cent from fp64 by now (measured: 5 %); the CUDA
        # path must be no further from fp64 than three times that
        for k in (0, 1):
            assert abs(la[k] - l64[k]) <= 3.0 * abs(lo[k] - l64[k]) + 1e-3 * abs(l64[k]), (k, la, lo, l64)
    finally:
        agent.ctx.set_gemm_mode(0)


def _torchrun(script, *args, nproc=2, timeout=600):
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr",
           "127.0.0.1", "--master-port", "29617", os.path.join(ROOT, script), *args]
    return subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (one process per GPU over NVLink)")
@pytest.mark.parametrize("script,args", [("tools/dp_peer_check.py", ("--updates", "12", "--batch", "1024", "--no-timing")),
                                         ("tools/dp_cql_check.py", ("--updates", "12", "--no-timing"))])
def test_data_parallel_on_real_gpus(script, args):
    """The data-parallel update on real hardware (2 processes, 2 GPUs): the tools assert that the replicas stay
    bitwise identical, that the in-kernel gradient sum equals the NCCL all-reduce path and that the data-parallel
    run reproduces the single-process update at the same global batch."""
    r = _torchrun(script, *args)
    assert r.returncode == 0, (r.stdout[-3000:], r.stderr[-3000:])
    assert "OK" in r.stdout


def _final(stdout):
    import ast
    import re
    return [ast.literal_eval(m) for m in re.findall(r"\{'updates'[^}]*\}", stdout)]     # ranks may share a line


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (one process per GPU over NVLink)")
@pytest.mark.parametrize("backbone", ["none", "fasternet"])
def test_sorl_train_cli_under_torchrun(backbone, tmp_path):
    """The documented WORLD_SIZE > 1 mode of the trainer itself: `torchrun --nproc-per-node 2
    sorl_train.py`.  Every rank must build its agent on its own GPU (process group and device first), finish, and report
    the same losses; without a backbone the run must also reproduce the single-process run at the same global batch
    (same permutation, each rank takes its slice of every minibatch, gradients summed in-kernel over NVLink)."""
    flags = ["--backbone", backbone, "--batch_size", "256", "--dataset_rows", "8192", "--updates", "24", "--log_every", "8",
             "--train_steps", "24", "--episodes", "1", "--weights_dir", str(tmp_path)]
    r = _torchrun("sorl_train.py", *flags)
    assert r.returncode == 0, (r.stdout[-3000:], r.stderr[-3000:])
    outs = _final(r.stdout)
    assert len(outs) == 2 and all(o["updates"] == 24 for o in outs), r.stdout[-2000:]
    assert all(np.isfinite([o["v_loss"], o["g_loss"]]).all() for o in outs)
    assert outs[0]["v_loss"] == outs[1]["v_loss"] and outs[0]["g_loss"] == outs[1]["g_loss"]     # replicas agree
    if backbone == "none":
        one = subprocess.run([sys.executable, os.path.join(ROOT, "sorl_train.py"), *flags], cwd=ROOT, capture_output=True,
                             text=True, timeout=600)
        assert one.returncode == 0, (one.stdout[-3000:], one.stderr[-3000:])
        single = _final(one.stdout)[0]
        for k in ("v_loss", "g_loss"):
            assert abs(outs[0][k] - single[k]) <= 1e-4 * abs(single[k]), (k, outs[0], single)
