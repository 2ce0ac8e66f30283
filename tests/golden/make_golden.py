"""Generate tests/golden/*.npz by running the UNMODIFIED reference (imported from /root/reference).

Run in the build container only:   python tests/golden/make_golden.py
The reference cannot travel to the GPU box, so its outputs are committed as small fixtures.
Inputs are regenerated at test time from oracle.synth (seeded numpy RNG); the fixtures hold the
reference's OUTPUTS: per-step losses, final parameters (small cases) or per-tensor checksums
(real-dimension cases), sampled indices, LR sequences.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import refimport, synth  # noqa: E402
from oracle.buffer import slice_row  # noqa: E402
from oracle.steps import dims  # noqa: E402

torch.set_num_threads(1)  # fixtures must not depend on the thread count of the build box


def checksum(sd):
    out = {}
    for k, v in sd.items():
        v = v.detach().double().flatten()
        out[k] = np.array([v.sum().item(), v.abs().sum().item(), *v[:6].tolist()])
    return out


def flat(prefix, d):
    return {f"{prefix}/{k}": (v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in d.items()}


def run_agent(kind, S, H, B, steps, max_steps, seed, A=2, tau=0.9, alpha=10.0, full=True):
    args = dims(S, H, 2, False, 256, A)
    cls = refimport.por_cls() if kind == "por" else refimport.sorl_cls()
    agent = cls(args, max_steps=max_steps, tau=tau, alpha=alpha)
    agent.load_state_dict(synth.det_state_dict(agent, seed))
    rows = torch.from_numpy(synth.laser_rows(B * steps, S, seed + 1, A))
    losses = []
    for i in range(steps):
        b = slice_row(rows[i * B:(i + 1) * B], S, A)
        if kind == "por":
            losses.append(agent.por_residual_update(b["observations"], b["next_observations"], b["rewards"], b["terminals"]))
        else:
            losses.append(agent.update(b["observations"], b["actions"], b["rewards"], b["next_observations"], b["terminals"]))
    sched = agent.goal_lr_schedule if kind == "por" else agent.lr_schedule
    out = dict(meta=np.array([S, H, B, steps, max_steps, seed, A]), hyper=np.array([tau, alpha]),
               losses=np.array(losses, np.float64), last_lr=np.array(sched.get_last_lr(), np.float64))
    out.update(flat("final", agent.state_dict()) if full else flat("sum", checksum(agent.state_dict())))
    return out


def run_cql(S, A, B, n_rows, iters, freq, seed):
    t = refimport.make_cql_trainer(S, A, update_target_freq=freq)
    t.q_network.load_state_dict(synth.det_state_dict(t.q_network, seed))
    t.target_network.load_state_dict(t.q_network.state_dict())
    t.batch_size = B
    rows = synth.laser_rows(n_rows, S, seed + 1)
    acts = synth.discrete_actions(n_rows, A, seed + 2)
    b = slice_row(rows, S)
    for i in range(n_rows):
        t.replay_buffer.push(b["observations"][i], acts[i], b["rewards"][i], b["next_observations"][i], bool(b["terminals"][i]))
    np.random.seed(seed + 3)
    state = np.random.get_state()
    idx = [np.random.choice(n_rows, B, replace=False) for _ in range(iters)]
    np.random.set_state(state)
    losses = t.train_offline(num_iterations=iters)
    out = dict(meta=np.array([S, A, B, n_rows, iters, freq, seed]), losses=np.array(losses, np.float64), indices=np.stack(idx))
    out.update(flat("q", t.q_network.state_dict()))
    out.update(flat("target", t.target_network.state_dict()))
    return out


def run_buffer(seed):
    RB = refimport.replay_buffer_cls()
    rb = RB(50, (7,), torch.device("cpu"))
    rng = np.random.RandomState(seed)
    for i in range(70):  # wraps the ring
        rb.push(rng.rand(7).astype(np.float32), int(rng.randint(5)), float(rng.randn()), rng.rand(7).astype(np.float32), bool(rng.rand() < 0.2))
    np.random.seed(seed)
    s, a, r, s2, d = rb.sample(16)
    return dict(seed=np.array(seed), states=s.numpy(), actions=a.numpy(), rewards=r.numpy(), next_states=s2.numpy(), dones=d.numpy(),
                size=np.array(len(rb)), position=np.array(rb.position))


def run_cosine(base, t_max, n):
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.Adam([p], lr=base)
    sch = torch.optim.lr_scheduler.CosineAnnealingLR(opt, t_max)
    lrs = []
    for _ in range(n):
        lrs.append(opt.param_groups[0]["lr"])
        opt.step()
        sch.step()
    return dict(base=np.array(base), t_max=np.array(t_max), lrs=np.array(lrs, np.float64))


def run_encoder(seed, B):
    """The unmodified reference FasterNet(3, 256) on deterministic weights; DropPath masks injected by replacing
    the module-level drop_path function (agent/fasternet.py:76-93) with one that replays recorded factors."""
    from oracle import encoder
    FN = refimport.fasternet_cls()
    import agent.fasternet as rf
    st, keeps = encoder.encoder_inputs(seed, B)
    cm = refimport.costmap_fn()(st.clone())
    m = FN(3, 256)
    keys = list(m.state_dict().keys())
    m.load_state_dict(encoder.det_encoder_state_dict(seed + 1))
    queue = [k for k in keeps if k is not None]
    orig = rf.drop_path

    def replay(x, drop_prob=0., training=False, scale_by_keep=True):
        if drop_prob == 0. or not training:
            return x
        return x * queue.pop(0).view(-1, 1, 1, 1).to(x.dtype)

    rf.drop_path = replay
    try:
        x = st.clone()
        feats = [m(x).detach().numpy()]
    finally:
        rf.drop_path = orig
    sd = m.state_dict()
    out = dict(meta=np.array([seed, B]), keys=np.array(keys), features=feats[0], clamped_state=x.numpy(),
               costmap_nonzero=np.argwhere(cm.numpy() != 0).astype(np.int32))
    out.update(flat("stats", {k: v for k, v in sd.items() if "running" in k or "tracked" in k}))
    torch.manual_seed(seed)
    init = FN(3, 256).state_dict()
    out.update(flat("init", {k: np.array([v.double().sum().item(), v.double().abs().sum().item()])
                             for k, v in init.items() if v.is_floating_point()}))
    return out


PM_CASES = {   # name: (seed, B, model kwargs)
    "a": (61, 8, dict(points=64, class_num=10, embed_dim=8, k_neighbors=[8, 8, 8, 8])),                # defaults: center, xyz
    "b": (62, 6, dict(points=48, class_num=7, embed_dim=8, k_neighbors=[6, 6], reducers=[2, 2], dim_expansion=[2, 2],
                      pre_blocks=[1, 2], pos_blocks=[2, 1], normalize="anchor", use_xyz=False, bias=False,
                      res_expansion=0.5)),
}


def pointmlp_inputs(seed, B, points):
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(B, 3, points, generator=g) * 4 - 2
    x[:, 2] = (torch.rand(B, points, generator=g) < 0.1).float()      # a flag-like third coordinate
    return x


def run_pointmlp(name):
    """The unmodified reference PointMLPModel in train mode.  Its two random draws are replayed from recorded
    values: torch.randint (FPS start, feature_pcl.py:81) and the classifier's Dropout masks."""
    from oracle import pointmlp
    seed, B, kw = PM_CASES[name]
    rm = refimport.pointmlp_module()
    torch.manual_seed(seed)
    m = rm.PointMLPModel(**kw)
    keys = list(m.state_dict().keys())
    m.load_state_dict(pointmlp.randomize_(m.state_dict(), seed + 1))
    sd0 = {k: v.clone() for k, v in m.state_dict().items()}
    x = pointmlp_inputs(seed, B, kw["points"])
    n_stages = len(kw["k_neighbors"])
    g = torch.Generator().manual_seed(seed + 2)
    anchors, starts = kw["points"], []
    for red in kw.get("reducers", [2, 2, 2, 2]):
        starts.append(torch.randint(0, anchors, (B,), generator=g))
        anchors //= red
    starts = torch.stack(starts)
    drops = [torch.bernoulli(torch.full((B, w), 0.5), generator=g) / 0.5 for w in (512, 256)]
    queue, dq = list(starts), list(drops)
    orig_randint, orig_dropout = torch.randint, torch.nn.functional.dropout

    def randint(lo, hi, size, **k):
        nonlocal queue, dq
        v = queue.pop(0)
        assert tuple(size) == tuple(v.shape) and int(v.max()) < hi
        return v.clone()

    def dropout(inp, p=0.5, training=True, inplace=False):
        nonlocal queue, dq
        assert training and p == 0.5
        return inp * dq.pop(0)

    m64 = rm.PointMLPModel(**kw).double()
    m64.load_state_dict({k: (v.double() if v.is_floating_point() else v) for k, v in sd0.items()})
    torch.randint, torch.nn.functional.dropout = randint, dropout
    try:
        m.train()
        out = m(x.clone()).detach()
        queue, dq = list(starts), [d.double() for d in drops]
        out64 = m64.train()(x.double()).detach()
    finally:
        torch.randint, torch.nn.functional.dropout = orig_randint, orig_dropout
    assert not queue and not dq
    cfg = dict(points=kw["points"], k_neighbors=kw["k_neighbors"], reducers=kw.get("reducers", [2, 2, 2, 2]),
               use_xyz=kw.get("use_xyz", True), normalize=kw.get("normalize", "center"))
    rec = []
    sd = {k: v.clone() for k, v in sd0.items()}
    mine = pointmlp.pointmlp_forward(sd, x.clone(), cfg, starts, drops, record=rec)
    margin = min(r["margin"] for r in rec if "margin" in r)
    err = float((mine - out).norm() / out.norm())
    sd64 = {k: (v.double() if v.is_floating_point() else v.clone()) for k, v in sd0.items()}
    mine64 = pointmlp.pointmlp_forward(sd64, x.double(), cfg, starts, [d.double() for d in drops])
    err64 = float((mine64 - out64).norm() / out64.norm())
    gap = float((out.double() - out64).norm() / out64.norm())
    print(f"pointmlp_{name}: oracle vs reference rel-L2 fp32 {err:.2e}, fp64 {err64:.2e}; reference fp32 vs fp64 {gap:.2e}; "
          f"smallest kNN margin {margin:.2e}")
    assert err < 1e-3 and err64 < 1e-10 and margin > 1e-4, (err, err64, margin)
    ref_sd = m.state_dict()
    for k in ref_sd:
        if "running" in k:
            assert torch.allclose(ref_sd[k], sd[k], rtol=2e-3, atol=1e-4), k
    d = dict(meta=np.array([seed, B]), keys=np.array(keys), x=x.numpy(), fps_start=starts.numpy(),
             drop1=drops[0].numpy(), drop2=drops[1].numpy(), out=out.numpy(), out64=out64.numpy(), kwargs=np.array(repr(kw)))
    d.update(flat("sd0", sd0))
    d.update(flat("stats", {k: v for k, v in ref_sd.items() if "running" in k or "tracked" in k}))
    for i, r in enumerate(x for x in rec if "fps" in x):
        d[f"fps.{i}"] = r["fps"].numpy().astype(np.int32)
        d[f"knn.{i}"] = np.sort(r["knn"].numpy().astype(np.int32), axis=-1)
    return d


def main():
    assert refimport.available(), "needs /root/reference"
    if len(sys.argv) > 1 and sys.argv[1] == "pointmlp":
        for name in PM_CASES:
            path = os.path.join(HERE, f"pointmlp_{name}.npz")
            np.savez_compressed(path, **run_pointmlp(name))
            print(f"pointmlp_{name}: {os.path.getsize(path) / 1024:.1f} KiB")
        return
    if len(sys.argv) > 1 and sys.argv[1] == "encoder":
        d = run_encoder(51, 6)
        path = os.path.join(HERE, "encoder_small.npz")
        np.savez_compressed(path, **d)
        print(f"encoder_small: {os.path.getsize(path) / 1024:.1f} KiB")
        return
    cases = {
        "por_small": run_agent("por", 22, 32, 16, 5, 3, 11),
        "sorl_small": run_agent("sorl", 22, 32, 16, 5, 3, 12),
        "sorl_small_tau5": run_agent("sorl", 22, 32, 16, 3, 1000, 13, tau=0.5, alpha=3.0),
        "por_real": run_agent("por", 365, 1024, 64, 2, 1000, 21, full=False),
        "sorl_real": run_agent("sorl", 362, 512, 64, 2, 1000, 22, full=False),
        "cql_small": run_cql(22, 5, 16, 200, 5, 2, 31),
        "cql_real": run_cql(362, 5, 64, 500, 3, 10, 32),
        "replay_buffer": run_buffer(41),
        "cosine_lr": run_cosine(1e-4, 4, 14),
        "encoder_small": run_encoder(51, 6),
    }
    for name, d in cases.items():
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **d)
        print(f"{name}: {os.path.getsize(path) / 1024:.1f} KiB")


if __name__ == "__main__":
    main()
