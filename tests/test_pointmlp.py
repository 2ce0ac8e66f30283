"""PointMLP scan encoder (agent/feature_pcl.py): oracle vs the reference's fixtures (CPU), CUDA vs oracle (GPU).

Fixtures tests/golden/pointmlp_{a,b}.npz come from the UNMODIFIED reference PointMLPModel in train mode
(tests/golden/make_golden.py pointmlp): `out` is its fp32 forward, `out64` the same module in float64.
Tolerances (relative L2 of the (B, class_num) output):
  GOLD_TOL  1e-3   fp32 restatement vs the reference's fp32 output: the reference's own fp32 forward sits up to
                   1.5e-5 from its fp64 forward on these cases (tiny-batch BatchNorm amplifies rounding ~1e3 x)
  PIN_TOL   1e-9   fp64 restatement vs the reference in fp64 (pins the algorithm)
  FEAT_TOL  2e-4   CUDA path (3xTF32 contractions, fp64 BatchNorm sums) vs the fp64 yardstick
Indices (farthest-point sampling, kNN sets) are bit-exact.
"""
import ast
import os

import numpy as np
import pytest
import torch

from oracle import pointmlp as opm
from oracle import refimport

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD_TOL, PIN_TOL, FEAT_TOL = 1e-3, 1e-9, 2e-4


def rel(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def load(name):
    z = np.load(os.path.join(HERE, "golden", f"pointmlp_{name}.npz"), allow_pickle=False)
    kw = ast.literal_eval(str(z["kwargs"]))
    sd0 = {k[4:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("sd0/")}
    cfg = dict(points=kw["points"], k_neighbors=kw["k_neighbors"], reducers=kw.get("reducers", [2, 2, 2, 2]),
               use_xyz=kw.get("use_xyz", True), normalize=kw.get("normalize", "center"))
    return z, kw, sd0, cfg


def ordered(z, sd):
    return {k: sd[k] for k in [str(s) for s in z["keys"]]}


# ------------------------------------------------------------------------------------------ CPU: the oracle is pinned
@pytest.mark.parametrize("name", ["a", "b"])
def test_oracle_matches_reference_fixture(name):
    z, kw, sd0, cfg = load(name)
    starts = torch.from_numpy(z["fps_start"])
    drops = [torch.from_numpy(z["drop1"]), torch.from_numpy(z["drop2"])]
    rec = []
    sd = {k: v.clone() for k, v in sd0.items()}
    out = opm.pointmlp_forward(sd, torch.from_numpy(z["x"]), cfg, starts, drops, record=rec)
    assert rel(out, z["out"]) <= GOLD_TOL
    sd64 = {k: (v.double() if v.is_floating_point() else v.clone()) for k, v in sd0.items()}
    out64 = opm.pointmlp_forward(sd64, torch.from_numpy(z["x"]).double(), cfg, starts, [d.double() for d in drops])
    assert rel(out64, z["out64"]) <= PIN_TOL
    stages = [r for r in rec if "fps" in r]
    for i, r in enumerate(stages):
        assert np.array_equal(r["fps"].numpy().astype(np.int32), z[f"fps.{i}"])
        assert np.array_equal(np.sort(r["knn"].numpy().astype(np.int32), axis=-1), z[f"knn.{i}"])
    for k in z.files:
        if k.startswith("stats/") and "running" in k:
            np.testing.assert_allclose(sd[k[6:]].numpy(), z[k], rtol=2e-3, atol=1e-4)


@pytest.mark.parametrize("name", ["a", "b"])
def test_module_state_dict_layout_matches_reference(name):
    from porl_b200.agent.feature_pcl import PointMLPModel
    z, kw, sd0, _ = load(name)
    m = PointMLPModel(**kw)
    keys = list(m.state_dict().keys())
    assert keys == [str(s) for s in z["keys"]]
    for k, v in m.state_dict().items():
        assert tuple(v.shape) == tuple(sd0[k].shape), k
    m.load_state_dict(sd0)                                   # strict


@pytest.mark.skipif(not refimport.available(), reason="reference not mounted")
def test_module_init_equals_reference_init():
    """Same torch seed -> same initial weights as the reference class (same parameter creation order)."""
    from porl_b200.agent.feature_pcl import PointMLPModel
    rm = refimport.pointmlp_module()
    kw = dict(points=64, class_num=12, embed_dim=8, k_neighbors=[8, 8, 8, 8])
    torch.manual_seed(5)
    a = rm.PointMLPModel(**kw).state_dict()
    torch.manual_seed(5)
    b = PointMLPModel(**kw).state_dict()
    assert list(a) == list(b)
    assert all(torch.equal(a[k], b[k]) for k in a)


def test_module_rejects_what_is_not_built():
    from porl_b200.agent.feature_pcl import PointMLPModel
    with pytest.raises(NotImplementedError):
        PointMLPModel(points=64, embed_dim=8, groups=2)
    with pytest.raises(NotImplementedError):
        PointMLPModel(points=64, embed_dim=8, activation="gelu")
    m = PointMLPModel(points=16, embed_dim=8, k_neighbors=[4], reducers=[2], dim_expansion=[2], pre_blocks=[1], pos_blocks=[1])
    with pytest.raises(Exception, match="CUDA"):
        m(torch.zeros(2, 3, 16))


# ------------------------------------------------------------------------------------------ GPU
@pytest.fixture(scope="module")
def env():
    from porl_b200 import _C
    dev = torch.device("cuda:0")
    return _C, _C.context(dev), dev


def cloud(B, N, seed, scan=False):
    g = torch.Generator().manual_seed(seed)
    if not scan:
        return torch.rand(B, N, 3, generator=g) * 4 - 2
    from oracle.synth import laser_rows
    st = torch.from_numpy(laser_rows(B, 362, seed)[:, :362].copy())
    return opm.scan_to_points(st).permute(0, 2, 1).contiguous()


@pytest.mark.gpu
@pytest.mark.parametrize("B,N,npoint,scan", [(4, 64, 32, False), (3, 1024, 512, False), (5, 361, 180, True), (2, 37, 37, False)])
def test_fps_bit_exact(env, B, N, npoint, scan):
    _C, ctx, dev = env
    xyz = cloud(B, N, 100 + N, scan)
    start = torch.randint(0, N, (B,), generator=torch.Generator().manual_seed(N))
    want = opm.fps(xyz, npoint, start)
    idx = torch.zeros(B, npoint, dtype=torch.int32, device=dev)
    xyz_d, start_d = xyz.to(dev), start.to(dev)              # named: a temporary's block would be reused at once
    _C.check(_C.lib.porl_fps(ctx.handle, xyz_d.data_ptr(), B, N, npoint, start_d.data_ptr(), idx.data_ptr(),
                             _C.stream_ptr()))
    assert torch.equal(idx.cpu().long(), want)


@pytest.mark.gpu
@pytest.mark.parametrize("B,N,S,k", [(3, 64, 32, 8), (2, 1024, 512, 32), (4, 361, 180, 32), (2, 40, 7, 40)])
def test_knn_sets_match_topk(env, B, N, S, k):
    _C, ctx, dev = env
    xyz = cloud(B, N, 200 + N)
    q = xyz[:, torch.randperm(N, generator=torch.Generator().manual_seed(k))[:S]].contiguous()
    assert opm.knn_margin(k, xyz, q) > 1e-6                  # the neighbour set does not hinge on one ulp
    want = torch.sort(opm.knn(k, xyz, q), dim=-1)[0]
    idx = torch.zeros(B, S, k, dtype=torch.int32, device=dev)
    xyz_d, q_d = xyz.to(dev), q.to(dev)
    _C.check(_C.lib.porl_knn(ctx.handle, xyz_d.data_ptr(), q_d.data_ptr(), B, N, S, k, idx.data_ptr(),
                             _C.stream_ptr()))
    got = idx.cpu().long()
    assert torch.equal(torch.sort(got, dim=-1)[0], want)
    d = opm.square_distance(q, xyz).gather(-1, got)          # returned in ascending distance
    assert bool((d[..., 1:] >= d[..., :-1]).all())


@pytest.mark.gpu
def test_scan_to_points(env):
    from oracle.synth import laser_rows
    from porl_b200.agent.feature_pcl import scan_to_points
    _C, ctx, dev = env
    st = torch.from_numpy(laser_rows(16, 362, 9)[:, :362].copy())
    st[0, 5] = 9.0                                           # "no return" beam
    want = opm.scan_to_points(st.clone())
    got = scan_to_points(st.to(dev)).cpu()
    assert got.shape == (16, 3, 361)
    np.testing.assert_allclose(got.numpy(), want.numpy(), rtol=0, atol=2e-6)
    assert float(got[0, 0, 5]) == 0.0 and float(got[0, 2, 360]) == 1.0


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["a", "b"])
def test_forward_matches_reference_fixture(env, name):
    from porl_b200.agent.feature_pcl import PointMLPModel
    _C, ctx, dev = env
    z, kw, sd0, cfg = load(name)
    m = PointMLPModel(**kw)
    m.load_state_dict(sd0)
    m = m.to(dev).train()
    starts = torch.from_numpy(z["fps_start"])
    drops = [torch.from_numpy(z["drop1"]), torch.from_numpy(z["drop2"])]
    out = m(torch.from_numpy(z["x"]).to(dev), fps_start=starts, drops=drops)
    torch.cuda.synchronize()
    assert int(ctx.status_tensor()[0]) == 0
    err = rel(out, z["out64"])
    print(f"pointmlp_{name}: CUDA vs reference fp64 {err:.2e} (reference fp32 vs fp64 {rel(z['out'], z['out64']):.2e})")
    assert err <= FEAT_TOL
    sd = m.state_dict()
    for k in z.files:
        if k.startswith("stats/"):
            ref = z[k]
            if "tracked" in k:
                assert int(sd[k[6:]]) == int(ref)
            else:
                np.testing.assert_allclose(sd[k[6:]].cpu().numpy(), ref, rtol=2e-3, atol=1e-4)
    # a second forward reuses the workspace and keeps counting
    m(torch.from_numpy(z["x"]).to(dev), fps_start=starts, drops=drops)
    assert int(m.embedding.net[1].num_batches_tracked) == 2


@pytest.mark.gpu
def test_forward_scan_sized_model_vs_oracle(env):
    """PointMLPModel(points=361) on scan point sets at the reference's default widths (64..1024 channels, k = 32):
    CUDA vs the fp32 oracle on the same weights, start indices and dropout multipliers."""
    from oracle.synth import laser_rows
    from porl_b200.agent.feature_pcl import PointMLPModel, scan_to_points
    _C, ctx, dev = env
    B = 6
    torch.manual_seed(3)
    m = PointMLPModel(points=361, class_num=256)
    sd0 = opm.randomize_({k: v.clone() for k, v in m.state_dict().items()}, 4)
    m.load_state_dict(sd0)
    m = m.to(dev).train()
    g = torch.Generator().manual_seed(8)
    x = (torch.rand(B, 3, 361, generator=g) * 4 - 2).to(dev)  # generic positions: no exactly tied neighbour distances
    starts, drops = m.draw(B, dev)
    out = m(x, fps_start=starts, drops=drops)
    torch.cuda.synchronize()
    cfg = dict(points=361, k_neighbors=[32] * 4, reducers=[2] * 4, use_xyz=True, normalize="center")
    # the fp32 oracle is the yardstick: FPS / kNN decisions follow the reference's fp32 arithmetic
    rec = []
    sd32 = {k: v.clone() for k, v in sd0.items()}
    want = opm.pointmlp_forward(sd32, x.cpu(), cfg, starts.cpu(), [d.cpu() for d in drops], record=rec)
    err = rel(out, want)
    margin = min(r["margin"] for r in rec if "margin" in r)
    print(f"pointmlp 361-point model: CUDA vs fp32 oracle {err:.2e} (smallest kNN margin {margin:.1e})")
    assert margin > 1e-6
    assert err <= 1e-3
    for k, v in m.state_dict().items():
        if "running" in k:
            np.testing.assert_allclose(v.cpu().numpy(), sd32[k].numpy(), rtol=2e-3, atol=1e-4, err_msg=k)
    assert int(ctx.status_tensor()[0]) == 0
    st = torch.from_numpy(laser_rows(B, 362, 21)[:, :362].copy())
    feats = m(scan_to_points(st.to(dev)))      # scan point sets: ties between beams at equal range are broken by index
    assert feats.shape == (B, 256) and bool(torch.isfinite(feats).all())
