"""Device index sampler (porl_sample_permutation) and dataset ingest.

The sampler has no reference code to match index for index (the reference shuffles on the host with
torch.randperm / np.random.choice); its contract is checked as properties -- bijection of [0, n), distinct rows in
a batch, one visit per row and epoch, a different order per key -- and the kernel is pinned bit for bit to the
CPU restatement oracle/sampler.py.  The ingest loaders are checked against the reference's own read path
(np.loadtxt(...).reshape(100, -1)[row] of dataloader/dataloader_bk.py:18-32)."""
import os

import numpy as np
import pytest
import torch

from oracle import sampler as osampler, synth


@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 17, 256, 1000, 65537])
def test_oracle_permutation_is_a_bijection(n):
    p = osampler.permutation(n, 11, 3, 0, n)
    assert np.array_equal(np.sort(p), np.arange(n))
    q = osampler.permutation(n, 11, 4, 0, n)
    assert n < 17 or not np.array_equal(p, q)                        # another key, another order
    w = osampler.permutation(n, 11, 3, n - 1, 3)                     # windows wrap modulo n
    assert w[0] == p[n - 1] and w[1] == p[0]


def test_oracle_permutation_looks_uniform():
    n, trials = 1000, 400
    first = np.array([osampler.permutation(n, s, 0, 0, 1)[0] for s in range(trials)])
    hist = np.bincount(first // 100, minlength=10)                   # 10 bins, expectation 40 each
    assert hist.min() >= 15 and hist.max() <= 70
    # position of row 0 over keys is spread as well
    pos = np.array([int(np.nonzero(osampler.permutation(n, s, 1, 0, n) == 0)[0][0]) for s in range(100)])
    assert pos.std() > 200


def test_ingest_csv_and_dict_shards_match_reference_read_path(tmp_path):
    from porl_b200.dataloader import ingest
    S, A = 362, 2
    rows = synth.laser_rows(250, S, 3, A)
    for k in range(2):                                               # two full 100-row shards (the reference's unit)
        np.savetxt(tmp_path / f"dataset_{k}.csv", rows[k * 100:(k + 1) * 100].reshape(-1)[None, :], delimiter=",")
    got = ingest.load_csv_shards(str(tmp_path), S, A)
    assert got.shape == (200, 2 * S + 2 + A) and got.dtype == np.float32
    for index in (0, 57, 100, 199):                                  # dataloader_bk.py:14-32
        ref = np.loadtxt(tmp_path / f"dataset_{index // 100}.csv", delimiter=",", dtype=float).reshape(100, -1)[index % 100]
        assert np.array_equal(got[index], ref.astype(np.float32))
    assert np.array_equal(ingest.load_directory(str(tmp_path), S, A), got)

    d = tmp_path / "dicts"
    d.mkdir()
    for k in range(2):                                               # collect.py:46-56
        shard = [{"state": rows[i, :S], "next_state": rows[i, S + 1:2 * S + 1], "reward": float(rows[i, S]),
                  "done": bool(rows[i, 2 * S + 1]), "truncated": False} for i in range(k * 100, (k + 1) * 100)]
        torch.save(shard, d / f"dataset_0_{k}.pt")
    got2 = ingest.load_dict_shards(str(d), S, A)
    assert got2.shape == (200, 2 * S + 2 + A)
    assert np.array_equal(got2[:, :2 * S + 2], rows[:200, :2 * S + 2]) and not got2[:, -A:].any()

    a = tmp_path / "a_star"
    a.mkdir()
    lab = torch.rand(200, 361)
    torch.save(lab[:100].clone(), a / "dataset_0.pt")
    torch.save(lab[100:].clone(), a / "dataset_1.pt")
    assert torch.equal(ingest.load_astar_shards(str(a)), lab)


@pytest.mark.gpu
@pytest.mark.parametrize("n,first,count", [(1, 0, 5), (2, 1, 4), (5, 0, 5), (1000, 990, 20), (1_000_003, 0, 4096),
                                           (1_000_003, 999_000, 2000), (10_000_000, 123, 65536), ((1 << 33) + 5, 7, 1000)])
def test_cuda_permutation_matches_oracle(n, first, count):
    from porl_b200 import _C
    dev = torch.device("cuda:0")
    ctx = _C.context(dev)
    out = torch.empty(count, dtype=torch.int64, device=dev)
    for seed, stream in ((0, 0), (2024, 17), ((1 << 63) + 5, (1 << 40) + 1)):
        _C.check(_C.lib.porl_sample_permutation(ctx.handle, n, seed, stream, first, count, out.data_ptr(), _C.stream_ptr()))
        assert np.array_equal(out.cpu().numpy(), osampler.permutation(n, seed, stream, first, count))


@pytest.mark.gpu
def test_cuda_device_sampler_epoch_contract():
    from porl_b200.dataloader.dataset import DeviceSampler
    n, bs = 1_000_003, 4096
    s = DeviceSampler(n, bs, "cuda:0", seed=5)
    batches = list(s)
    assert len(batches) == len(s) == (n + bs - 1) // bs and batches[-1].numel() == n % bs
    allidx = torch.cat(batches)
    assert torch.equal(torch.sort(allidx).values, torch.arange(n, device="cuda:0"))   # every row exactly once
    second = torch.cat(list(s))
    assert not torch.equal(allidx, second) and torch.equal(torch.sort(second).values, torch.arange(n, device="cuda:0"))
    d = s.draw(256)
    assert d.unique().numel() == 256 and int(d.min()) >= 0 and int(d.max()) < n
    assert DeviceSampler(10, 4, "cuda:0", drop_last=True).__len__() == 2
