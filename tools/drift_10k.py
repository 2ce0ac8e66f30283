"""Free-running drift over 10,000 updates (BASELINE.json north_star: "drift bound stated over 10k steps").

    python tools/drift_10k.py [--kind sorl|por] [--updates 10000] [--batch 256] [--every 1000] [--out FILE]

Three runs from identical weights on identical minibatches, nothing re-synchronised in between:
  gpu   porl_b200 agent, tcgen05 3xTF32 contractions (what bench.py times),
  ref32 the oracle restatement of the reference update in fp32 on the CPU (agent/sorl.py:78-128, agent/por.py:73-112;
        pinned to the unmodified reference by tests/test_oracle_golden.py),
  ref64 the same in fp64: the yardstick.
Every `--every` updates it prints (and appends to --out as JSON lines) the relative L2 distance of all parameters to
ref64 for gpu and for ref32, and the relative loss differences.  The update is chaotic at the ulp level (a ReLU unit of
one sample landing on the other side moves the gradient by O(1/B), Adam normalises steps to +-lr), so no fp32
implementation tracks fp64 bit for bit; what can be stated and is asserted here is that the CUDA path drifts from fp64
no faster than the reference's own fp32 arithmetic does: d_gpu <= 3 * d_ref32 at every checkpoint.
The dims are the trainers' defaults: SORL S=362 H=512 (sorl_train.py:85,93), POR S=365 H=1024 (por_train.py:127,139).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import synth  # noqa: E402
from oracle.buffer import slice_row  # noqa: E402
from oracle.steps import OraclePOR, OracleSORL, dims  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--kind", default="sorl", choices=["sorl", "por"])
    ap.add_argument("--updates", type=int, default=10_000)
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--every", type=int, default=1000)
    ap.add_argument("--rows", type=int, default=65_536)
    ap.add_argument("--gemm-mode", type=int, default=1)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    from porl_b200 import POR, SORL
    kind = a.kind
    S, H, A = (362, 512, 2) if kind == "sorl" else (365, 1024, 2)
    args = dims(S, H, 2, False, 256, A)
    hyper = dict(max_steps=a.updates, tau=0.9, alpha=10.0)        # cosine schedule over the whole run
    dev = torch.device("cuda:0")
    agent = (POR if kind == "por" else SORL)(args, device=dev, **hyper)
    agent.ctx.set_gemm_mode(a.gemm_mode)
    ocls = OraclePOR if kind == "por" else OracleSORL
    o32, o64 = ocls(args, **hyper), ocls(args, **hyper)
    sd = synth.det_state_dict(o32, 9)
    for o in (o32, o64):
        o.load_state_dict(sd)
    o64.double()
    agent.load_state_dict(sd)
    rows = torch.from_numpy(synth.laser_rows(a.rows, S, 23, A))
    rows_dev = rows.to(dev)
    g = torch.Generator().manual_seed(2000)
    torch.set_num_threads(max(1, (os.cpu_count() or 2) // 2))
    pool = ThreadPoolExecutor(2)

    def step(agent_, b):
        if kind == "por":
            return agent_.por_residual_update(b["observations"], b["next_observations"], b["rewards"], b["terminals"])
        return agent_.update(b["observations"], b["actions"], b["rewards"], b["next_observations"], b["terminals"])

    flat = lambda m: torch.cat([v.double().flatten().cpu() for v in m.state_dict().values()])
    out = open(a.out, "a") if a.out else None
    t0, worst_ratio, perm, pos = time.time(), 0.0, None, 0
    for i in range(1, a.updates + 1):
        if perm is None or pos + a.batch > perm.numel():          # epoch permutations, like DataLoader(shuffle=True)
            perm, pos = torch.randperm(a.rows, generator=g), 0
        idx = perm[pos:pos + a.batch]
        pos += a.batch
        chunk = rows[idx]
        f32 = pool.submit(step, o32, slice_row(chunk.clone(), S, A))
        f64 = pool.submit(step, o64, slice_row(chunk.double(), S, A))
        lg = step(agent, slice_row(rows_dev[idx.to(dev)], S, A))
        l32, l64 = f32.result(), f64.result()
        if i % a.every == 0 or i == a.updates:
            ref = flat(o64)
            d_gpu = float((flat(agent) - ref).norm() / ref.norm())
            d_ref = float((flat(o32) - ref).norm() / ref.norm())
            rec = {"kind": kind, "S": S, "H": H, "batch": a.batch, "updates": i, "param_rel_l2_gpu_vs_fp64": d_gpu,
                   "param_rel_l2_ref32_vs_fp64": d_ref, "ratio": d_gpu / max(d_ref, 1e-30),
                   "v_loss": {"gpu": lg[0], "ref32": l32[0], "fp64": l64[0]},
                   "policy_loss": {"gpu": lg[1], "ref32": l32[1], "fp64": l64[1]},
                   "loss_rel_gpu_vs_fp64": [abs(lg[k] - l64[k]) / abs(l64[k]) for k in (0, 1)],
                   "loss_rel_ref32_vs_fp64": [abs(l32[k] - l64[k]) / abs(l64[k]) for k in (0, 1)],
                   "gemm_mode": a.gemm_mode, "seconds": round(time.time() - t0, 1)}
            worst_ratio = max(worst_ratio, rec["ratio"])
            line = json.dumps(rec)
            print(line, flush=True)
            if out:
                out.write(line + "\n")
                out.flush()
            assert d_gpu <= 3.0 * d_ref + 1e-6, f"update {i}: gpu drifts {d_gpu:.3e} from fp64, reference fp32 {d_ref:.3e}"
    print(f"OK drift_10k {kind}: {a.updates} free-running updates, worst d_gpu / d_ref32 = {worst_ratio:.2f}", flush=True)


if __name__ == "__main__":
    main()
