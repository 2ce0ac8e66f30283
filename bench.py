#!/usr/bin/env python
"""Headline benchmark: offline-RL gradient updates/sec of the porl update step on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[1]): SORL, MLP-only heads (S=362, H=512, A=2; sorl_train_v0.py's
variant of agent/sorl.py), synthetic 1M-transition LaserScan+goal dataset resident in HBM, batch
4096 per GPU.  One "step" = one gradient update = minibatch gather (K1) + value step + Polyak +
policy step (K2..K7).  N > 1: data-parallel, one process per GPU, per-GPU batch fixed (weak
scaling), one NCCL all-reduce per optimizer group; `value` counts 4096-sample minibatches/s over
all ranks.  Prints ONE JSON line (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

S, H, A, N_HIDDEN = 362, 512, 2, 2
METRIC = "offline-RL gradient updates/sec"
UNIT = "updates/s"


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=200)
    p.add_argument("--warmup", type=int, default=10)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--batch", type=int, default=4096)
    p.add_argument("--rows", type=int, default=1_000_000)
    p.add_argument("--gemm-mode", type=int, default=int(os.environ.get("PORL_GEMM_MODE", "1")),
                   help="0 fp32 FFMA contractions, 1 tcgen05 3xTF32 where shapes allow")
    p.add_argument("--graph", type=int, default=1, help="replay the update as a CUDA graph")
    p.add_argument("--cpu-seconds", type=float, default=15.0, help="budget of the cpu_baseline leg")
    p.add_argument("--e2e-steps", type=int, default=0, help="0 = same as --steps")
    p.add_argument("--also-batch", type=int, default=256, help="second batch size reported under 'also'")
    p.add_argument("--other-configs", type=int, default=1, help="also time BASELINE configs 1 (POR) and 3 (CQL)")
    p.add_argument("--pointmlp-batch", type=int, default=128,
                   help="batch of the PointMLP (agent/feature_pcl.py) forward reported under 'also_pointmlp' (0: skip)")
    p.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                   help="N > 1: weak = --batch rows per GPU (global batch grows with N); strong = --batch rows in total")
    p.add_argument("--cfg4-batch", type=int, default=-1,
                   help="N > 1: per-GPU batch of the POR + FasterNet leg (BASELINE config 4: 65536 over 8 GPUs); "
                        "-1 = 8192 when N == 8, else skipped; 0 = skip")
    p.add_argument("--ceilings", type=int, default=1, help="measure the TF32 ceilings (cuBLAS TF32, own kernel 1-pass)")
    p.add_argument("--encoder-batch", type=int, default=256,
                   help="batch of the SORL + FasterNet scan-encoder workload reported under 'also_encoder' (0: skip)")
    return p.parse_args()


def workload_name(batch):
    return f"sorl_mlp_S{S}_H{H}_rows1M_b{batch}"


def hyper():
    from types import SimpleNamespace
    args = SimpleNamespace(state_size=S, hidden_dim=H, n_hidden=N_HIDDEN, layer_norm=False, feature_dim=256, action_size=A)
    return args, dict(max_steps=1_000_000, tau=0.9, alpha=10.0)


def synth_rows_torch(n, device, seed, state_size=None):
    """Synthetic LaserScan+goal transitions generated with torch on `device`; state_size > 362 puts
    pose-like filler columns U(-3, 3) between the scan and the goal (the S = 365 rows of por_train.py)."""
    state_size = state_size or S
    g = torch.Generator(device=device).manual_seed(seed)
    r = lambda *s: torch.rand(*s, generator=g, device=device)
    scan = 0.13 + (3.5 - 0.13) * r(n, 360)
    u = r(n, 360)
    scan = torch.where(u < 0.25, torch.full_like(scan, 3.5), scan)
    scan = torch.where(u > 0.97, torch.full_like(scan, 10.0), scan)
    rad, th = 0.2 + 3.3 * r(n), (2 * r(n) - 1) * 3.14159265
    goal = torch.stack([rad * torch.cos(th), rad * torch.sin(th)], 1)
    if state_size > 362:
        goal = torch.cat([6 * r(n, state_size - 362) - 3, goal], 1)
    nscan = (scan + 0.05 * torch.randn(n, 360, generator=g, device=device)).clamp(0.13, 3.5)
    nscan = torch.where(scan == 3.5, scan, nscan)
    nscan = torch.where(scan == 10.0, scan, nscan)
    ngoal = goal + 0.05 * torch.randn(n, goal.shape[1], generator=g, device=device)
    done = (r(n) < 0.01).float()
    sign = torch.where(r(n) < 0.5, -5.0, 5.0)
    rew = 0.1 * torch.randn(n, generator=g, device=device) + done * sign
    act = 2 * r(n, A) - 1
    return torch.cat([scan, goal, rew[:, None], nscan, ngoal, done[:, None], act], 1).float()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for nm, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                continue
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        return {"hbm_gbs": d["hbm_gbs"], "bf16_burst": d["bf16_tflops"], "bf16_sustained": d["bf16_tflops_sustained"],
                "source": "measured"}
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "source": "fallback"}


def measure_ceilings(ctx, device):
    """Measured denominators of the 3xTF32 contraction ("the bench measures them the same way"):
    cuBLAS TF32 (torch.matmul with allow_tf32, 8192^3, best of 5 -- a library call used ONLY as the yardstick) and this
    repo's tcgen05 kernel issuing one TF32 pass instead of three (porl_ctx_set_tc_terms(1)) on the same problem.  The
    3xTF32 mode issues three MMAs per product: its ceiling is a third of the single-pass TF32 rate."""
    import ctypes as C
    from porl_b200 import _C
    n = 8192
    g = torch.Generator(device=device).manual_seed(5)
    x = torch.randn(n, n, generator=g, device=device)
    w = torch.randn(n, n, generator=g, device=device)
    out = torch.empty(n, n, device=device)
    flop = 2.0 * n * n * n

    def best(fn, reps=5):
        fn()
        torch.cuda.synchronize()
        t = []
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            t.append(e0.elapsed_time(e1))
        return flop / (min(t) * 1e-3) * 1e-12

    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        cublas_tf32 = best(lambda: torch.matmul(x, w.t(), out=out))
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old
    cublas_fp32 = best(lambda: torch.matmul(x, w.t(), out=out), reps=2)

    def own():
        _C.check(_C.lib.porl_gemm(ctx.handle, 1, 1, n, n, n, x.data_ptr(), n, w.data_ptr(), n, out.data_ptr(), n, None, 0,
                                  None, 0, n, None, None, 0, _C.stream_ptr()))
    mode = ctx.gemm_mode
    ctx.set_gemm_mode(1)
    try:
        own3 = best(own)
        ctx.set_tc_terms(1)
        own1 = best(own)
    finally:
        ctx.set_tc_terms(3)
        ctx.set_gemm_mode(mode)
    del x, w, out
    form = mma_form_rates()
    return {"mma_form": form, "shape": "8192^3 fp32 operands, best of 5, CUDA events", "cublas_tf32_tflops": cublas_tf32,
            "cublas_fp32_tflops": cublas_fp32, "own_kernel_1xtf32_tflops": own1, "own_kernel_3xtf32_tflops": own3,
            "ceiling_3xtf32_tflops": cublas_tf32 / 3.0,
            "note": "3xTF32 issues 3 MMAs per product: its ceiling is measured cuBLAS TF32 / 3"}


def mma_form_rates():
    """Issue rate of the contraction kernel's own MMA form, measured in this run by the stand-alone probe built beside the
    library (tools/micro/umma_rate.cu --json: 12 tcgen05.mma.kind::tf32 M=128 K=8 + one commit per k-block, one issuing
    warp per SM, 148 SMs).  form_ceiling_3xtf32_tflops = what a 128x128-tile SS kernel can reach with three MMAs per
    product at the maximum SM clock; the cuBLAS-derived ceiling above it is that of N=256 / cta_group::2 tiles."""
    exe = os.path.join(os.path.dirname(os.path.abspath(__file__)), "tools", "micro", "umma_rate")
    if not os.path.exists(exe):
        return {"unavailable": "tools/micro/umma_rate not built"}
    try:
        r = subprocess.run([exe, "--json"], capture_output=True, text=True, timeout=60)
        d = json.loads(r.stdout.strip().splitlines()[-1])
    except Exception as e:  # noqa: BLE001 -- a probe failure must not lose the bench line
        return {"unavailable": f"{type(e).__name__}: {e}"}
    try:
        mhz = float(subprocess.run(["nvidia-smi", "--query-gpu=clocks.max.sm", "--format=csv,noheader,nounits", "-i", "0"],
                                   capture_output=True, text=True, timeout=20).stdout.split()[0])
    except Exception:  # noqa: BLE001
        mhz = 1965.0
    for k in ("ss_n128", "ss_n256", "ts_n128", "ts_n256"):
        d[f"{k}_3xtf32_tflops"] = d[f"{k}_fma_per_clk_sm"] * 2.0 * 148 * mhz * 1e6 / 3.0 * 1e-12
    d["sm_mhz"] = mhz
    d["form_ceiling_3xtf32_tflops"] = d["ss_n128_3xtf32_tflops"]
    d["how"] = "tools/micro/umma_rate --json; kernel form = ss_n128 (both operands in shared memory, N = 128)"
    return d


def traffic_record(gemm_mode):
    """dram bytes per launch of gemm_tc_kernel from the committed `ncu --set full` capture, only if that capture was
    taken from the kernel source as it is now (sha256 of csrc/gemm_tc.cu stored beside it); a stale capture reads null."""
    import hashlib
    if gemm_mode != 1:
        return None, "no capture for the FFMA mode"
    try:
        with open(os.path.join(ROOT, "porl_b200", "csrc", "gemm_tc.cu"), "rb") as f:
            sha = hashlib.sha256(f.read()).hexdigest()
        best = None
        for name in sorted(os.listdir(os.path.join(ROOT, "profiles"))):
            if name.endswith("_gemm_tc_traffic.json"):
                with open(os.path.join(ROOT, "profiles", name)) as f:
                    d = json.load(f)
                if d.get("source_sha256") == sha:
                    best = (d["mean_bytes_per_launch"], f"profiles/{name} (ncu --set full, source hash matches)")
        return best if best else (None, "no ncu capture of the current gemm_tc.cu (source hash differs): traffic not claimed")
    except Exception as e:
        return None, f"traffic record unreadable: {e}"


# ------------------------------------------------------------------------------------------ CPU arm
def cpu_reference_updates(batch, steps, warmup, seconds=None):
    """The reference's CPU path for this workload: the oracle restatement of agent/sorl.py (eager
    torch fp32, all host threads), on a bounded sample of the same synthetic data.  Returns
    (updates/s, n_timed, threads)."""
    from oracle.buffer import slice_row
    from oracle.steps import OracleSORL
    args, hp = hyper()
    torch.manual_seed(1000)
    agent = OracleSORL(args, **hp)
    n_distinct = min(max(steps, 1), 8)
    rows = synth_rows_torch(batch * n_distinct, torch.device("cpu"), 20260102)
    def one(i):
        b = slice_row(rows[(i % n_distinct) * batch:((i % n_distinct) + 1) * batch], S, A)
        agent.update(b["observations"], b["actions"], b["rewards"], b["next_observations"], b["terminals"])
    for i in range(warmup):
        one(i)
    t0 = time.perf_counter()
    n = 0
    while n < steps:
        one(n)
        n += 1
        if seconds is not None and time.perf_counter() - t0 > seconds:
            break
    dt = time.perf_counter() - t0
    return n / dt, n, torch.get_num_threads(), dt


def cpu_reference_encoder(batch, seconds):
    """CPU path of the SORL + FasterNet agent (config 2b): oracle restatement (agent/sorl.py:78-128 over
    oracle.encoder's fasternet_forward), eager torch fp32 on all host threads, a bounded number of updates at a
    small batch (B=32 on the host).  Forward-only encoder, i.e. LESS work than the reference,
    which also back-propagates through it.  Returns (updates/s, samples/s, n)."""
    from oracle import encoder
    from oracle.buffer import slice_row
    from oracle.steps import OracleSORL
    args, hp = hyper()
    torch.manual_seed(1000)
    sd = encoder.det_encoder_state_dict(1)

    class Backbone:
        def __call__(self, x):
            with torch.no_grad():
                return encoder.fasternet_forward(sd, x, None)

    agent = OracleSORL(args, backbone=Backbone(), **hp)
    rows = synth_rows_torch(batch * 2, torch.device("cpu"), 20260102)
    def one(i):
        b = slice_row(rows[(i % 2) * batch:((i % 2) + 1) * batch].clone(), S, A)
        agent.update(b["observations"], b["actions"], b["rewards"], b["next_observations"], b["terminals"])
    one(0)
    t0, n = time.perf_counter(), 0
    while n < 1 or time.perf_counter() - t0 < seconds:
        one(n)
        n += 1
    dt = time.perf_counter() - t0
    return n / dt, n * batch / dt, n


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    torch.set_num_threads(os.cpu_count() or 1)
    ups, n, threads, dt = cpu_reference_updates(a.batch, a.steps, max(a.warmup, 1))
    sample = f"{n} updates of batch {a.batch} (oracle restatement of agent/sorl.py, eager torch fp32 CPU)"
    line = {"impl": "reference", "metric": METRIC, "value": ups, "unit": UNIT, "n_gpus": a.gpus, "steps": n,
            "warmup": max(a.warmup, 1), "ms_per_step": 1e3 * dt / n, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(a.batch), "algo": "SORL", "batch": a.batch, "state_size": S,
                       "hidden_dim": H, "device": "cpu"},
            "cpu_baseline": {"value": ups, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": ups, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------ B200 arm
class Runner:
    """Dataset replica + agent + (optionally graph-captured) step for one batch size on this rank."""

    def __init__(self, a, batch, device, rank, world, table, backbone=False):
        from porl_b200 import SORL, TransitionDataset
        args, hp = hyper()
        self.a, self.B, self.device, self.rank, self.world = a, batch, device, rank, world
        self.ds = table
        torch.manual_seed(1000)                      # identical initial weights on every rank
        if backbone:                                 # the literal sorl_train.py:29-32 agent (BASELINE config 2b)
            from porl_b200.agent.fasternet import FasterNet
            hp = dict(hp, backbone=FasterNet(3, args.feature_dim))
        self.agent = SORL(args, device=device, **hp)
        self.agent.ctx.set_gemm_mode(a.gemm_mode)
        self.pb = self.ds.new_batch(batch)
        self.idx_static = torch.zeros(batch, dtype=torch.int64, device=device)
        self.graph = None

    def indices_global(self, n_steps, seed):
        """(n_steps, world * B): the global minibatches; rank r's share is columns [r * B, (r + 1) * B)."""
        g = torch.Generator(device=self.device).manual_seed(seed)
        total = n_steps * self.B * self.world
        chunks = []
        while total > 0:
            perm = torch.randperm(len(self.ds), generator=g, device=self.device)
            chunks.append(perm[:total])
            total -= chunks[-1].numel()
        return torch.cat(chunks).view(n_steps, self.world * self.B)

    def indices(self, n_steps, seed):
        g = torch.Generator(device=self.device).manual_seed(seed)
        total = n_steps * self.B * self.world
        chunks = []
        while total > 0:                              # epoch permutations (without replacement)
            perm = torch.randperm(len(self.ds), generator=g, device=self.device)
            chunks.append(perm[:total])
            total -= chunks[-1].numel()
        flat = torch.cat(chunks).view(n_steps, self.world, self.B)
        return flat[:, self.rank].contiguous()

    def step(self, idx):
        self.ds.gather_packed(idx, self.pb)
        self.agent.update_packed(self.pb)

    def capture(self):
        """CUDA graph of gather + update on a static index buffer (single process, or DP with the gradient sum
        fused into the Adam kernel; the NCCL all-reduce fallback stays outside graphs here)."""
        self.step(self.idx_static)                    # allocate workspaces outside capture
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        l0 = self.agent.ctx.launches
        with torch.cuda.graph(g):
            self.step(self.idx_static)
        self.launches_per_step = self.agent.ctx.launches - l0    # kernels recorded in the graph = launched per replay
        self.graph = g

    def run(self, idx_all):
        if self.graph is None:
            for i in range(idx_all.shape[0]):
                self.step(idx_all[i])
        else:
            for i in range(idx_all.shape[0]):
                self.idx_static.copy_(idx_all[i])
                self.graph.replay()


def timed(runner, idx_warm, idx_timed, dist_on):
    import torch.distributed as dist
    runner.run(idx_warm)
    if dist_on:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = runner.agent.ctx.launches
    e0.record()
    runner.run(idx_timed)
    e1.record()
    torch.cuda.synchronize()
    if dist_on:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    launches = runner.agent.ctx.launches - l0
    if dist_on:
        t = torch.tensor([ms], device=runner.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
    return ms, launches


def kernel_table(prof, pk):
    """Per-category achieved rate of the launches recorded by porl_profile_*: flop categories against the measured
    bf16 peak (and the fp32 SIMT pipe for the FFMA conv), byte categories against the measured HBM bandwidth."""
    out = {}
    flop_cats = {"contraction": pk["bf16_sustained"] * 1e12, "pconv": 74.0e12}
    for name, v in prof.items():
        if v["launches"] == 0:
            continue
        rate = v["work"] / (v["ms"] * 1e-3) if v["ms"] > 0 else 0.0
        is_flop = name in flop_cats
        peak = flop_cats[name] if is_flop else pk["hbm_gbs"] * 1e9
        out[name] = {"launches": v["launches"], "ms_total": v["ms"], "achieved": rate / (1e12 if is_flop else 1e9),
                     "unit": "TFLOP/s" if is_flop else "GB/s", "frac_of_peak": rate / peak}
        if name == "pconv":
            out[name]["peak_note"] = "fp32 FFMA pipe, 148 SMs x 128 lanes x 2 x 1.965 GHz = 74 TFLOP/s (nominal)"
    return out


def pointmlp_flops(points=361, embed=64, k=32, stages=4, blocks=2):
    """Algorithmic flop of one PointMLPModel(points, default widths) forward per sample: every Conv1d / Linear as 2 M N K."""
    f, n, c = 2.0 * points * 3 * embed, points, embed
    for _ in range(stages):
        n, out = n // 2, c * 2
        f += 2.0 * n * k * ((2 * c + 3) * out + blocks * 2 * out * out)      # transfer + pre blocks on (S*k) rows
        f += 2.0 * n * blocks * 2 * out * out                                 # pos blocks on S rows
        c = out
    return f + 2.0 * (c * 512 + 512 * 256 + 256 * 256)


def pointmlp_leg(a, device, ctx, pk):
    """The PointMLP scan encoder named by north_star (agent/feature_pcl.py), forward in train mode on scan point sets:
    scan_to_points + PointMLPModel(points=361, class_num=256).  Unwired in the reference's agents, so
    it is reported beside the update metric, not inside it."""
    from porl_b200.agent.feature_pcl import PointMLPModel, scan_to_points
    Bp = a.pointmlp_batch
    torch.manual_seed(1234)
    model = PointMLPModel(points=361, class_num=256).to(device).train()
    states = synth_rows_torch(Bp, device, 77)[:, :S].contiguous()

    def fwd():
        return model(scan_to_points(states))

    l0 = ctx.launches
    fwd()
    per = ctx.launches - l0
    ms = time_updates(fwd, 2, 5)
    flop = pointmlp_flops() * Bp
    out = {"workload": f"pointmlp_points361_E64_k32_b{Bp}", "forward_ms": ms, "samples_per_s": Bp / (ms * 1e-3),
           "gpu_launches_per_forward": per, "achieved_tflops": flop / (ms * 1e-3) * 1e-12,
           "gflop_per_sample": pointmlp_flops() * 1e-9}
    ctx.profile(True)
    fwd()
    torch.cuda.synchronize()
    out["kernels"] = kernel_table(ctx.profile_read(), pk)
    ctx.profile(False)
    if a.cpu_seconds > 0:
        from oracle import pointmlp as opm
        torch.set_num_threads(os.cpu_count() or 1)
        sd = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
        cfg = dict(points=361, k_neighbors=[32] * 4, reducers=[2] * 4, use_xyz=True, normalize="center")
        xb = scan_to_points(states[:8]).cpu()
        starts = torch.zeros(4, 8, dtype=torch.long)
        t0, n = time.perf_counter(), 0
        while n < 1 or time.perf_counter() - t0 < a.cpu_seconds / 3:
            opm.pointmlp_forward(sd, xb, cfg, starts, None)
            n += 1
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = {"samples_per_s": 8 * n / dt, "batch": 8, "cores": torch.get_num_threads(), "kind": "port",
                               "sample": f"{n} forwards of batch 8 (oracle restatement of PointMLPModel.forward, eager torch fp32)"}
    return out


def time_updates(step_fn, n_warm, n_timed):
    for _ in range(n_warm):
        step_fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n_timed):
        step_fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n_timed


def graph_of(step_fn):
    """CUDA graph of one update (workspaces are allocated by a first eager call)."""
    step_fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        step_fn()
    return g


def other_configs(a, device, table):
    """BASELINE configs 1 (POR, S=365, H=1024, batch 256 and 4096) and 3 (discrete CQL, QNetwork 362-64-128-64-5, batch
    4096): device-timed updates/s, fresh random rows every step, one CUDA graph replay per update (--graph 1)."""
    from types import SimpleNamespace
    from porl_b200 import CQLTrainer, POR, QNetwork, TransitionDataset
    out = {}
    g = torch.Generator(device=device).manual_seed(7)
    # config 1: POR on the 365-float rows of por_train.py (100k rows)
    rows = synth_rows_torch(100_000, device, 20260101, state_size=365)
    ds = TransitionDataset(rows, 365, A, device)
    args = SimpleNamespace(state_size=365, hidden_dim=1024, n_hidden=N_HIDDEN, layer_norm=False, feature_dim=256, action_size=A)
    for B in (256, 4096):
        torch.manual_seed(1000)
        agent = POR(args, max_steps=1_000_000, tau=0.9, alpha=10.0, device=device)
        agent.ctx.set_gemm_mode(a.gemm_mode)
        pb = ds.new_batch(B)
        idx = torch.zeros(B, dtype=torch.int64, device=device)

        def update():
            ds.gather_packed(idx, pb)
            agent.update_packed(pb)

        l0 = agent.ctx.launches
        graph = graph_of(update) if a.graph else None
        launches = (agent.ctx.launches - l0) // 2 if a.graph else None

        def step():
            torch.randint(0, len(ds), (B,), generator=g, device=device, out=idx)
            graph.replay() if graph is not None else update()

        ms = time_updates(step, 5, 50 if B == 256 else 20)
        agent.read_losses()
        out[f"por_mlp_S365_H1024_rows100k_b{B}"] = {"value": 1e3 / ms, "unit": UNIT, "ms_per_step": ms,
                                                     "achieved_tflops": 37.0e6 * B / (ms * 1e-3) * 1e-12,
                                                     "cuda_graph": bool(a.graph), "launches_per_step": launches}
        del agent, graph
    del ds, rows
    # config 3: discrete CQL on a replay buffer filled from the SORL table
    t = CQLTrainer(state_size=S, action_size=5, gamma=0.99, epsilon=1.0, epsilon_min=0.05, epsilon_decay=0.99,
                   update_target_freq=10, device=device, network=QNetwork, log_dir="/tmp/porl_b200_bench_logs")
    t.ctx.set_gemm_mode(a.gemm_mode)
    n = min(len(table), 1_000_000)
    from porl_b200.buffer.replaybuffer import ReplayBuffer
    t.replay_buffer = ReplayBuffer(n, (S,), device)
    idx = torch.arange(n, device=device)
    rowsb = table.gather(idx)
    t.replay_buffer.push_batch(rowsb[:, :S], torch.randint(0, 5, (n,), generator=g, device=device), rowsb[:, S],
                               rowsb[:, S + 1:2 * S + 1], rowsb[:, 2 * S + 1])
    del rowsb, idx
    for B in (4096, 256):
        idx = torch.zeros(B, dtype=torch.int64, device=device)
        update = lambda: t.learn_indices(idx, hard_every=10)
        l0 = t.ctx.launches
        graph = graph_of(update) if a.graph else None
        launches = (t.ctx.launches - l0) // 2 if a.graph else None

        def cql_step():
            torch.randint(0, n, (B,), generator=g, device=device, out=idx)
            graph.replay() if graph is not None else update()

        ms = time_updates(cql_step, 10, 300)
        t.read_losses()
        prof = None
        if t.fused:                                   # kernel time of the fused update, CUDA events inside the library
            t.ctx.profile(True)
            for _ in range(20):
                torch.randint(0, n, (B,), generator=g, device=device, out=idx)
                update()
            torch.cuda.synchronize()
            pr = t.ctx.profile_read()
            t.ctx.profile(False)
            c = pr["contraction"]
            us = c["ms"] / max(c["launches"], 1) * 1e3
            prof = {"kernel": "cql_fused_kernel (gather + 2 forwards + head + backward, "
                              + ("mma.sync 3xTF32 tiles)" if a.gemm_mode == 1 else "fp32 FFMA tiles)"),
                    "us_per_launch": us, "achieved_tflops": c["work"] / (c["ms"] * 1e-3) * 1e-12,
                    "frac_of_ffma_peak": c["work"] / (c["ms"] * 1e-3) / 74.0e12,
                    "gather_gbs": 2.0 * B * S * 4 / (us * 1e-6) * 1e-9}
        out[f"cql_q362-64-128-64-5_rows1M_b{B}"] = {
            "value": 1e3 / ms, "unit": UNIT, "ms_per_step": ms, "cuda_graph": bool(a.graph),
            "launches_per_step": launches, "fused_kernel": prof,
            "note": "one persistent kernel per update (csrc/cql_fused.cu) + partial sum + Adam" if t.fused
                    else "layer-by-layer kernels"}
        del graph
    return out


def param_checksum(agent):
    """(int64 sum of the fp32 bit patterns, float64 sum) over both optimizer groups: equal digests <=> bitwise equal
    replicas for all practical purposes; printed so that runs can be compared."""
    flat = torch.cat([agent.v_group.params, agent.p_group.params])
    return int(flat.view(torch.int32).to(torch.int64).sum()), float(flat.double().sum())


def cql_dp_leg(a, device, rank, world):
    """BASELINE config 3 under bench.py --gpus N: discrete CQL (QNetwork 362-64-128-64-5), global batch 4096 split over
    the ranks (strong) and 4096 rows per rank (weak); fused gradient kernel per rank, gradient sum inside the Adam
    kernel over NVLink peer memory, one CUDA graph per update; device-timed, max over ranks."""
    import torch.distributed as dist
    from porl_b200 import CQLTrainer, QNetwork
    from porl_b200.buffer.replaybuffer import ReplayBuffer
    n = 200_000
    torch.manual_seed(1000)
    t = CQLTrainer(state_size=S, action_size=5, gamma=0.99, epsilon=1.0, epsilon_min=0.05, epsilon_decay=0.99,
                   update_target_freq=10, device=device, network=QNetwork, log_dir="/tmp/porl_b200_bench_logs")
    t.ctx.set_gemm_mode(a.gemm_mode)
    g = torch.Generator(device=device).manual_seed(3)
    t.replay_buffer = ReplayBuffer(n, (S,), device)
    st = 3.5 * torch.rand(n, S, generator=g, device=device)
    t.replay_buffer.push_batch(st, torch.randint(0, 5, (n,), generator=g, device=device),
                               0.1 * torch.randn(n, generator=g, device=device),
                               st + 0.05 * torch.randn(n, S, generator=g, device=device),
                               (torch.rand(n, generator=g, device=device) < 0.01).float())
    del st
    out = {"gradient_sum": "inside the Adam kernel (NVLink peer memory)" if t.q_group.peers is not None else "NCCL all-reduce"}
    graphable = t.q_group.peers is not None
    for label, per_rank in (("strong_global4096", 4096 // world), ("weak_4096_per_gpu", 4096)):
        idx = torch.zeros(per_rank, dtype=torch.int64, device=device)
        upd = lambda: t.learn_indices(idx, hard_every=10)
        graph = graph_of(upd) if (a.graph and graphable) else None
        gi = torch.Generator(device=device).manual_seed(100 + rank)

        def step():
            torch.randint(0, n, (per_rank,), generator=gi, device=device, out=idx)
            graph.replay() if graph is not None else upd()

        for _ in range(10):
            step()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(300):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=device)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        t.read_losses()
        ups = 300 / (float(ms[0]) * 1e-3)
        out[label] = {"gradient_updates_per_s": ups, "samples_per_s": ups * per_rank * world, "rows_per_gpu": per_rank,
                      "global_batch": per_rank * world, "ms_per_update": float(ms[0]) / 300, "cuda_graph": graph is not None}
        del graph
    digest = torch.tensor([int(t.q_group.params.view(torch.int32).to(torch.int64).sum())], device=device)
    all_d = [torch.zeros_like(digest) for _ in range(world)]
    dist.all_gather(all_d, digest)
    out["replicas_bitwise_identical"] = all(int(d) == int(all_d[0]) for d in all_d)
    return out


def config4_leg(a, device, rank, world, table, per_gpu):
    """BASELINE config 4: POR with the FasterNet scan encoder, large batch data-parallel (65536 over 8 GPUs = 8192 per
    GPU).  Every BatchNorm of the encoder sees the GLOBAL batch (cross-rank sums between the statistics and the
    normalisation), gradients of the heads are summed over the ranks; device-timed, max over ranks."""
    import torch.distributed as dist
    from types import SimpleNamespace
    from porl_b200 import POR
    from porl_b200.agent.fasternet import FasterNet
    args = SimpleNamespace(state_size=S, hidden_dim=1024, n_hidden=N_HIDDEN, layer_norm=False, feature_dim=256, action_size=A)
    torch.manual_seed(1000)
    agent = POR(args, max_steps=1_000_000, tau=0.9, alpha=10.0, device=device, backbone=FasterNet(3, args.feature_dim))
    agent.ctx.set_gemm_mode(a.gemm_mode)
    pb = table.new_batch(per_gpu)           # (data parallel: the encoder picks up the default process group by itself)
    g = torch.Generator(device=device).manual_seed(4000 + rank)
    idx = torch.zeros(per_gpu, dtype=torch.int64, device=device)

    def step():
        torch.randint(0, len(table), (per_gpu,), generator=g, device=device, out=idx)
        table.gather_packed(idx, pb)
        agent.update_packed(pb)

    ctx = agent.ctx
    for _ in range(2):
        step()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    n = 5
    l0 = ctx.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    losses = agent.read_losses()
    ms_step = float(ms[0]) / n
    out = {"workload": f"por_fasternet_S{S}_F256_H1024_b{per_gpu}x{world}", "global_batch": per_gpu * world,
           "rows_per_gpu": per_gpu, "gradient_updates_per_s": 1e3 / ms_step, "samples_per_s": per_gpu * world * 1e3 / ms_step,
           "ms_per_update": ms_step, "gpu_launches_per_update": (ctx.launches - l0) // n, "final_losses": losses,
           # per sample: two encoder forwards (0.860 GMAC each) + the heads on 256 features
           # (ten value-net and three policy-net forward-equivalents: 18.2 MMAC)
           "achieved_tflops_per_gpu": (2 * 2 * 0.860e9 + 36.3e6) * per_gpu / (ms_step * 1e-3) * 1e-12,
           "sync_batchnorm": "per-channel sums all-reduced across ranks between statistics and normalisation" if world > 1 else "single process"}
    ctx.profile(True)
    step()
    torch.cuda.synchronize()
    prof = ctx.profile_read()
    ctx.profile(False)
    tot = sum(v["ms"] for v in prof.values())
    out["kernel_time_fractions"] = {k: round(v["ms"] / tot, 4) for k, v in prof.items() if v["launches"]}
    out["kernel_ms"] = {k: round(v["ms"], 3) for k, v in prof.items() if v["launches"]}
    return out


def run_b200(a):
    from porl_b200 import TransitionDataset, _C, dist as pdist
    rank, local, world = pdist.init_from_env()
    if world != a.gpus and rank == 0 and world > 1:
        print(f"warning: --gpus {a.gpus} but WORLD_SIZE {world}", file=sys.stderr)
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    dist_on = world > 1
    pk = peaks()
    strong = dist_on and a.scaling == "strong"
    if strong and a.batch % world:
        raise SystemExit(f"--scaling strong: batch {a.batch} is not divisible by {world} ranks")
    b_local = a.batch // world if strong else a.batch

    rows = synth_rows_torch(a.rows, device, 20260102)          # same replica on every rank
    table = TransitionDataset(rows, S, A, device)
    del rows
    runner = Runner(a, b_local, device, rank, world, table)
    K, W = a.steps, max(a.warmup, 3)
    peer_reduce = dist_on and all(g.peers is not None for g in (runner.agent.v_group, runner.agent.p_group))
    use_graph = bool(a.graph) and (not dist_on or peer_reduce)   # the peer-sum Adam kernel is an ordinary launch
    if use_graph:
        runner.capture()
    idx_warm, idx_timed = runner.indices(W, 2000), runner.indices(K, 2001)

    with ClockSampler(local) as clk:
        ms, launches = timed(runner, idx_warm, idx_timed, dist_on)
    v_loss, g_loss = runner.agent.read_losses()
    clocks = clk.summary()
    updates_per_s = K / (ms * 1e-3)                 # optimizer steps per second (every rank takes the same step)
    # `value`: weak scaling counts batch-`--batch` minibatches per second over all ranks (per-GPU work fixed, the
    # global batch grows with N); strong scaling keeps the global batch, so value = gradient updates per second
    value = updates_per_s if strong else world * updates_per_s
    digest, fsum = param_checksum(runner.agent)
    identical = True
    if dist_on:
        import torch.distributed as dist
        d = torch.tensor([digest], device=device)
        all_d = [torch.zeros_like(d) for _ in range(world)]
        dist.all_gather(all_d, d)
        identical = all(int(x) == int(all_d[0]) for x in all_d)
        assert identical, f"data-parallel replicas diverged: parameter digests {[int(x) for x in all_d]}"

    line = None
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong" if strong else "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "gradient_updates_per_s": updates_per_s, "samples_per_s": updates_per_s * b_local * world,
                "value_definition": ("gradient updates/s at a fixed global batch" if strong or not dist_on else
                                     f"batch-{a.batch} minibatches/s summed over {world} ranks (weak scaling: every "
                                     f"optimizer step consumes {world} of them; gradient_updates_per_s is the step rate)"),
                "config": {"workload": workload_name(a.batch), "algo": "SORL (agent/sorl.py, MLP-only heads)",
                           "batch_per_gpu": b_local, "global_batch": b_local * world, "rows": a.rows,
                           "state_size": S, "hidden_dim": H, "parallelism": f"dp{world}" if dist_on else "single",
                           "contraction": "tcgen05 3xTF32" if a.gemm_mode == 1 else "fp32 FFMA",
                           "cuda_graph": use_graph, "inputs_larger_than_l2": True,
                           "programmatic_dependent_launch": os.environ.get("PORL_PDL", "1") != "0",
                           "gradient_sum": ("none (single process)" if not dist_on else
                                            "peer-memory loads fused into the Adam kernel (NVLink P2P)" if peer_reduce
                                            else "NCCL all-reduce per optimizer group"),
                           "l2_note": "each step gathers fresh random rows of a 2.9 GB HBM table (>> 126 MB L2)"},
                "clocks": clocks, "gpu_launches": launches * (1 if not use_graph else 1),
                "final_losses": {"v_loss": v_loss, "g_loss": g_loss},
                "replicas_bitwise_identical": identical, "param_checksum": {"int64_bit_sum": digest, "f64_sum": fsum}}
        if use_graph:
            # graph replays re-issue the captured launches; count them from the capture
            line["gpu_launches"] = runner.launches_per_step * K if hasattr(runner, "launches_per_step") else launches

    # ---- single-process extras on rank 0: roofline, e2e, cpu baseline, second batch size
    if not dist_on:
        ctx = runner.agent.ctx
        # launches per step (eager) for the gpu_launches claim under graph replay
        l0 = ctx.launches
        runner.step(idx_timed[0])
        per_step = ctx.launches - l0
        line["gpu_launches"] = per_step * K
        # per-kernel timing with events around every launch (eager, after the timed region)
        ctx.profile(True)
        for i in range(min(10, K)):
            runner.step(idx_timed[i])
        torch.cuda.synchronize()
        prof = ctx.profile_read()
        ctx.profile(False)
        c = prof["contraction"]
        achieved = c["work"] / (c["ms"] * 1e-3) / 1e12 if c["ms"] > 0 else 0.0
        peak = pk["bf16_sustained"]
        traffic, traffic_note = traffic_record(a.gemm_mode)
        ceil = measure_ceilings(ctx, device) if a.ceilings else None
        ceiling3 = ceil["ceiling_3xtf32_tflops"] if ceil else peak / 6.0
        line["roofline"] = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                            "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_note,
                            "frac_of_3xtf32_ceiling": achieved / ceiling3,
                            "ceiling_3xtf32_tflops": ceiling3,
                            "frac_of_mma_form_rate": (achieved / ceil["mma_form"]["form_ceiling_3xtf32_tflops"]
                                                      if ceil and "form_ceiling_3xtf32_tflops" in ceil.get("mma_form", {}) else None),
                            "ceiling_source": "measured in this run: cuBLAS TF32 8192^3 / 3" if ceil else "bf16 sustained / 6 (not measured)",
                            "measured_ceilings": ceil,
                            "kernel": "gemm_tc_kernel (tcgen05 3xTF32)" if a.gemm_mode == 1 else "gemm_simt_kernel (fp32 FFMA)",
                            "peak_source": f"MEASURED_PEAKS.json bf16_tflops_sustained ({pk['source']})",
                            "launches_per_step": c["launches"] / max(1, min(10, K)),
                            "share_of_step_ms": c["ms"] / max(1e-9, sum(v["ms"] for v in prof.values()))}
        kern = kernel_table(prof, pk)
        line["kernels"] = kern

        # e2e: the public API with HOST buffers: pinned rows -> H2D -> agent.update_async() -> losses D2H, every step.
        # Four device row buffers: while update(i) runs, the rows of steps i+1 .. i+3 travel host -> device on a side
        # stream (a DataLoader with pin_memory and prefetch_factor 3), and the losses of update(i-1) are read on the host
        # -- one step late, each through its own event -- so the compute stream never drains.  Every step's H2D copy and
        # its loss read-back lie inside the timed region.  SORL.update() (the reference's blocking signature) is timed
        # beside it.
        Ke = a.e2e_steps or K
        ring, depth = 8, int(os.environ.get("PORL_E2E_DEPTH", "3"))
        host = torch.empty(ring, a.batch, table.row_len, dtype=torch.float32).pin_memory()
        src_idx = runner.indices(ring, 2002)
        for r in range(ring):
            host[r].copy_(table.gather(src_idx[r]).cpu())
        nbuf = depth + 1
        dev_rows = [torch.empty(a.batch, table.row_len, dtype=torch.float32, device=device) for _ in range(nbuf)]
        copy_stream = torch.cuda.Stream(device=device)
        arrived = [torch.cuda.Event() for _ in range(nbuf)]
        consumed = [torch.cuda.Event() for _ in range(nbuf)]
        main = torch.cuda.current_stream()

        def prefetch(i):
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(consumed[i % nbuf])        # the update that last read this buffer has been issued and run
                dev_rows[i % nbuf].copy_(host[i % ring], non_blocking=True)
                arrived[i % nbuf].record(copy_stream)

        def views(rows):
            return rows[:, :S], rows[:, 2 * S + 2:], rows[:, S], rows[:, S + 1:2 * S + 1], rows[:, 2 * S + 1]

        def e2e_loop(n, blocking):
            for bslot in range(nbuf):
                consumed[bslot].record(main)
            for j in range(depth):
                prefetch(j)
            pending, losses = None, None
            for i in range(n):
                main.wait_event(arrived[i % nbuf])
                if blocking:
                    losses = runner.agent.update(*views(dev_rows[i % nbuf]))
                else:
                    h = runner.agent.update_async(*views(dev_rows[i % nbuf]))
                consumed[i % nbuf].record(main)
                prefetch(i + depth)
                if not blocking:
                    if pending is not None:
                        losses = pending.result()            # losses of update(i-1), read while update(i) runs
                    pending = h
            if pending is not None:
                losses = pending.result()
            torch.cuda.synchronize()
            return losses

        e2e_loop(5, False)
        t0 = time.perf_counter()
        e2e_loop(Ke, False)
        dt = time.perf_counter() - t0
        e2e_loop(3, True)
        t0 = time.perf_counter()
        e2e_loop(Ke, True)
        dt_block = time.perf_counter() - t0
        line["e2e"] = {"value": Ke / dt, "unit": UNIT, "h2d_bytes_per_step": a.batch * table.row_len * 4,
                       "d2h_bytes_per_step": 12, "steps": Ke,
                       "h2d_overlap": f"rows of steps i+1 .. i+{depth} are copied on a side stream while update(i) runs",
                       "api": "SORL.update_async(observations, actions, rewards, next_observations, terminals) -> handle; "
                              "handle.result() -> (float, float), read one step late (every step's losses are read)",
                       "blocking_api": {"value": Ke / dt_block, "unit": UNIT,
                                        "api": "SORL.update(...) -> (float, float): the reference's signature, one host "
                                               "synchronisation per step"},
                       "frac_of_device_value": (Ke / dt) / value}
        # the host link: the same pinned -> device copy alone (what bounds e2e once an update is faster than its rows' copy)
        # (the loop's own ring of pinned buffers, 95 MB, so that the host side reads from DRAM as it does in the loop)
        nbytes = a.batch * table.row_len * 4
        dev_rows[0].copy_(host[0], non_blocking=True)
        torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for j in range(3 * ring):
            dev_rows[j % nbuf].copy_(host[j % ring], non_blocking=True)
        ev1.record()
        torch.cuda.synchronize()
        link = nbytes * 3 * ring / (ev0.elapsed_time(ev1) * 1e-3) * 1e-9
        line["e2e"]["h2d_gbs_in_loop"] = nbytes * (Ke / dt) * 1e-9
        line["e2e"]["h2d_link_gbs"] = link
        line["e2e"]["link_bound_updates_per_s"] = link * 1e9 / nbytes

        # second batch size of the metric (batch 256)
        if a.also_batch and a.also_batch != a.batch:
            r2 = Runner(a, a.also_batch, device, rank, world, table)
            if use_graph:
                r2.capture()
            ms2, _ = timed(r2, r2.indices(W, 2003), r2.indices(K * 4, 2004), False)
            line["also"] = {"workload": workload_name(a.also_batch), "value": K * 4 / (ms2 * 1e-3), "unit": UNIT,
                            "ms_per_step": ms2 / (K * 4)}

        # the literal sorl_train.py agent: FasterNet scan encoder in front of the heads (config 2b).  Single process: the
        # whole step (gather, two encoder forwards with their DropPath draws from torch's device generator, heads) is
        # replayed as one CUDA graph; data parallel: eager (the encoder's BatchNorm sums go through a torch.distributed
        # callback)
        if a.encoder_batch:
            Be = a.encoder_batch
            r3 = Runner(a, Be, device, rank, world, table, backbone=True)
            enc_graph = use_graph and not dist_on
            if enc_graph:
                r3.run(r3.indices(2, 2008))           # eager calls allocate the workspaces
                r3.capture()
            ms3, l3 = timed(r3, r3.indices(3, 2005), r3.indices(10, 2006), False)
            r3.agent.read_losses()                    # raises on device flags (scan range, NaN)
            flop = 3.449e9 * Be                       # 2 encoder forwards + heads, algorithmic
            line["also_encoder"] = {"workload": f"sorl_fasternet_S{S}_F256_H{H}_rows1M_b{Be}", "value": 10 / (ms3 * 1e-3),
                                    "unit": UNIT, "ms_per_step": ms3 / 10, "gpu_launches": l3,
                                    "achieved_tflops": flop / (ms3 / 10 * 1e-3) * 1e-12, "graph": bool(enc_graph)}
            if enc_graph:
                line["also_encoder"]["gpu_launches"] = r3.launches_per_step * 10
                r3.graph = None                       # the per-kernel table below needs eager launches
            ctx.profile(True)
            r3.run(r3.indices(2, 2007))
            torch.cuda.synchronize()
            line["also_encoder"]["kernels"] = kernel_table(ctx.profile_read(), pk)
            ctx.profile(False)
            del r3
        if a.pointmlp_batch:
            line["also_pointmlp"] = pointmlp_leg(a, device, ctx, pk)
        if a.other_configs:
            line["also_configs"] = other_configs(a, device, table)

        # CPU baseline beside it (bounded sample; reported, not the target)
        if a.cpu_seconds > 0:
            torch.set_num_threads(os.cpu_count() or 1)
            ups, n, threads, _ = cpu_reference_updates(a.batch, 10_000, 1, seconds=a.cpu_seconds)
            line["cpu_baseline"] = {"value": ups, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": f"{n} updates of batch {a.batch} in <= {a.cpu_seconds:.0f} s "
                                              "(oracle restatement of agent/sorl.py, eager torch fp32)"}
            if a.encoder_batch and "also_encoder" in line:
                ups_e, sps_e, n_e = cpu_reference_encoder(32, a.cpu_seconds / 3)
                line["also_encoder"]["cpu_baseline"] = {
                    "value": ups_e, "unit": UNIT, "samples_per_s": sps_e, "batch": 32, "cores": threads, "kind": "port",
                    "sample": f"{n_e} updates of batch 32 (oracle SORL + forward-only FasterNet; the reference also "
                              "back-propagates through the encoder)"}
                line["also_encoder"]["samples_per_s"] = line["also_encoder"]["value"] * a.encoder_batch
    else:
        cql = cql_dp_leg(a, device, rank, world) if a.other_configs else None
        per4 = a.cfg4_batch if a.cfg4_batch >= 0 else (8192 if world == 8 else 0)
        cfg4 = config4_leg(a, device, rank, world, table, per4) if per4 > 0 else None
        if rank == 0:
            # the host-buffer end-to-end figure is an N=1 measurement; at N>1 the inputs are device-resident
            line["e2e"] = {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                           "copies_declared": False,
                           "note": "multi-GPU run: device-resident inputs, NOT an end-to-end figure (this repeats the "
                                   "device-timed value); the host-buffer e2e is measured at N=1"}
            if cql is not None:
                line["also_cql_dp"] = cql
            if cfg4 is not None:
                line["also_config4"] = cfg4
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist_on:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


def main():
    # the contract is ONE JSON line on stdout: NCCL's version banner (NCCL_DEBUG=VERSION) would precede it
    if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
        os.environ["NCCL_DEBUG"] = "WARN"
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)


if __name__ == "__main__":
    main()
