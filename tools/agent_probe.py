"""Per-tensor gradient comparison of one agent update (gemm mode 1 vs the CPU oracle).
   python tools/agent_probe.py por 365 256 2 64 [mode]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from oracle import synth
from oracle.buffer import slice_row
from test_gpu_agents import make_agent, step, rel_l2

kind, S, H, A, B = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
mode = int(sys.argv[6]) if len(sys.argv) > 6 else 1
agent, oracle = make_agent(kind, S, H, A, 1000, 0.8, 3.0 if kind == "sorl" else 10.0, 5, gemm_mode=mode)
rows = torch.from_numpy(synth.laser_rows(B * 3, S, 17, A))
chunk = rows[:B]
want = step(oracle, kind, slice_row(chunk.clone(), S, A))
got = step(agent, kind, slice_row(chunk.to("cuda:0"), S, A))
print("losses", got, want, "tc launches", agent.ctx.tc_launches)
ref = {k: p.grad for k, p in oracle.named_parameters() if p.grad is not None}
for k, p in agent.named_parameters():
    if k in ref:
        d = (p.grad.cpu().double() - ref[k].double()).abs()
        bad = (d > 1e-5 * ref[k].abs().max()).nonzero()
        print(f"{k:28s} rel_l2 {rel_l2(p.grad, ref[k]):.2e} max|d| {float(d.max()):.2e} ref_max {float(ref[k].abs().max()):.2e} "
              f"n_bad {len(bad)} first_bad {bad[:3].tolist()}")
