"""Text summaries of ncu captures for profiles/ (the .ncu-rep files themselves stay in gpurun_out/).

    python -m tools.ncu_summary full   gpurun_out/x.ncu-rep  "<command line>"  > profiles/x_ncu.txt
    python -m tools.ncu_summary traffic gpurun_out/x.ncu-rep kernel_name       > profiles/x_traffic.json
    python -m tools.ncu_summary list   gpurun_out/launches.csv [LAST_N | START:COUNT] > profiles/x_step.txt
"""
from __future__ import annotations

import collections
import csv
import io
import json
import re
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__cluster_size", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tc.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",   # legacy HMMA path (mma.sync)
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__cycles_active.avg",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
    "smsp__average_warp_latency_issue_stalled_short_scoreboard.ratio",
    "smsp__average_warp_latency_issue_stalled_barrier.ratio",
    "smsp__average_warp_latency_issue_stalled_mio_throttle.ratio",
    "smsp__average_warp_latency_issue_stalled_math_pipe_throttle.ratio",
    "smsp__average_warp_latency_issue_stalled_not_selected.ratio",
    "smsp__average_warp_latency_issue_stalled_wait.ratio",
    "smsp__average_warp_latency_issue_stalled_sleeping.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
]


def raw_rows(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    return rows[0], rows[1], rows[2:]


def short(name):
    name = re.sub(r"\(.*", "", name)
    return name.split("::")[-1]


def full(rep, cmd):
    hdr, units, rows = raw_rows(rep)
    print(f"# {cmd}")
    print(f"# selected metrics of `ncu -i {rep.split('/')[-1]} --page raw --csv` (full report not committed)")
    kcol = hdr.index("Kernel Name")
    for i, r in enumerate(rows):
        print(f"\n== {short(r[kcol])}  (launch id {i})")
        for m in METRICS:
            if m in hdr:
                j = hdr.index(m)
                print(f"  {m:<92s} {r[j]:>18s} {units[j]}")


def traffic(rep, kernel):
    hdr, units, rows = raw_rows(rep)
    rd, wr, du = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("gpu__time_duration.sum")
    scale = {"Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Gbyte": 1e9}
    per = [int(float(r[rd].replace(",", "")) * scale[units[rd]] + float(r[wr].replace(",", "")) * scale[units[wr]])
           for r in rows if kernel in r[hdr.index("Kernel Name")]]
    dur = [float(r[du].replace(",", "")) for r in rows if kernel in r[hdr.index("Kernel Name")]]
    import hashlib, os
    src = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "porl_b200", "csrc", "gemm_tc.cu")
    sha = hashlib.sha256(open(src, "rb").read()).hexdigest() if "gemm_tc" in kernel else None
    print(json.dumps({"kernel": kernel, "source": f"{rep.split('/')[-1]} (ncu --set full --clock-control none)",
                      "source_sha256": sha,
                      "dram_bytes_per_launch": per, "duration_us_under_ncu": dur,
                      "mean_bytes_per_launch": sum(per) // max(len(per), 1)}, indent=1))


def launch_list(path, last_n=None):
    rows = []
    with open(path) as f:
        lines = [ln for ln in f if ln.startswith('"')]
    rd = csv.DictReader(io.StringIO("".join(lines)))
    for r in rd:
        if r["Metric Name"] == "gpu__time_duration.sum":
            rows.append((short(r["Kernel Name"]), float(r["Metric Value"].replace(",", "")) / 1e3, r["Grid Size"], r["Block Size"]))
    if last_n and ":" in str(last_n):            # START:COUNT window of launch indices (one step of the loop)
        a, c = (int(v) for v in last_n.split(":"))
        rows = rows[a:a + c]
    elif last_n:
        rows = rows[-int(last_n):]
    tot = sum(t for _, t, _, _ in rows)
    agg = collections.OrderedDict()
    for n, t, _, _ in rows:
        c, s = agg.get(n, (0, 0.0))
        agg[n] = (c + 1, s + t)
    print(f"# {path.split('/')[-1]}: {len(rows)} launches, {tot:.1f} us serialised (cold-cache ncu replay; shares, not absolutes)")
    print(f"{'kernel':<44s} {'launches':>8s} {'us':>10s} {'share':>7s}")
    for n, (c, s) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{n[:44]:<44s} {c:>8d} {s:>10.1f} {s / tot:>7.3f}")
    print("\n# in launch order")
    for n, t, g, b in rows:
        print(f"{n[:44]:<44s} {t:>9.2f} us  grid {g} block {b}")


if __name__ == "__main__":
    mode = sys.argv[1]
    if mode == "full":
        full(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
    elif mode == "traffic":
        traffic(sys.argv[2], sys.argv[3])
    else:
        launch_list(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else None)
