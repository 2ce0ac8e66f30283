"""Per-phase share of the warp-state samples of one kernel: the SASS of `ncu --page source --csv` is cut at every
BAR.SYNC, and samples / executed instructions / HMMA / LDS counts are summed per segment.

    python -m tools.ncu_phases gpurun_out/x.ncu-rep >> profiles/x_ncu.txt
"""
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(out)))
    h, rows = r[1], r[2:]
    col = {k: h.index(k) for k in ("Source", "# Samples", "Instructions Executed", "L1 Wavefronts Shared",
                                  "L1 Wavefronts Shared Ideal")}
    new = lambda start: dict(start=start, static=0, samples=0, executed=0, hmma=0, ffma=0, lds=0, stg=0, wf=0, wfi=0)
    segs, cur, total = [], new(0), 0
    num = lambda x: int(x) if x and x.isdigit() else 0
    for k, x in enumerate(rows):
        src, n, e = x[col["Source"]], num(x[col["# Samples"]]), num(x[col["Instructions Executed"]])
        cur["samples"] += n; cur["static"] += 1; cur["executed"] += e; total += n
        cur["wf"] += num(x[col["L1 Wavefronts Shared"]]); cur["wfi"] += num(x[col["L1 Wavefronts Shared Ideal"]])
        for key, pat in (("hmma", "HMMA"), ("ffma", "FFMA"), ("lds", "LDS"), ("stg", "STG")):
            if pat in src:
                cur[key] += e
        if "BAR.SYNC" in src or "EXIT" in src:
            segs.append(cur); cur = new(k + 1)
    print(f"\n== phases between barriers of {rows and r[0][1]} ({total} samples; static SASS instructions {len(rows)})")
    print(f"{'seg':>3} {'first':>6} {'static':>6} {'share':>6} {'executed':>9} {'HMMA':>7} {'FFMA':>8} {'LDS':>7} {'STG':>6} "
          f"{'smem wavefronts':>15} {'ideal':>8}")
    for i, c in enumerate(segs):
        if c["samples"] * 200 >= total:
            print(f"{i:3d} {c['start']:6d} {c['static']:6d} {100 * c['samples'] / total:5.1f}% {c['executed']:9d} {c['hmma']:7d} "
                  f"{c['ffma']:8d} {c['lds']:7d} {c['stg']:6d} {c['wf']:15d} {c['wfi']:8d}")


if __name__ == "__main__":
    main()
