"""SASS instruction-count listing of every kernel in the in-tree sm_100a objects (porl_b200/build/*.o).

    python tools/sass_counts.py > profiles/rNN_sass_counts.txt

cuobjdump -sass runs on the build machine (no GPU needed).  The counts are static (instructions in the binary, not
executed): they prove which hardware paths a kernel uses -- UTCHMMA (tcgen05.mma), UTMALDG / UTMASTG (TMA tensor load /
store), LDTM / STTM (tcgen05.ld / .st), UTCBAR (tcgen05.commit), SYNCS (mbarrier), HMMA (legacy warp-level mma.sync)."""
import collections
import hashlib
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ["UTCHMMA", "UTCBAR", "UTMALDG", "UTMASTG", "LDTM", "STTM", "SYNCS", "HMMA", "FFMA", "DFMA", "MUFU", "LDS", "STS", "LDG",
        "STG", "ATOMG", "RED", "BAR", "MEMBAR", "R2UR", "ELECT"]


def per_function(obj):
    txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    funcs, cur = {}, None
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur:
            funcs[cur][m.group(1).split(".")[0]] += 1
    return funcs


def demangle(name):
    d = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    d = d.replace("(anonymous namespace)::", "").replace("porl::", "")
    return re.sub(r"\(.*", "", d).replace("void ", "")


def main():
    src = open(os.path.join(ROOT, "porl_b200", "csrc", "gemm_tc.cu"), "rb").read()
    print("# SASS instruction counts per kernel: cuobjdump -sass of porl_b200/build/*.o (sm_100a), static counts")
    print(f"# gemm_tc.cu sha256 {hashlib.sha256(src).hexdigest()}")
    print("# UTCHMMA tcgen05.mma | UTMALDG / UTMASTG TMA tensor load / store | LDTM / STTM tcgen05.ld / .st | UTCBAR tcgen05.commit |")
    print("# SYNCS mbarrier | HMMA warp-level mma.sync (legacy path: the fused CQL kernel only) | R2UR / ELECT single-thread issue paths")
    print(f"{'kernel':<56} {'instrs':>7}  counts")
    build = os.path.join(ROOT, "porl_b200", "build")
    for obj in sorted(os.listdir(build)):
        if not obj.endswith(".o"):
            continue
        rows = []
        for name, c in per_function(os.path.join(build, obj)).items():
            tot = sum(c.values())
            if tot >= 50:
                rows.append((tot, demangle(name)[:55], "  ".join(f"{k}={c[k]}" for k in KEYS if c[k])))
        if rows:
            print(f"## {obj}")
        for tot, name, counts in sorted(rows, reverse=True):
            print(f"{name:<56} {tot:>7}  {counts}")


if __name__ == "__main__":
    main()
