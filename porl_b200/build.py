"""Build recipe for libporl_b200.so (sm_100a only, in-tree so that it travels to the GPU box).

`python -m porl_b200.build` or `porl_b200.build.build()`; __graft_entry__.build() calls this.
nvcc cross-compiles without a GPU.  The library links the static CUDA runtime and has no
dependency on torch's C++ ABI: it is loaded with ctypes (porl_b200/_C.py).
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libporl_b200.so")
SOURCES = ["api.cu", "cql_fused.cu", "encoder.cu", "gather.cu", "gemm_simt.cu", "gemm_tc.cu", "head.cu", "loss.cu", "mlp.cu", "optim.cu", "sampler.cu", "pointmlp.cu", "step.cu", "transpose.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "--threads", "0"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libporl_b200.so cannot be built")


def _stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "porl_b200.h")]
    probe = os.path.join(HERE, "..", "tools", "micro", "umma_rate.cu")
    if os.path.exists(probe):
        if not os.path.exists(probe[:-3]):
            return True
        deps.append(probe)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every CUDA source for sm_100a into porl_b200/libporl_b200.so; returns its path."""
    if not force and not _stale():
        return LIB
    objs = []
    build_dir = os.path.join(HERE, "build")
    os.makedirs(build_dir, exist_ok=True)
    nvcc = _nvcc()
    procs = []
    for src in SOURCES:
        obj = os.path.join(build_dir, src.replace(".cu", ".o"))
        objs.append(obj)
        cmd = [nvcc, *NVCC_FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for src, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode != 0:
            sys.stderr.write(out)
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}")
    link = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB, *objs]
    subprocess.run(link, check=True)
    # the tcgen05 instruction-rate probe bench.py runs beside the library (tools/micro/umma_rate.cu): a stand-alone binary
    probe_src = os.path.join(HERE, "..", "tools", "micro", "umma_rate.cu")
    if os.path.exists(probe_src):
        subprocess.run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-o", probe_src[:-3], probe_src], check=True)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
