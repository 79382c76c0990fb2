"""Builds trase_b200/libtrase_b200.so (sm_100a) with plain nvcc calls.

In-tree build so the shared library travels with the repository snapshot to
the GPU box.  No JIT, no torch cpp_extension: the library has a C ABI and no
torch dependency.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
BUILD = os.path.join(HERE, "csrc", "build")
LIB = os.path.join(HERE, "libtrase_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
# default -fmad=true, no fast-math: integer parity with the reference depends on it
FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden"]
# debugging / measurement builds: extra nvcc flags (e.g. -DTB_FU_PROF=1 for tools/fused_mlp_stalls.py) and another output
# name, so that the shipped library is never overwritten by such a build
FLAGS += os.environ.get("TRASE_B200_EXTRA_NVCC_FLAGS", "").split()
if os.environ.get("TRASE_B200_LIB_OUT"):
    LIB = os.path.join(HERE, os.environ["TRASE_B200_LIB_OUT"])
    BUILD = os.path.join(HERE, "csrc", "build_" + os.path.splitext(os.environ["TRASE_B200_LIB_OUT"])[0])

SOURCES = ["api.cu", "preprocess.cu", "binning.cu", "render.cu", "render_mma.cu", "no_alloc.cu", "mlp_gemm.cu", "assemble.cu", "mlp_prep.cu", "loss.cu", "mlp_fused.cu", "optim.cu", "contrastive.cu"]
HEADERS = [os.path.join(CSRC, "common.cuh"), os.path.join(CSRC, "ptx.cuh"), os.path.join(CSRC, "render_common.cuh"), os.path.join(HERE, "..", "include", "trase_b200.h")]


def _stale(out: str, deps: list[str]) -> bool:
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(d) > t for d in deps)


def build(verbose: bool = False, force: bool = False) -> str:
    os.makedirs(BUILD, exist_ok=True)
    extra = ["-Xptxas", "-v"] if verbose else []
    objs, jobs = [], []
    for src in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(BUILD, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _stale(o, [s] + HEADERS):
            jobs.append([NVCC, *ARCH, *FLAGS, *extra, "-c", s, "-o", o])

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))
        return r.stderr

    with ThreadPoolExecutor(max_workers=len(jobs) or 1) as ex:
        for log in ex.map(run, jobs):
            if verbose and log:
                print(log, file=sys.stderr)
    if force or jobs or _stale(LIB, objs):
        run([NVCC, *ARCH, "-shared", "-o", LIB, *objs, "-Xcompiler", "-fvisibility=hidden"])
    return LIB


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
