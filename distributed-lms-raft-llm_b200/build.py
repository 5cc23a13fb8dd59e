"""Builds libtutor_b200.so in-tree with nvcc for sm_100a (no JIT cache: the .so travels with the repo)."""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libtutor_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--expt-relaxed-constexpr",
         "-Xcompiler", "-fPIC", "-Xptxas", "-v"]
if os.environ.get("TUTOR_EXPERIMENTS"):  # A/B knobs for tools/ (csrc/common.cuh); never set for the shipped library
    FLAGS.append("-DTUTOR_EXPERIMENTS")
FLAGS += os.environ.get("TUTOR_EXTRA_NVCC_FLAGS", "").split()  # one-off timing experiments (tools/)


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cpp")))


def _newest_input():
    files = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "tutor_b200.h"),
                                                                __file__]
    return max(os.path.getmtime(f) for f in files)


def up_to_date():
    return os.path.exists(LIB) and os.path.getmtime(LIB) >= _newest_input()


def _compile(src, force=False):
    obj = os.path.join(OBJ, os.path.splitext(src)[0] + ".o")
    hdr_time = max(os.path.getmtime(os.path.join(CSRC, f)) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh", ".inc")))
    hdr_time = max(hdr_time, os.path.getmtime(os.path.join(HERE, "..", "include", "tutor_b200.h")))
    if not force and os.path.exists(obj) and os.path.getmtime(obj) >= max(os.path.getmtime(os.path.join(CSRC, src)), hdr_time):
        return obj, ""
    cmd = [NVCC, *FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
    p = subprocess.run(cmd, capture_output=True, text=True)
    if p.returncode != 0:
        raise RuntimeError(f"nvcc failed on {src}:\n{p.stdout}\n{p.stderr}")
    return obj, p.stderr


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and up_to_date():
        return LIB
    if not os.path.exists(NVCC):
        raise RuntimeError(f"nvcc not found at {NVCC}; libtutor_b200.so must be prebuilt")
    os.makedirs(OBJ, exist_ok=True)
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        results = list(ex.map(lambda s: _compile(s, force), sources()))
    if verbose:
        for _, log in results:
            sys.stderr.write(log)
    objs = [o for o, _ in results]
    cmd = [NVCC, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-ldl"]
    p = subprocess.run(cmd, capture_output=True, text=True)
    if p.returncode != 0:
        raise RuntimeError(f"link failed:\n{p.stdout}\n{p.stderr}")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
