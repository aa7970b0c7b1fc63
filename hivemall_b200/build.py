"""In-tree build of libhivemall_b200.so (nvcc, sm_100a only) and of the CPU oracle.

The shared library is plain CUDA C++ behind the C-ABI of include/hivemall_b200.h; it
does not link torch.  `python -m hivemall_b200.build` rebuilds when sources are newer.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "hivemall_b200", "csrc")
LIB = os.path.join(ROOT, "hivemall_b200", "libhivemall_b200.so")
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_LIB = os.path.join(ORACLE_DIR, "_build", "libhm_oracle.so")

SOURCES = ["hm_options.cpp", "hm_table.cpp", "hm_hash.cu", "hm_sort.cu", "hm_linear.cu", "hm_fm.cu", "hm_learner.cu"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-fmad=false",            # every fp32 operation is rounded on its own
    "-Xcompiler", "-fPIC,-O2,-Wall",
    "-shared",
    "-cudart", "shared",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: hivemall_b200 is CUDA-only and cannot be built without it")


def _newer(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force: bool = False, verbose: bool = False) -> str:
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    deps = srcs + [os.path.join(CSRC, "hm_internal.h"), os.path.join(ROOT, "include", "hivemall_b200.h")]
    deps += [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    if not force and not _newer(LIB, deps):
        return LIB
    cmd = [_nvcc(), *NVCC_FLAGS, "-I", os.path.join(ROOT, "include"), "-o", LIB, *srcs]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stdout + res.stderr)
    return LIB


def build_oracle(force: bool = False) -> str:
    deps = [os.path.join(ORACLE_DIR, "hm_oracle.c"), os.path.join(ORACLE_DIR, "hm_oracle.h")]
    if not force and not _newer(ORACLE_LIB, deps):
        return ORACLE_LIB
    res = subprocess.run(["make", "-C", ORACLE_DIR, "-B"], capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + res.stdout + res.stderr)
    return ORACLE_LIB


if __name__ == "__main__":
    force = "--force" in sys.argv
    print(build_library(force=force, verbose="-v" in sys.argv))
    print(build_oracle(force=force))
