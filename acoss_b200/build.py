"""Build libacoss_b200.so in-tree with nvcc for sm_100a:  python -m acoss_b200.build

nvcc cross-compiles without a GPU, so this runs in the build container; the resulting
acoss_b200/lib/libacoss_b200.so is git-ignored but travels to the GPU box with the snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")
OUT = os.path.join(HERE, "lib", "libacoss_b200.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC"]


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC)) + [os.path.join(INCLUDE, "acoss_b200.h")]


def up_to_date() -> bool:
    return os.path.exists(OUT) and all(os.path.getmtime(s) <= os.path.getmtime(OUT) for s in sources())


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and up_to_date():
        return OUT
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        if os.path.exists(OUT):
            return OUT  # GPU box without a toolkit: use the prebuilt library that travelled
        raise RuntimeError("nvcc not found and no prebuilt %s" % OUT)
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    defines = ["-D" + d for d in os.environ.get("ACOSS_NVCC_DEFINES", "").split(",") if d]  # e.g. ACOSS_S6_PROFILE
    cmd = [nvcc] + NVCC_FLAGS + defines + (["-Xptxas", "-v"] if verbose else []) + ["-o", OUT, os.path.join(CSRC, "api.cu")]
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
