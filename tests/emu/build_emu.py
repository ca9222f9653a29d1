"""tests/emu/build_emu.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Compiles the product's CUDA sources (acoss_b200/csrc/api.cu and the kernels it includes)
with g++ against tests/emu/cuda_emu.h into tests/emu/_build/libacoss_emu.so, so the kernel
logic can be stepped on a machine without a GPU.  Only tests/test_emu_*.py load the result.
"""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "acoss_b200", "csrc")
OUT = os.path.join(HERE, "_build", "libacoss_emu.so")


def build(force: bool = False) -> str:
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if not f.startswith("_")] + [os.path.join(HERE, "cuda_emu.h"),
                                                                os.path.join(ROOT, "include", "acoss_b200.h")]
    if not force and os.path.exists(OUT) and all(os.path.getmtime(s) <= os.path.getmtime(OUT) for s in srcs):
        return OUT
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    cxx = shutil.which("g++")
    if cxx is None:
        raise RuntimeError("g++ not found")
    units = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))
    cmd = [cxx, "-std=c++17", "-O1", "-g", "-shared", "-fPIC", "-x", "c++", "-DACOSS_EMU", "-I", HERE, "-I", CSRC] + units + [
        "-o", OUT]
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build(force=True))
