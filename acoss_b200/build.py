"""Build libacoss_b200.so in-tree with nvcc for sm_100a:  python -m acoss_b200.build

nvcc cross-compiles without a GPU, so this runs in the build container; the resulting
acoss_b200/lib/libacoss_b200.so is git-ignored but travels to the GPU box with the snapshot.
Every .cu under csrc/ is one translation unit, compiled in parallel into csrc/_obj/ and rebuilt only
when it or a header changed.
"""
from __future__ import annotations

import concurrent.futures
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "_obj")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")
OUT = os.path.join(HERE, "lib", "libacoss_b200.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC"]


def units():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def headers():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))) + [
        os.path.join(INCLUDE, "acoss_b200.h")]


def sources():
    return units() + headers()


def up_to_date() -> bool:
    return os.path.exists(OUT) and all(os.path.getmtime(s) <= os.path.getmtime(OUT) for s in sources())


def _defines():
    return ["-D" + d for d in os.environ.get("ACOSS_NVCC_DEFINES", "").split(",") if d]  # e.g. ACOSS_S7_STATS


def build(force: bool = False, verbose: bool = False) -> str:
    tag = ",".join(_defines())
    tagfile = os.path.join(OBJ, "defines.txt")
    if os.path.exists(OUT) and os.path.exists(tagfile) and open(tagfile).read() != tag:
        force = True  # built with other -D options
    if not force and up_to_date():
        return OUT
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        if os.path.exists(OUT):
            return OUT  # GPU box without a toolkit: use the prebuilt library that travelled
        raise RuntimeError("nvcc not found and no prebuilt %s" % OUT)
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    os.makedirs(OBJ, exist_ok=True)
    hdr_time = max(os.path.getmtime(h) for h in headers())
    jobs = []
    for cu in units():
        obj = os.path.join(OBJ, os.path.basename(cu)[:-3] + ".o")
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(cu), hdr_time):
            jobs.append((cu, obj))

    def compile_one(job):
        cu, obj = job
        cmd = [nvcc] + NVCC_FLAGS + _defines() + (["-Xptxas", "-v"] if verbose else []) + ["-c", cu, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        return cu, r.returncode, r.stdout + r.stderr

    failed = False
    with concurrent.futures.ThreadPoolExecutor(max_workers=max(1, len(jobs))) as ex:
        for cu, rc, log in ex.map(compile_one, jobs):
            if verbose or rc != 0:
                sys.stderr.write(log)
            if rc != 0:
                failed = True
    if failed:
        raise RuntimeError("nvcc failed")
    objs = [os.path.join(OBJ, os.path.basename(cu)[:-3] + ".o") for cu in units()]
    subprocess.check_call([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", OUT] + objs)
    open(tagfile, "w").write(tag)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
