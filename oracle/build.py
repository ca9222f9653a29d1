"""oracle/build.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Builds the two CPU checkers:

* ``oracle/_build/liboracle_seqalign.so``  from ``oracle/seqalign_oracle.c``
  (our C restatement of the reference DPs);
* ``oracle/_ref/libseqalign_ref.so``       from the reference's OWN source file
  ``/root/reference/benchmarking/SequenceAlignment.c`` compiled where it lies
  (never copied into this repo) with the flag the reference's setup.py uses
  (``-Ofast``, benchmarking/setup.py:45).  Only possible where /root/reference
  exists; on the GPU box the prebuilt file travels with the snapshot.

Both directories are git-ignored.  Run ``python -m oracle.build``.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
BUILD_DIR = os.path.join(HERE, "_build")
REF_DIR = os.path.join(HERE, "_ref")
ORACLE_SO = os.path.join(BUILD_DIR, "liboracle_seqalign.so")
REF_SO = os.path.join(REF_DIR, "libseqalign_ref.so")
REF_SRC = "/root/reference/benchmarking/SequenceAlignment.c"


def _newer(src: str, dst: str) -> bool:
    return (not os.path.exists(dst)) or os.path.getmtime(src) > os.path.getmtime(dst)


def build_oracle(force: bool = False) -> str:
    os.makedirs(BUILD_DIR, exist_ok=True)
    src = os.path.join(HERE, "seqalign_oracle.c")
    if force or _newer(src, ORACLE_SO):
        cc = shutil.which("gcc") or shutil.which("cc")
        if cc is None:
            raise RuntimeError("no C compiler found for the oracle")
        # strict IEEE: no -ffast-math, so the restatement is the plain recurrence
        tmp = "%s.%d.tmp" % (ORACLE_SO, os.getpid())   # several worker processes may build at once: publish atomically
        subprocess.check_call([cc, "-O2", "-std=c11", "-fPIC", "-shared", src, "-o", tmp])
        os.replace(tmp, ORACLE_SO)
    return ORACLE_SO


def build_ref(force: bool = False) -> str | None:
    """Compile the reference's own C file in place. Returns None when the
    reference tree is not on this machine and nothing was prebuilt."""
    if not os.path.exists(REF_SRC):
        return REF_SO if os.path.exists(REF_SO) else None
    os.makedirs(REF_DIR, exist_ok=True)
    if force or _newer(REF_SRC, REF_SO):
        cc = shutil.which("gcc") or shutil.which("cc")
        tmp = "%s.%d.tmp" % (REF_SO, os.getpid())
        subprocess.check_call([cc, "-Ofast", "-x", "c", "-w", "-fPIC", "-shared", REF_SRC, "-o", tmp])
        os.replace(tmp, REF_SO)
    return REF_SO


def main() -> int:
    print("oracle :", build_oracle(force="--force" in sys.argv))
    print("ref    :", build_ref(force="--force" in sys.argv))
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
