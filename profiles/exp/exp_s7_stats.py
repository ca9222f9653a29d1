"""Selection statistics of the gen-7 sweep on the bench workload (library built with ACOSS_NVCC_DEFINES=ACOSS_S7_STATS).
Prints passes per row and the stage times for a batch; usage: python profiles/exp/exp_s7_stats.py [pairs] [gen]"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from acoss_b200 import _native, synth  # noqa: E402
import bench  # noqa: E402

n_pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
gen = int(sys.argv[2]) if len(sys.argv) > 2 else 7
songs, _ = synth.make_song_set("whatisacover", seed=10000, n_songs=10000)
ctx = _native.Context(0)
ctx.upload_songs(songs)
ctx.set_option("sweep_gen", gen)
out = (ctypes.c_uint64 * 8)()
for step in range(3):
    pairs = bench.step_pairs(10000, n_pairs, step, 0, 1)
    ctx.lib.acoss_debug_sweep7_stats(ctx._h, out, 1)
    ctx.serra09_pairs(pairs)
    ctx.lib.acoss_debug_sweep7_stats(ctx._h, out, 0)
    rows, cnt, full, unsafe = [int(out[i]) for i in range(4)]
    ms = ctx.last_stage_ms()
    print("step %d: %s  rows %d count passes/row %.2f full passes/row %.2f unsafe rows %d" % (
        step, {k: round(v, 2) for k, v in ms.items()}, rows, cnt / max(rows, 1), full / max(rows, 1), unsafe), flush=True)
