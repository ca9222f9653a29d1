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
out = (ctypes.c_uint64 * 16)()
for step in range(3):
    pairs = bench.step_pairs(10000, n_pairs, step, 0, 1)
    ctx.lib.acoss_debug_sweep7_stats(ctx._h, out, 1)
    ctx.serra09_pairs(pairs)
    ctx.lib.acoss_debug_sweep7_stats(ctx._h, out, 0)
    rows, unsafe, strips, cyc, s2, sel, s1, wait = [int(out[i]) for i in range(8)]
    ms = ctx.last_stage_ms()
    tot = max(cyc, 1)
    print("step %d: %s  rows %d unsafe %d strips %d; warp cycles: stage2 %.1f%% selection %.1f%% stage1 %.1f%% barrier-wait %.1f%% "
          "rest %.1f%%; per strip and warp: total %.0f sel %.0f s1 %.0f s2 %.0f" % (
              step, {k: round(v, 2) for k, v in ms.items()}, rows, unsafe, strips, 100 * s2 / tot, 100 * sel / tot, 100 * s1 / tot,
              100 * wait / tot, 100 * (tot - s2 - sel - s1 - wait) / tot, cyc / max(strips, 1) / 12, sel / max(strips, 1) / 12,
              s1 / max(strips, 1) / 12, s2 / max(strips, 1) / 12), flush=True)
