"""gen-4 sweep with profiling knobs: bit 30 = no selection, bit 29 = no per-row barrier (racy)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, bench
from acoss_b200 import _native
songs, _ = bench.make_workload(3000)
songs = [s for s in songs if 385 <= s["chroma"].shape[1] <= 512]
n = len(songs)
ctx = _native.Context(0); ctx.upload_songs(songs)
for dbg in (0, 1 << 30, (1 << 30) | (1 << 29)):
    ctx.set_option("debug_flags", dbg)
    for s in range(2):
        ctx.serra09_pairs(bench.step_pairs(n, 12000, s, 0, 1))
    print('debug %x' % dbg, {k: round(v, 2) for k, v in ctx.last_stage_ms().items()})
