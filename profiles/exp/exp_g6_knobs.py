"""gen 5 vs gen 6 on one launch bucket (songs of 385..504 frames), with and without selection (debug bit 30); bit 28 = gen 6 without the histogram selection."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, bench
from acoss_b200 import _native
lo, hi = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (385, 504)
songs, _ = bench.make_workload(3000)
songs = [s for s in songs if lo <= s["chroma"].shape[1] <= hi]
n = len(songs)
ctx = _native.Context(0); ctx.upload_songs(songs)
for gen in (5, 6):
    ctx.set_option("sweep_gen", gen)
    for dbg in ((0, 1 << 30) if gen == 5 else (0, 1 << 28, 1 << 30)):
        ctx.set_option("debug_flags", dbg)
        for s in range(2):
            r = ctx.serra09_pairs(bench.step_pairs(n, 12000, s, 0, 1))
        print('gen', gen, 'debug %x' % dbg, {k: round(v, 2) for k, v in ctx.last_stage_ms().items()}, float(r['qmax_raw'].sum()))
