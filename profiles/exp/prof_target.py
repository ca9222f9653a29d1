"""Small fixed workload for ncu captures: songs of 385..512 frames only (one launch bucket,
128 threads), one serra09_pairs call -> one MODE 0 sweep, one MODE 2 sweep, one DP launch."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, bench
from acoss_b200 import _native
n_pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 12000
songs, _ = bench.make_workload(3000)
songs = [s for s in songs if 385 <= s["chroma"].shape[1] <= 512]
n = len(songs)
ctx = _native.Context(0); ctx.upload_songs(songs)
for a in sys.argv[2:]:
    k, v = a.split("=")
    ctx.set_option(k, int(v))
r = ctx.serra09_pairs(bench.step_pairs(n, n_pairs, 0, 0, 1))
print(n, 'songs', {k: round(v, 2) for k, v in ctx.last_stage_ms().items()}, float(r['qmax_raw'].sum()), float(r['dmax_raw'].sum()))
