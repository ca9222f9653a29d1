import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, bench
from acoss_b200 import _native
songs, _ = bench.make_workload(2000)
ctx = _native.Context(0); ctx.upload_songs(songs)
ctx.set_option("debug_flags", 1 << 30)
for cpl in (16, 8, 4):
    ctx.set_option("dp_max_cpl", cpl)
    for s in range(3):
        ctx.serra09_pairs(bench.step_pairs(2000, 50000, s, 0, 1))
    print('dp_max_cpl', cpl, ctx.last_stage_ms()['dp'])
