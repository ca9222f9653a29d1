import sys, time
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, bench
from acoss_b200 import _native
songs,_=bench.make_workload(2000)
ctx=_native.Context(0); ctx.upload_songs(songs)
for dbg in (0, 1<<30):
    ctx.set_option("debug_flags", dbg)
    for s in range(3):
        p=bench.step_pairs(2000,50000,s,0,1)
        ctx.serra09_pairs(p)
    print('debug',dbg, ctx.last_stage_ms())
