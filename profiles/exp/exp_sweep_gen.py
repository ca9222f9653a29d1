import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, bench
from acoss_b200 import _native
songs, _ = bench.make_workload(2000)
ctx = _native.Context(0); ctx.upload_songs(songs)
ref = None
for gen in (2, 3):
    for dbg in (0, 1 << 30):
        ctx.set_option("sweep_gen", gen); ctx.set_option("debug_flags", dbg)
        for s in range(3):
            r = ctx.serra09_pairs(bench.step_pairs(2000, 50000, s, 0, 1))
        if dbg == 0:
            if ref is None: ref = r
            else: print('identical to gen 2:', np.array_equal(ref['qmax_raw'], r['qmax_raw']), np.array_equal(ref['dmax_raw'], r['dmax_raw']))
        print('gen', gen, 'noselect' if dbg else 'full', {k: round(v, 1) for k, v in ctx.last_stage_ms().items()})
