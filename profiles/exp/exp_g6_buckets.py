"""gen 5 vs gen 6 per launch bucket of gen 6.  debug bits: 27 = first threshold from the separator
history instead of the order-statistic estimate, 28 = histogram selection on, 30 = no selection."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, bench
from acoss_b200 import _native
allsongs, _ = bench.make_workload(6000)
ctx = _native.Context(0)
for lo, hi in ((60, 119), (120, 247), (248, 375), (376, 503), (504, 631)):
    songs = [s for s in allsongs if lo <= s["chroma"].shape[1] <= hi][:900]
    n = len(songs)
    if n < 20:
        continue
    ctx.upload_songs(songs)
    for gen, dbg in ((5, 0), (6, 0), (6, 1 << 27), (6, 1 << 28), (6, 1 << 30)):
        ctx.set_option("sweep_gen", gen)
        ctx.set_option("debug_flags", dbg)
        for s in range(2):
            r = ctx.serra09_pairs(bench.step_pairs(n, 12000, s, 0, 1))
        ms = ctx.last_stage_ms()
        print('%4d..%4d n=%4d gen %d dbg %8x thresholds %6.2f crp %6.2f dp %5.2f  sum %.1f' % (
            lo, hi, n, gen, dbg, ms['thresholds'], ms['crp'], ms['dp'], float(r['qmax_raw'].sum())))
