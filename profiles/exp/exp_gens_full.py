"""All sweep generations on the full-length workload (2000 songs, 50000 pairs per call)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, bench
from acoss_b200 import _native
songs, _ = bench.make_workload(2000)
ctx = _native.Context(0); ctx.upload_songs(songs)
ref = None
for gen in (2, 4, 5):
    ctx.set_option("sweep_gen", gen)
    for s in range(3):
        r = ctx.serra09_pairs(bench.step_pairs(2000, 50000, s, 0, 1))
    if ref is None: ref = r
    same = np.array_equal(ref['qmax_raw'], r['qmax_raw']) and np.array_equal(ref['dmax_raw'], r['dmax_raw'])
    print('gen', gen, {k: round(v, 1) for k, v in ctx.last_stage_ms().items()}, 'same as gen 2:', same)
