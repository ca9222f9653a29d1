"""Where a k_sweep6 period goes: per-warp cycle counters (debug flag bit 26) for one launch bucket."""
import os, sys, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, bench
from acoss_b200 import _native
lo, hi = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (376, 503)
songs, _ = bench.make_workload(6000)
songs = [s for s in songs if lo <= s["chroma"].shape[1] <= hi][:900]
n = len(songs)
ctx = _native.Context(0); ctx.upload_songs(songs)
ctx.set_option("sweep_gen", 6)
f = ctx.lib.acoss_debug_sweep6_cycles
out = (ctypes.c_uint64 * 8)()
for extra, name in ((0, "default"), (1 << 27, "estimator"), (1 << 30, "no selection")):
    ctx.set_option("debug_flags", (1 << 26) | extra)
    ctx.serra09_pairs(bench.step_pairs(n, 6000, 0, 0, 1))
    f(ctx._h, out, 1)
    ctx.serra09_pairs(bench.step_pairs(n, 6000, 1, 0, 1))
    f(ctx._h, out, 1)
    v = [int(x) for x in out]
    nw = 2 * 16  # both sweeps counted; warps per CTA
    per = max(v[4], 1)
    print("%-12s periods %d items %d | per period per warp (cycles): stage2 %.0f  selection %.0f  stage1 %.0f  barrier wait %.0f | stage ms %s"
          % (name, v[4], v[5], v[0] / per / 16, v[1] / per / 16, v[2] / per / 16, v[3] / per / 16,
             {k: round(x, 2) for k, x in ctx.last_stage_ms().items()}))
