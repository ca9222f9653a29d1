"""Selection statistics of the default sweep (library built with ACOSS_NVCC_DEFINES=ACOSS_S7_STATS; run with
ACOSS_S4_SEL_STATS=1).  usage: ACOSS_S4_SEL_STATS=1 python profiles/exp/exp_s4_selstats.py [pairs]"""
import ctypes, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from acoss_b200 import _native, synth  # noqa: E402
import bench  # noqa: E402
n_pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
songs, _ = synth.make_song_set("whatisacover", seed=10000, n_songs=10000)
ctx = _native.Context(0)
ctx.upload_songs(songs)
out = (ctypes.c_uint64 * 16)()
pairs = bench.step_pairs(10000, n_pairs, 0, 0, 1)
ctx.serra09_pairs(pairs)
ctx.lib.acoss_debug_sweep7_stats(ctx._h, out, 1)
ctx.serra09_pairs(pairs)
ctx.lib.acoss_debug_sweep7_stats(ctx._h, out, 0)
v = [int(out[i]) for i in range(16)]
calls = max(v[4], 1)
print("rows %d unsafe %d | selections %d: passes over the keys/row %.2f (fused + counting) candidate-pop rounds/row %.2f failed "
      "pops/row %.3f boundary passes/row %.3f fallbacks to the bracketing search/row %.4f estimator calls/row %.3f "
      "no-guess starts/row %.4f" % (
          v[0], v[1], v[4], v[5] / calls, v[6] / calls, v[7] / calls, v[8] / calls, v[9] / calls, v[10] / calls, v[11] / calls))
