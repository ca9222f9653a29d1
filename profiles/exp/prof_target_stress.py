"""Small fixed workload for ncu captures of the tiled long-song sweep: 20 songs of 4096 frames, m = 20, 190 pairs,
one serra09_pairs call -> k_sweep_tile<12,3>, k_select_rows_cta<16>, k_emit_crp, k_dp."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from acoss_b200 import _native, synth
songs, _ = synth.make_song_set("covers80", seed=4096, n_songs=20, fixed_length=4096)
ctx = _native.Context(0); ctx.upload_songs(songs)
pairs = np.array([(i, j) for i in range(20) for j in range(i + 1, 20)], np.int32)
r = ctx.serra09_pairs(pairs, m=20, kappa=0.095)
print({k: round(v, 2) for k, v in ctx.last_stage_ms().items()}, float(r['qmax_raw'].sum()), float(r['dmax_raw'].sum()))
