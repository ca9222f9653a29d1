"""Tiny workload for compute-sanitizer (racecheck / memcheck): a few pairs through every launch bucket of the
default sweep (gen 6: 128 / 256 / 512 / 640 threads, gen 5, gen 4) and the DP."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from acoss_b200 import _native, synth
songs = []
for lo, hi, n in ((60, 110, 3), (130, 240, 3), (300, 500, 3), (520, 630, 3), (700, 900, 2), (1000, 1300, 2)):
    songs += synth.make_song_set("covers80", seed=lo, n_songs=n, length_range=(lo, hi))[0]
rng = np.random.default_rng(1)
pairs = np.array([(i, j) for i in range(len(songs)) for j in range(i + 1, len(songs))], np.int32)
pairs = pairs[rng.permutation(len(pairs))[:40]]
ctx = _native.Context(0)
ctx.upload_songs(songs)
for a in sys.argv[1:]:
    k, v = a.split("=")
    ctx.set_option(k, int(v))
r = ctx.serra09_pairs(pairs)
print(len(pairs), "pairs", float(r["qmax_raw"].sum()), float(r["dmax_raw"].sum()))
