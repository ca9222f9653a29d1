"""Tiny workload for compute-sanitizer (racecheck / memcheck): a few pairs through every launch bucket of the
default sweep (gen 7: the eight column capacities 128 .. 1024; gens 4 / 2 and the tiled path for longer songs), the
DP, and the static-scratch DP entry."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from acoss_b200 import _native, synth
songs = []
for lo, hi, n in ((60, 110, 2), (130, 240, 2), (260, 370, 2), (380, 500, 2), (520, 630, 2), (640, 750, 2), (770, 880, 2), (900, 1010, 2),
                  (1030, 1300, 2), (1330, 1400, 1), (1420, 1500, 1)):
    songs += synth.make_song_set("covers80", seed=lo, n_songs=n, length_range=(lo, hi))[0]
rng = np.random.default_rng(1)
pairs = np.array([(i, j) for i in range(len(songs)) for j in range(i + 1, len(songs))], np.int32)
pairs = pairs[rng.permutation(len(pairs))[:60]]
ctx = _native.Context(0)
ctx.upload_songs(songs)
for a in sys.argv[1:]:
    k, v = a.split("=")
    ctx.set_option(k, int(v))
r = ctx.serra09_pairs(pairs)
print(len(pairs), "pairs", float(r["qmax_raw"].sum()), float(r["dmax_raw"].sum()))
