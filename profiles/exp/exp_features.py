"""Throughput of acoss_aggregate_songs (x40 block median + global chroma) against the numpy restatement
of Serra09.load_features on the host: 2000 songs of ~18 000 full-rate frames (3.5 min at 86 fps), float32."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from acoss_b200 import _native
from oracle import features_oracle as fo
rng = np.random.default_rng(0)
S = 2000
songs = [rng.random((int(n), 12), dtype=np.float32) for n in rng.integers(12000, 24000, S)]
frames = sum(a.shape[0] for a in songs)
ctx = _native.Context(0)
ctx.aggregate_songs(songs[:400], want_gchroma=True)   # warm-up: allocates the page-locked staging buffers
t0 = time.perf_counter()
agg, g = ctx.aggregate_songs(songs, want_gchroma=True)
t1 = time.perf_counter()
kms, kbytes = ctx.last_aggregate_stats()
n_cpu = 100
t2 = time.perf_counter()
ref = [fo.load_features(a) for a in songs[:n_cpu]]
t3 = time.perf_counter()
ok = all(np.array_equal(agg[i], ref[i]["chroma"]) and np.array_equal(g[i], ref[i]["gchroma"]) for i in range(n_cpu))
print("songs %d, %.1f M full-rate frames (%.2f GB): GPU path incl. host packing + PCIe %.3f s = %.0f songs/s; "
      "numpy on one core %.1f songs/s; bit-identical on the first %d songs: %s"
      % (S, frames / 1e6, frames * 48 / 1e9, t1 - t0, S / (t1 - t0), n_cpu / (t3 - t2), n_cpu, ok))
print("k_feat_aggregate: %.3f ms for %.2f GB read + written = %.0f GB/s (CUDA events around the launches)"
      % (kms, kbytes / 1e9, kbytes / 1e9 / (kms / 1e3)))
