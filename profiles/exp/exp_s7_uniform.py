"""Per-strip cycle breakdown of k_sweep7 on a uniform workload (every song L frames, one capacity class), library built
with ACOSS_NVCC_DEFINES=ACOSS_S7_STATS.  usage: python profiles/exp/exp_s7_uniform.py L n_pairs [debug_flags ...]"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from acoss_b200 import _native, synth  # noqa: E402

L = int(sys.argv[1]) if len(sys.argv) > 1 else 480
n_pairs = int(sys.argv[2]) if len(sys.argv) > 2 else 6000
flagsets = [int(x, 0) for x in sys.argv[3:]] or [0]
songs, _ = synth.make_song_set("whatisacover", seed=1, n_songs=400, fixed_length=L)
ctx = _native.Context(0)
ctx.upload_songs(songs)
rng = np.random.default_rng(0)
pairs = rng.integers(0, 400, size=(n_pairs + 200, 2)).astype(np.int32)
pairs = pairs[pairs[:, 0] != pairs[:, 1]][:n_pairs]
out = (ctypes.c_uint64 * 16)()
cap = 128 if L + 9 <= 128 else 256 if L + 9 <= 256 else 512 if L + 9 <= 512 else 768 if L + 9 <= 768 else 1024
for fl in flagsets:
    ctx.set_option("debug_flags", fl)
    nw = {128: 4, 256: 8, 512: 8 if not (fl & (1 << 20)) else 16, 768: 12, 1024: 16}[cap]
    for step in range(2):
        ctx.lib.acoss_debug_sweep7_stats(ctx._h, out, 1)
        ctx.serra09_pairs(pairs)
        ctx.lib.acoss_debug_sweep7_stats(ctx._h, out, 0)
    v = [int(out[i]) for i in range(16)]
    rows, unsafe, strips, items = v[:4]
    names = ["setup", "stage1", "stage2", "period-barrier", "strip-init+barrier", "selection", "post-sel-barrier", "export/ballots",
             "fixups", "copy-out", "last-barrier", "item-fetch"]
    d = max(strips, 1) * nw
    ms = ctx.last_stage_ms()
    print("L %d flags 0x%x: %s | items %d strips %d rows/strip %.1f | cycles per strip (per warp): total %.0f | %s" % (
        L, fl, {k: round(x, 2) for k, x in ms.items()}, items, strips, rows / max(strips, 1), sum(v[4:]) / d,
        " ".join("%s %.0f" % (n, c / d) for n, c in zip(names, v[4:]))), flush=True)
