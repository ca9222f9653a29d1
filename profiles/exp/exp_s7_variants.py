"""Stage times of the default sweep under debug-flag variants (bench workload, 20 000-pair batches).
usage: python profiles/exp/exp_s7_variants.py [pairs] [flagset ...]   (each flagset an integer for "debug_flags")"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from acoss_b200 import _native, synth  # noqa: E402
import bench  # noqa: E402

n_pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
flagsets = [int(x, 0) for x in sys.argv[2:]] or [0]
songs, _ = synth.make_song_set("whatisacover", seed=10000, n_songs=10000)
ctx = _native.Context(0)
ctx.upload_songs(songs)
out = (ctypes.c_uint64 * 16)()
ref = None
if os.environ.get("ACOSS_GAP"):
    ctx.set_option("gap_units", int(os.environ["ACOSS_GAP"]))
    print("gap_units", os.environ["ACOSS_GAP"])
for fl in flagsets:
    ctx.set_option("debug_flags", fl)
    for step in range(3):
        pairs = bench.step_pairs(10000, n_pairs, step, 0, 1)
        if os.environ.get("ACOSS_MAXLEN"):   # drop the pairs with a song longer than this (cost of the long-song kernels)
            Ls = np.array([sg["chroma"].shape[1] for sg in songs])
            keep = (Ls[pairs[:, 0]] <= int(os.environ["ACOSS_MAXLEN"])) & (Ls[pairs[:, 1]] <= int(os.environ["ACOSS_MAXLEN"]))
            if step == 2:
                M = Ls[pairs] - 8
                print("kept %.4f of the pairs, %.4f of the cells" % (keep.mean(), (M[keep, 0] * M[keep, 1]).sum() / (M[:, 0] * M[:, 1]).sum()))
            pairs = pairs[keep]
        ctx.lib.acoss_debug_sweep7_stats(ctx._h, out, 1)
        r = ctx.serra09_pairs(pairs)
        ctx.lib.acoss_debug_sweep7_stats(ctx._h, out, 0)
        if step == 2:
            if ref is None:
                ref = r
            same = bool(np.array_equal(r["qmax_raw"], ref["qmax_raw"]) and np.array_equal(r["dmax_raw"], ref["dmax_raw"]))
            v = [int(out[i]) for i in range(16)]
            tot = max(sum(v[4:]), 1)
            names = ["setup", "stage1", "stage2", "period-barrier", "strip-init+barrier", "selection", "post-sel-barrier",
                     "export/ballots", "fixups", "copy-out", "last-barrier", "item-fetch"]
            ms = ctx.last_stage_ms()
            print("flags 0x%x: %s same_scores=%s | warp cycles %%: %s" % (
                fl, {k: round(x, 2) for k, x in ms.items()}, same, " ".join("%s %.1f" % (n, 100.0 * c / tot) for n, c in zip(names, v[4:]))),
                flush=True)
