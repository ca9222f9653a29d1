"""BASELINE configs[4], the long-song stress case: 4096-frame HPCP songs, m = 20, kappa = 0.095 (M = N = 4077,
1.66e7 cells per pair), and a mixed 1024..4096 set for variable-length batching.  Prints one JSON line per set:
pairs/s, cells/s and Qmax+Dmax GCUPS through Context.serra09_pairs (songs resident, pair list from the host,
scores back), the stage times of the library's own CUDA events, and the CPU oracle (numpy restatement of
CRPUtils.py + the reference's SequenceAlignment.c) on one core for a bounded sample."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from acoss_b200 import _native, synth

ctx = _native.Context(0)
m, kappa = 20, 0.095
for name, kw, n_songs in (("fixed4096", dict(fixed_length=4096), 48), ("mixed1024_4096", dict(length_range=(1024, 4096)), 64)):
    songs, _ = synth.make_song_set("covers80", seed=4096, n_songs=n_songs, **kw)
    ctx.upload_songs(songs)
    pairs = np.array([(i, j) for i in range(n_songs) for j in range(i + 1, n_songs)], np.int32)
    L = np.array([s["chroma"].shape[1] for s in songs])
    cells = float(((L[pairs[:, 0]] - m + 1) * (L[pairs[:, 1]] - m + 1)).sum())
    upd = float(((L[pairs[:, 0]] - m - 1) * (L[pairs[:, 1]] - m - 1)).sum() + ((L[pairs[:, 0]] - m - 2) * (L[pairs[:, 1]] - m - 2)).sum())
    ctx.serra09_pairs(pairs, m=m, kappa=kappa)   # warm-up at full size: scratch buffers are allocated once
    dt = 1e30
    for _ in range(2):
        t0 = time.perf_counter()
        r = ctx.serra09_pairs(pairs, m=m, kappa=kappa)
        dt = min(dt, time.perf_counter() - t0)
    st = ctx.last_stage_ms()
    out = {"config": "BASELINE configs[4] " + name, "songs": n_songs, "pairs": int(len(pairs)), "m": m, "kappa": kappa,
           "seconds": dt, "pairs_per_s": len(pairs) / dt, "cells_per_s": cells / dt, "qmax_dmax_gcups_whole_path": upd / dt / 1e9,
           "stage_ms_last_chunk": st}
    if name == "fixed4096":
        sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
        import parity_checks as pc
        t0 = time.perf_counter()
        o = pc.oracle_pair(songs, int(pairs[0, 0]), int(pairs[0, 1]), m, kappa, False)
        cpu = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": 1.0 / cpu, "unit": "pairs/s", "cores": 1, "kind": "port", "sample": "1 pair of 4096 x 4096 frames"}
        out["parity_on_sample"] = bool(r["qmax_raw"][0] == o["qmax_raw"] and r["dmax_raw"][0] == o["dmax_raw"])
    print(json.dumps(out), flush=True)
