"""tests/wide_oracle.py -- TEST INFRASTRUCTURE.  The oracle (oracle/crp_oracle.py + the DP checker) over many pairs on
all host cores, for the wide parity gate of SURVEY.md 8(d).  Workers are spawned (not forked: the test process owns a
CUDA context) and rebuild the seeded song set themselves from its generator arguments."""
from __future__ import annotations

import multiprocessing as mp
import os

import numpy as np

_W = {}


def build_songs(spec):
    """spec = ("set", shape, kwargs) | ("ranges", seed, [(lo, hi, count), ...]) | ("mfcc", seed, n, median)."""
    from acoss_b200 import synth
    kind = spec[0]
    if kind == "set":
        return synth.make_song_set(spec[1], **spec[2])[0]
    if kind == "ranges":
        songs = []
        for k, (lo, hi, cnt) in enumerate(spec[2]):
            songs += synth.make_song_set("covers80", seed=spec[1] + k, n_songs=cnt, length_range=(lo, hi))[0]
        return songs
    if kind == "mfcc":
        rng = np.random.default_rng(spec[1])
        songs = []
        for L in synth.draw_lengths(rng, spec[2], median=spec[3]):
            fr = rng.normal(size=(int(L), 13)).cumsum(axis=0).astype(np.float32) * 0.1 + rng.normal(size=(int(L), 13)).astype(np.float32)
            songs.append({"chroma": np.ascontiguousarray(fr.T), "gchroma": None})
        return songs
    raise ValueError(kind)


def _init(spec):
    os.environ["OPENBLAS_NUM_THREADS"] = "1"
    _W["songs"] = build_songs(spec)


def _chunk(args):
    from oracle import crp_oracle, dp
    chunk, m, kappa, do_oti, nonmutual = args
    songs = _W["songs"]
    out = []
    for i, j in chunk:
        Si, Sj = songs[int(i)], songs[int(j)]
        oti, B, S = crp_oracle.crp_pair(Si["chroma"].astype(np.float64), Sj["chroma"].astype(np.float64),
                                        Si["gchroma"], Sj["gchroma"], m, kappa, do_oti=do_oti, want_S=True)
        margin = min(crp_oracle.tie_margins(S, kappa))
        if nonmutual:
            B = crp_oracle.knn_mask(S, kappa)
            q, d = dp.qmax_only(B), dp.dmax_clean(B)
            dc = d
        else:
            q, d = dp.qmax_dmax_serra09(B)
            dc = dp.dmax_clean(B)
        out.append((oti, B.shape, np.packbits(B, axis=1).tobytes(), q, d, dc, margin))
    return out


class OraclePool:
    def __init__(self, spec, procs=None):
        self.procs = procs or os.cpu_count() or 1
        self.pool = mp.get_context("spawn").Pool(self.procs, initializer=_init, initargs=(spec,))

    def run(self, pairs, m=9, kappa=0.095, do_oti=True, nonmutual=False):
        pairs = np.asarray(pairs).reshape(-1, 2)
        chunks = [c for c in np.array_split(pairs, max(1, min(len(pairs), 8 * self.procs))) if len(c)]
        res = self.pool.map(_chunk, [(c, m, kappa, do_oti, nonmutual) for c in chunks], chunksize=1)
        return [r for chunk in res for r in chunk]

    def close(self):
        self.pool.close()
        self.pool.join()


def compare(ctx, songs, pairs, oracle_rows, m=9, kappa=0.095, flags=0, batch=1000, tie_margin=1e-9):
    """GPU (through the C ABI, CRPs included) against the oracle rows of OraclePool.run.  Pairs whose k-th / (k+1)-th
    neighbour distances are closer than tie_margin (relative) are "tied": the reference's own choice depends on its
    float64 rounding noise there, so they are compared but only counted.  Returns (#pairs, #tied, #tied that differ)."""
    from acoss_b200 import _native
    pairs = np.asarray(pairs, np.int32).reshape(-1, 2)
    ctx.upload_songs(songs)
    tied = tied_diff = 0
    clean = bool(flags & _native.FLAG_DMAX_CLEAN)
    for b0 in range(0, len(pairs), batch):
        sub = pairs[b0:b0 + batch]
        r = ctx.serra09_pairs(sub, m=m, kappa=kappa, flags=flags, want_crp=True)
        for p in range(len(sub)):
            oti, shape, packed, q, d, dc, margin = oracle_rows[b0 + p]
            where = "pair %d = (%d, %d) shape %s flags %d" % (b0 + p, sub[p, 0], sub[p, 1], shape, flags)
            assert r["oti"][p] == oti, where
            assert tuple(r["shape"][p]) == tuple(shape), where
            ref = np.unpackbits(np.frombuffer(packed, np.uint8).reshape(shape[0], -1), axis=1)[:, :shape[1]]
            if not (margin > tie_margin):
                tied += 1
                if not np.array_equal(r["crp"][p], ref):
                    tied_diff += 1
                continue
            assert np.array_equal(r["crp"][p], ref), where + ": %d CRP bits differ" % int((r["crp"][p] != ref).sum())
            assert r["qmax_raw"][p] == q, where + " qmax %r vs %r" % (r["qmax_raw"][p], q)
            assert r["dmax_raw"][p] == (dc if clean else d), where + " dmax %r vs %r" % (r["dmax_raw"][p], dc if clean else d)
    return len(pairs), tied, tied_diff
