"""Seeded synthetic HPCP-like song sets shaped like the reference's datasets.

Shapes follow /root/reference/subsets/*.json: covers80 = 80
cliques x 2, whatisacover = 5000 x 2, benchmark = 1000 x 13 + 2000 x 1.  A song
is the output of Serra09.load_features (benchmarking/Serra09.py:96-156) for
the chroma path: 'chroma' (12, L) after the x40 median aggregation and
'gchroma' (12,) from global_chroma (Serra09.py:24-28).

Values are float32 (the hpcp dtype, preprocess/features.py:209-275) and are
jittered so that no two distances tie (the reference's tie order is
implementation-defined).
"""
from __future__ import annotations

import numpy as np

SHAPES = {
    # name: (n_cliques_multi, versions_per_clique, n_singletons)
    "covers80": (80, 2, 0),
    "whatisacover": (5000, 2, 0),
    "benchmark": (1000, 13, 2000),
}


def _base_sequence(rng, length):
    """Smoothed random walk over 12 pitch classes, max-normalised per frame."""
    steps = rng.normal(0.0, 1.0, size=(length + 16, 12))
    walk = np.cumsum(steps, axis=0)
    kernel = np.hanning(9)
    kernel /= kernel.sum()
    sm = np.stack([np.convolve(walk[:, b], kernel, mode="same") for b in range(12)], axis=1)
    sm = sm[8:8 + length]
    sm = sm - sm.min(axis=1, keepdims=True)
    # sharpen towards a few dominant pitch classes like an HPCP frame
    sm = sm ** 2
    sm /= np.maximum(sm.max(axis=1, keepdims=True), 1e-9)
    return sm


def _version_of(rng, base, length):
    """Cover version: pitch shift, +-15 % time warp by resampling, noise, jitter."""
    shift = int(rng.integers(0, 12))
    pos = np.linspace(0.0, base.shape[0] - 1.0, length)
    lo = np.floor(pos).astype(int)
    hi = np.minimum(lo + 1, base.shape[0] - 1)
    frac = (pos - lo)[:, None]
    v = (1.0 - frac) * base[lo] + frac * base[hi]
    v = np.roll(v, shift, axis=1)
    v = v + rng.uniform(0.0, 0.1, size=v.shape)
    v /= np.maximum(v.max(axis=1, keepdims=True), 1e-9)
    v = v + 1e-3 * rng.uniform(0.0, 1.0, size=v.shape)
    return v.astype(np.float32)


def draw_lengths(rng, n, median=450.0, sigma=0.35, lo=150, hi=1500):
    L = np.rint(rng.lognormal(np.log(median), sigma, size=n))
    return np.clip(L, lo, hi).astype(np.int64)


def global_chroma_f64(frames):
    """benchmarking/Serra09.py:24-28 on (L, 12) frames, returned as float64."""
    tot = frames.astype(np.float64).sum(axis=0)
    return tot / tot.max()


def make_song_set(shape="covers80", seed=80, n_songs=None, fixed_length=None,
                  length_range=None, median=450.0):
    """Returns (songs, labels): songs[i] = {'chroma': (12, L) float32,
    'gchroma': (12,) float64}, labels[i] = clique id string.

    n_songs truncates/limits the set (cliques are kept whole where possible);
    fixed_length forces every song to that many frames (stress config);
    length_range=(lo, hi) draws uniform lengths instead of the lognormal.
    """
    rng = np.random.default_rng(seed)
    multi, per, single = SHAPES[shape]
    clique_sizes = [per] * multi + [1] * single
    if n_songs is not None:
        out, tot = [], 0
        for c in clique_sizes:
            if tot >= n_songs:
                break
            c = min(c, n_songs - tot)
            out.append(c)
            tot += c
        clique_sizes = out
    songs, labels = [], []
    for cid, csz in enumerate(clique_sizes):
        if fixed_length is not None:
            lens = np.full(csz, fixed_length, dtype=np.int64)
        elif length_range is not None:
            lens = rng.integers(length_range[0], length_range[1] + 1, size=csz)
        else:
            lens = draw_lengths(rng, csz, median=median)
        base_len = int(max(lens.max(), 32) * 1.05) + 8
        base = _base_sequence(rng, base_len)
        for v in range(csz):
            L = int(lens[v])
            warp = rng.uniform(0.85, 1.0)
            src = base[: max(int(base_len * warp), 16)]
            frames = _version_of(rng, src, L)          # (L, 12) float32
            songs.append({"chroma": np.ascontiguousarray(frames.T),
                          "gchroma": global_chroma_f64(frames)})
            labels.append("clique_%05d" % cid)
    # shuffle so cliques are not contiguous (sorted glob order in the reference
    # is by file name, unrelated to clique)
    perm = rng.permutation(len(songs))
    songs = [songs[p] for p in perm]
    labels = [labels[p] for p in perm]
    return songs, labels


def pack_arena(songs):
    """Frames-major packing for the C ABI: frames (total, 12) float32,
    offsets int64 (n,), lengths int32 (n,), gchroma float64 (n, 12)."""
    lengths = np.array([s["chroma"].shape[1] for s in songs], dtype=np.int32)
    offsets = np.zeros(len(songs), dtype=np.int64)
    if len(songs) > 1:
        offsets[1:] = np.cumsum(lengths[:-1].astype(np.int64))
    frames = np.concatenate([np.ascontiguousarray(s["chroma"].T, dtype=np.float32) for s in songs], axis=0)
    g = np.stack([np.asarray(s["gchroma"], dtype=np.float64) for s in songs], axis=0)
    return np.ascontiguousarray(frames), offsets, lengths, np.ascontiguousarray(g)


def dense_affinity_matrix(rng, M, N):
    """Config 4 (EarlySNF.py:86): a dense (M, N) float32 'distance' = -W with W
    a smoothed random affinity in (0, 1], tie-free by construction."""
    W = rng.uniform(0.0, 1.0, size=(M + 6, N + 6))
    k = np.hanning(7)
    k /= k.sum()
    W = np.apply_along_axis(lambda r: np.convolve(r, k, mode="valid"), 1, W)
    W = np.apply_along_axis(lambda c: np.convolve(c, k, mode="valid"), 0, W)
    diag = np.exp(-0.5 * ((np.arange(M)[:, None] * (N / max(M, 1)) - np.arange(N)[None, :]) / 6.0) ** 2)
    W = 0.6 * W + 0.4 * diag * rng.uniform(0.5, 1.0, size=(M, N))
    W = W + 1e-4 * rng.uniform(0.0, 1.0, size=W.shape)
    W = W / W.max()
    return (-W).astype(np.float32)
