"""oracle/features_oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatement of the feature preparation in Serra09.load_features (benchmarking/Serra09.py:96-114)
with numpy: global_chroma (Serra09.py:24-28) and the x40 block aggregation the reference does with
librosa.util.sync(data, np.arange(0, n, fac), aggregate=np.median | np.mean) (Serra09.py:104, 110).

Parity pinning: global_chroma is the reference's own three lines and is checked against the
reference module's arithmetic by construction (np.sum(axis=0) / np.max).  librosa (pinned 0.6.3 in
requirements.txt) is a third-party dependency that is NOT under /root/reference, is not installed here
and cannot be fetched (no network), so its source lines cannot be quoted; the algorithm of
librosa.util.sync in the 0.6 series is restated below step by step (sync_slices):
  sync(data, idx, aggregate, pad=True, axis=-1)
    -> slices = util.index_to_slice(np.asarray(idx), 0, data.shape[axis], pad=True)
  index_to_slice(idx, idx_min, idx_max, step=None, pad=True)
    -> idx_fixed = util.fix_frames(idx, idx_min, idx_max, pad=pad)
       [slice(start, end, step) for (start, end) in zip(idx_fixed, idx_fixed[1:])]
  fix_frames(frames, x_min=0, x_max=None, pad=True)
    -> clip frames to [x_min, x_max]; with pad, prepend x_min and x_max to the list;
       keep x_min <= frames <= x_max; return np.unique(frames)   (sorted, duplicates removed)
  and every slice is reduced with aggregate(data[..., slice], axis=axis).
With idx = np.arange(0, n, fac) (Serra09.py:104,110) the boundaries are unique([0, n, 0, fac, 2 fac, ...]) =
[0, fac, ..., fac * floor((n-1)/fac), n]: ceil(n / fac) slices, the last one shorter when fac does not divide n,
exactly n / fac slices when it does (the padded boundary n coincides with no extra slice), one slice [0, n) when
n <= fac.  `sync` (the closed form used by the kernels' tests) is checked against `sync_slices` in
tests/test_features.py for those edge cases.  PARITY UNPINNED against librosa itself: nothing here can execute it.
Only tests/ and bench.py's cpu_baseline may import this module.
"""
import numpy as np


def fix_frames(frames, x_min=0, x_max=None, pad=True):
    """librosa.util.fix_frames (0.6 series), restated."""
    frames = np.asarray(frames)
    if np.any(frames < 0):
        raise ValueError("Negative frame index detected")
    if pad and (x_min is not None or x_max is not None):
        frames = np.clip(frames, x_min, x_max)
    if pad:
        pad_data = []
        if x_min is not None:
            pad_data.append(x_min)
        if x_max is not None:
            pad_data.append(x_max)
        frames = np.concatenate((np.asarray(pad_data, dtype=frames.dtype), frames))
    if x_min is not None:
        frames = frames[frames >= x_min]
    if x_max is not None:
        frames = frames[frames <= x_max]
    return np.unique(frames).astype(int)


def sync_slices(data, idx, aggregate):
    """librosa.util.sync(data, idx, aggregate=aggregate) for a (d, n) array through the restated index_to_slice /
    fix_frames: the literal form."""
    n = data.shape[-1]
    b = fix_frames(np.asarray(idx), 0, n, pad=True)
    out = np.empty((data.shape[0], len(b) - 1), dtype=data.dtype, order="F")
    for k, (s, e) in enumerate(zip(b, b[1:])):
        out[:, k] = aggregate(data[:, s:e], axis=-1)
    return out


def global_chroma(chroma):
    """Serra09.py:24-28 on a (frames, bins) array."""
    if chroma.shape[1] not in [12, 24, 36]:
        raise IOError("Wrong axis for the input chroma array. Expected shape '(frame_size, bin_size)'")
    g = np.sum(chroma, axis=0)
    return g / np.max(g)


def sync(data, fac, aggregate):
    """librosa.util.sync(data, np.arange(0, n, fac), aggregate=aggregate) for a (d, n) array."""
    d, n = data.shape
    starts = list(range(0, n, fac))
    out = np.empty((d, len(starts)), dtype=data.dtype, order="F")
    for k, s in enumerate(starts):
        out[:, k] = aggregate(data[:, s:min(s + fac, n)], axis=1)
    return out


def load_features(chroma, mfcc=None, fac=40):
    """The chroma / mfcc part of Serra09.load_features (Serra09.py:99-114).  chroma: (n, bins) at
    full rate; mfcc: (13, n) or None."""
    g = global_chroma(chroma)
    c = sync(chroma.T, fac, np.median)
    m = None
    if mfcc is not None:
        m = np.array(mfcc)
        m[np.isnan(m)] = 0
        m[np.isinf(m)] = 0
        m = sync(m, fac, np.mean)
        N = min(c.shape[1], m.shape[1])
        c = c[:, 0:N]
        m = m[:, 0:N]
    return {"gchroma": g, "chroma": c, "mfcc": m}
