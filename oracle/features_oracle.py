"""oracle/features_oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatement of the feature preparation in Serra09.load_features (benchmarking/Serra09.py:96-114)
with numpy: global_chroma (Serra09.py:24-28) and the x40 block aggregation the reference does with
librosa.util.sync(data, np.arange(0, n, fac), aggregate=np.median | np.mean) (Serra09.py:104, 110).

Parity pinning: global_chroma is the reference's own three lines and is checked against the
reference module's arithmetic by construction (np.sum(axis=0) / np.max).  librosa (pinned 0.6.3 in
requirements.txt) is NOT under /root/reference and not installed here, so the slicing convention of
util.sync -- slice k = frames [idx[k], idx[k+1]) with the last slice running to n, and the
aggregate applied along the time axis -- is restated from its published documentation:
PARITY UNPINNED for that convention; the aggregates themselves are numpy's own np.median/np.mean.
Only tests/ and bench.py's cpu_baseline may import this module.
"""
import numpy as np


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
