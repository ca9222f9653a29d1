"""oracle/crp_oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

numpy restatement of the cross-recurrence-plot half of the Serra09 path.
Each function names the reference lines it follows (paths relative to
/root/reference/).  dtype contract: synthetic chroma values are
float32-representable and are handed to the oracle upcast to float64, which is
the reference's default (`chroma_type='crema'`, float64).
"""
from __future__ import annotations

import itertools

import numpy as np

from . import dp as _dp


# ----------------------------------------------------------------------------
# per-song / per-pair numerics
# ----------------------------------------------------------------------------
def global_chroma(chroma_frames):
    """benchmarking/Serra09.py:24-28.  chroma_frames: (n_frames, bins)."""
    if chroma_frames.shape[1] not in (12, 24, 36):
        raise IOError("Wrong axis for the input chroma array. Expected shape '(frame_size, bin_size)'")
    tot = chroma_frames.sum(axis=0)
    return np.divide(tot, np.max(tot))


def oti(g1, g2):
    """benchmarking/CRPUtils.py:109-136.  argmax over circular shifts s of
    sum(roll(g1, s) * g2); first maximum wins (np.argmax)."""
    n = len(g1)
    agreement = np.zeros(n)
    for s in range(n):
        agreement[s] = np.sum(np.roll(g1, s) * g2)
    return int(np.argmax(agreement))


def cross_dist(X, Y):
    """benchmarking/CRPUtils.py:67-84.  X: (M, d), Y: (N, d) -> (M, N)
    Euclidean distances via the norm expansion, clamped at 0 before sqrt."""
    C = np.sum(X ** 2, 1)[:, None] + np.sum(Y ** 2, 1)[None, :] - 2 * X.dot(Y.T)
    C[C < 0] = 0
    return np.sqrt(C)


def diag_window(D, win):
    """benchmarking/CRPUtils.py:24-45.  Implicit delay embedding: for every
    diagonal, square, prefix-sum in float64, windowed difference, sqrt."""
    M = D.shape[0] - win + 1
    N = D.shape[1] - win + 1
    S = np.zeros((M, N))
    for off in range(-M + 1, N):
        sq = np.diagonal(D, off) ** 2           # input dtype, like np.diag(D, i)**2
        c = np.concatenate(([0.0], np.asarray(sq, dtype=np.float64)))
        c = np.cumsum(c)
        w = c[win:] - c[:-win]
        n = w.shape[0]
        if off >= 0:
            rows = np.arange(n)
            cols = rows + off
        else:
            cols = np.arange(n)
            rows = cols - off
        keep = (rows < M) & (cols < N)
        S[rows[keep], cols[keep]] = np.sqrt(w[keep])
    return S


def n_neighbours(kappa, n):
    """benchmarking/CRPUtils.py:188-193: int(np.round(kappa*n)) (half-to-even
    on the double product) for 0 < kappa < 1, kappa itself otherwise."""
    if kappa < 1:
        return int(np.round(kappa * n))
    return int(kappa)


def knn_mask(D, kappa):
    """benchmarking/CRPUtils.py:169-199.  Row-wise: mark the k smallest entries
    of every row (np.argpartition), k = n_neighbours(kappa, n_cols)."""
    rows, cols = D.shape
    if kappa == 0:
        return np.ones_like(D)
    k = n_neighbours(kappa, cols)
    J = np.argpartition(D, k, 1)[:, 0:k]
    B = np.zeros((rows, cols), dtype=np.uint8)
    B[np.arange(rows)[:, None], J] = 1
    return B


def mutual_mask(D, kappa):
    """benchmarking/CRPUtils.py:201-219."""
    return knn_mask(D, kappa) * (knn_mask(D.T, kappa).T)


def tie_margins(S, kappa):
    """Relative gap between the k-th and (k+1)-th smallest value of every row
    and every column of S (SURVEY.md 8c tie-freeness guard).  A pair is
    'tie-free' when both minima are > 0; tests demand a safety factor."""
    out = []
    for A in (S, S.T):
        k = n_neighbours(kappa, A.shape[1])
        if k <= 0 or k >= A.shape[1]:
            out.append(np.inf)
            continue
        P = np.partition(A, (k - 1, k), axis=1)
        gap = (P[:, k] - P[:, k - 1]) / np.maximum(P[:, k], 1e-300)
        out.append(float(gap.min()))
    return tuple(out)


# ----------------------------------------------------------------------------
# the pair pipeline
# ----------------------------------------------------------------------------
def crp_pair(chroma_i, chroma_j, g_i, g_j, m=9, kappa=0.095, do_oti=True, want_S=False):
    """benchmarking/Serra09.py:165-172 (chroma) / :177-180 (mfcc: do_oti=False).
    chroma_*: (bins, L) arrays, g_*: (bins,) global chroma.  Returns
    (oti, crp uint8 (M, N)[, S float64])."""
    shift = oti(g_i, g_j) if do_oti else 0
    C1 = np.roll(chroma_i, shift, axis=0)
    C2 = chroma_j
    csm = cross_dist(C1.T, C2.T)
    S = diag_window(csm, m)
    B = mutual_mask(S, kappa)
    if want_S:
        return shift, B, S
    return shift, B


def serra09_pair(chroma_i, chroma_j, g_i, g_j, m=9, kappa=0.095, do_oti=True, lib=None):
    """benchmarking/Serra09.py:165-175 for one pair.  Returns a dict with the
    raw DP maxima and the normalised scores (dp_max / (M+N), :174-175)."""
    shift, B = crp_pair(chroma_i, chroma_j, g_i, g_j, m, kappa, do_oti)
    M, N = B.shape
    q, d = _dp.qmax_dmax_serra09(B, lib)
    return {"oti": shift, "crp": B, "M": M, "N": N, "qmax_raw": q, "dmax_raw": d,
            "qmax": q / (M + N), "dmax": d / (M + N)}


def all_pairs(songs, m=9, kappa=0.095, lib=None):
    """benchmarking/CoverAlgorithm.py:166-183 with symmetric=True: pairs (i<j)
    in itertools.combinations order, float32 N x N matrices, `Ds += Ds.T`.
    songs: list of dicts {'chroma': (12, L), 'gchroma': (12,)}."""
    n = len(songs)
    Ds = {"chroma_qmax": np.zeros((n, n), np.float32), "chroma_dmax": np.zeros((n, n), np.float32)}
    for i, j in itertools.combinations(range(n), 2):
        r = serra09_pair(songs[i]["chroma"], songs[j]["chroma"], songs[i]["gchroma"],
                         songs[j]["gchroma"], m, kappa, True, lib)
        Ds["chroma_qmax"][i, j] = r["qmax"]
        Ds["chroma_dmax"][i, j] = r["dmax"]
    for key in Ds:
        Ds[key] += Ds[key].T
    return Ds


def pack_bits(B):
    """Pack a (M, N) 0/1 matrix row-wise into uint32 words, bit t of word w of
    row i = B[i, 32*w + t] (the layout the CUDA debug dump uses)."""
    M, N = B.shape
    W = (N + 31) // 32
    P = np.zeros((M, W * 32), dtype=np.uint8)
    P[:, :N] = B
    bits = P.reshape(M, W, 32).astype(np.uint32)
    return (bits << np.arange(32, dtype=np.uint32)[None, None, :]).sum(axis=2, dtype=np.uint64).astype(np.uint32)


def unpack_bits(P, N):
    M, W = P.shape
    bits = (P[:, :, None] >> np.arange(32, dtype=np.uint32)[None, None, :]) & 1
    return bits.reshape(M, W * 32)[:, :N].astype(np.uint8)
