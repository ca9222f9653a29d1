"""oracle/make_golden_ties.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Golden vectors for songs with EXACTLY repeated frames (digital silence at the song edges, a sustained frame in the
middle), produced by the reference itself (benchmarking/CRPUtils.py imported by path, SequenceAlignment.c compiled in
place) exactly like oracle/make_golden.py.  On such songs many window distances of a row / column are equal in exact
arithmetic; when that cluster straddles the k-th neighbour, np.argpartition (CRPUtils.py:191-199) still keeps exactly
k, choosing among the tied cells by the rounding noise of sliding_csm's cumsum differences (CRPUtils.py:38-43), while
the product keeps only the cells strictly below the k-th / (k+1)-th separator (deviation 1).
tests/test_ties.py measures what that costs; these vectors pin the reference side of the comparison.

Per pair: the reference's CRP, scores, and masks of the rows / columns whose k-th and (k+1)-th smallest distances are
closer than 1e-9 relative ("tied").  Run:  python -m oracle.make_golden_ties      (needs /root/reference)
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from acoss_b200 import synth  # noqa: E402
from oracle import crp_oracle, dp  # noqa: E402
from oracle.make_golden import load_reference_crputils, reference_pair  # noqa: E402

TIE_REL = 1e-9


def tied_lines(S, kappa):
    """Boolean masks (rows, cols) of the lines of S whose k-th / (k+1)-th smallest values are within TIE_REL."""
    out = []
    for A in (S, S.T):
        k = crp_oracle.n_neighbours(kappa, A.shape[1])
        if k <= 0 or k >= A.shape[1]:
            out.append(np.zeros(A.shape[0], bool))
            continue
        P = np.partition(A, (k - 1, k), axis=1)
        out.append((P[:, k] - P[:, k - 1]) <= TIE_REL * np.maximum(P[:, k], 1e-300))
    return out


def with_silence(song, head, tail):
    c = np.array(song["chroma"], dtype=np.float64)
    c[:, :head] = 0.0
    if tail:
        c[:, c.shape[1] - tail:] = 0.0
    return {"chroma": c, "gchroma": synth.global_chroma_f64(c.T)}


def with_sustain(song, at, run, frame):
    c = np.array(song["chroma"], dtype=np.float64)
    c[:, at:at + run] = frame[:, None]
    return {"chroma": c, "gchroma": synth.global_chroma_f64(c.T)}


def make_cases():
    """(tag, songs): four songs each, lengths 180..300, m = 9, kappa = 0.095 (k about 17..28)."""
    cases = []
    base, _ = synth.make_song_set("covers80", seed=41, n_songs=4, length_range=(180, 300))
    # short silences: the zero cluster is smaller than k, nothing straddles
    cases.append(("s", [with_silence(s, 5 + i, 7 + 2 * i) for i, s in enumerate(base)]))
    # long silences: rows / columns inside the silence see more than k equal distances
    cases.append(("l", [with_silence(s, 0 if i % 2 else 45, 50 + 6 * i) for i, s in enumerate(base)]))
    # one sustained frame (the same in all four songs), runs longer than k + m
    frame = np.array(base[0]["chroma"][:, 60], dtype=np.float64)
    cases.append(("r", [with_sustain(s, 70 + 9 * i, 44 + 3 * i, frame) for i, s in enumerate(base)]))
    return cases


def main():
    R = load_reference_crputils()
    ref = dp.ref_lib()
    assert ref is not None, "reference C file could not be compiled"
    m, kappa = 9, 0.095
    out = {"m": np.int32(m), "kappa": np.float64(kappa)}
    tags = []
    for tag, songs in make_cases():
        n = len(songs)
        for si, s in enumerate(songs):
            out["%s_chroma_%d" % (tag, si)] = s["chroma"]
            out["%s_gchroma_%d" % (tag, si)] = s["gchroma"]
        pi = 0
        for i in range(n):
            for j in range(i + 1, n):
                oti, B, S, q, d, dcl, sw = reference_pair(R, ref, songs[i], songs[j], m, kappa)
                rows, cols = tied_lines(S, kappa)
                key = "%s_p%d" % (tag, pi)
                out[key + "_ij"] = np.array([i, j], np.int32)
                out[key + "_oti"] = np.int32(oti)
                out[key + "_crp"] = crp_oracle.pack_bits(B)
                out[key + "_shape"] = np.array(B.shape, np.int32)
                out[key + "_scores"] = np.array([q, d, dcl, sw], np.float64)
                out[key + "_tied_rows"] = np.packbits(rows)
                out[key + "_tied_cols"] = np.packbits(cols)
                print("%s: shape %s oti %d tied rows %d cols %d  ones %d  qmax %g dmax %g" % (
                    key, B.shape, oti, rows.sum(), cols.sum(), int(B.sum()), q, d))
                pi += 1
        tags.append((tag, n, pi))
    out["cases"] = np.array(tags, dtype=[("tag", "U4"), ("n", "i4"), ("npairs", "i4")])
    path = os.path.join(ROOT, "tests", "golden", "ties_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
