"""oracle/make_golden.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Generates tests/golden/serra09_golden.npz by running THE REFERENCE ITSELF in
this container: /root/reference/benchmarking/CRPUtils.py imported by path and
/root/reference/benchmarking/SequenceAlignment.c compiled in place
(oracle/_ref, see oracle/build.py), composed exactly as
benchmarking/Serra09.py:165-175 composes them (Serra09.py cannot be imported:
it needs deepdish/librosa/kymatio at module level).

The reference ships no golden vectors; these are the pin for the oracle.
Run:  python -m oracle.make_golden      (needs /root/reference)
"""
from __future__ import annotations

import importlib.util
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from acoss_b200 import synth  # noqa: E402
from oracle import crp_oracle, dp  # noqa: E402

REF_CRP = "/root/reference/benchmarking/CRPUtils.py"


def load_reference_crputils():
    spec = importlib.util.spec_from_file_location("ref_CRPUtils", REF_CRP)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def reference_pair(R, ref, Si, Sj, m, kappa):
    """benchmarking/Serra09.py:165-175, calling the reference's functions."""
    oti = R.get_oti(Si["gchroma"], Sj["gchroma"])
    C1 = np.roll(Si["chroma"].astype(np.float64), oti, axis=0)
    C2 = Sj["chroma"].astype(np.float64)
    csm = R.get_csm(C1.T, C2.T)
    csm = R.sliding_csm(csm, m)
    S = csm
    csm = R.csm_to_binary_mutual(csm, kappa)
    M, N = csm.shape
    D = np.zeros(M * N, dtype=np.float32)
    flat = np.ascontiguousarray(csm.flatten(), dtype=np.uint8)
    q = ref.qmax(flat, D, M, N)
    d = ref.dmax(flat, D, M, N)
    Dc = np.zeros(M * N, dtype=np.float32)
    d_clean = ref.dmax(flat, Dc, M, N)
    Dsw = np.zeros((M + 1) * (N + 1), dtype=np.float32)
    sw = ref.sw(flat, Dsw, M, N)
    return int(oti), csm.astype(np.uint8), S, q, d, d_clean, sw


def main():
    R = load_reference_crputils()
    ref = dp.ref_lib()
    assert ref is not None, "reference C file could not be compiled"
    out = {}
    cases = [
        # (tag, seed, n_songs, length_range, m, kappa)
        ("a", 11, 6, (40, 90), 9, 0.095),
        ("b", 12, 4, (150, 260), 9, 0.095),
        ("c", 13, 4, (30, 70), 5, 0.1),
        ("d", 14, 3, (100, 140), 20, 0.095),
        ("e", 15, 4, (11, 16), 9, 0.095),    # tiny: M,N in 3..8, k in {0,1}
    ]
    meta = []
    for tag, seed, n, lr, m, kappa in cases:
        songs, labels = synth.make_song_set("covers80", seed=seed, n_songs=n, length_range=lr)
        out["%s_lengths" % tag] = np.array([s["chroma"].shape[1] for s in songs], np.int32)
        for si, s in enumerate(songs):
            out["%s_chroma_%d" % (tag, si)] = s["chroma"]
            out["%s_gchroma_%d" % (tag, si)] = s["gchroma"]
        pi = 0
        for i in range(n):
            for j in range(i + 1, n):
                oti, B, S, q, d, dcl, sw = reference_pair(R, ref, songs[i], songs[j], m, kappa)
                rm, cm = crp_oracle.tie_margins(S, kappa)
                key = "%s_p%d" % (tag, pi)
                out[key + "_ij"] = np.array([i, j], np.int32)
                out[key + "_oti"] = np.int32(oti)
                out[key + "_crp"] = crp_oracle.pack_bits(B)
                out[key + "_shape"] = np.array(B.shape, np.int32)
                out[key + "_scores"] = np.array([q, d, dcl, sw], np.float64)
                out[key + "_tiemargin"] = np.array([rm, cm], np.float64)
                pi += 1
        meta.append((tag, seed, n, lr[0], lr[1], m, kappa, pi))
    out["cases"] = np.array([(t, s, n, a, b, m, k, p) for (t, s, n, a, b, m, k, p) in meta],
                            dtype=[("tag", "U4"), ("seed", "i4"), ("n", "i4"), ("lo", "i4"), ("hi", "i4"),
                                   ("m", "i4"), ("kappa", "f8"), ("npairs", "i4")])
    # bare DP known answers on random binary matrices (incl. degenerate sizes)
    rng = np.random.default_rng(2024)
    shapes = [(1, 5), (2, 2), (3, 3), (4, 4), (3, 40), (40, 3), (4, 9), (17, 33), (64, 64), (65, 31),
              (100, 257), (300, 280)]
    out["dp_shapes"] = np.array(shapes, np.int32)
    for t, (M, N) in enumerate(shapes):
        dens = [0.06, 0.3][t % 2]
        B = (rng.uniform(size=(M, N)) < dens).astype(np.uint8)
        # add some diagonal structure
        for _ in range(3):
            i0, j0 = rng.integers(0, M), rng.integers(0, N)
            ln = int(rng.integers(1, max(2, min(M, N))))
            ii = np.arange(ln)
            ok = (i0 + ii < M) & (j0 + ii < N)
            B[i0 + ii[ok], j0 + ii[ok]] = 1
        flat = np.ascontiguousarray(B.flatten())
        D = np.zeros(M * N, np.float32)
        q = ref.qmax(flat, D, M, N)
        d = ref.dmax(flat, D, M, N)
        Dc = np.zeros(M * N, np.float32)
        dcl = ref.dmax(flat, Dc, M, N)
        Dsw = np.zeros((M + 1) * (N + 1), np.float32)
        sw = ref.sw(flat, Dsw, M, N)
        out["dp_%d_crp" % t] = crp_oracle.pack_bits(B)
        out["dp_%d_scores" % t] = np.array([q, d, dcl, sw], np.float64)
    path = os.path.join(ROOT, "tests", "golden", "serra09_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
