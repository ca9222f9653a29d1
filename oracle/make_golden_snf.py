"""oracle/make_golden_snf.py -- TEST INFRASTRUCTURE.  Golden vectors for the SNF front end (row F3), produced by the
REFERENCE ITSELF in the build container: /root/reference/benchmarking/SimilarityFusion.py (imported with matplotlib,
mpl_toolkits and librosa stubbed: they are import-only there), CRPUtils.py and SequenceAlignment.c (oracle/_ref),
composed exactly as EarlySNF.similarity does (EarlySNF.py:44-57, 85-90) with (chroma, mfcc) as the two feature kinds
(the scattering SSMs need kymatio, see oracle/snf_oracle.py).  Writes tests/golden/snf_golden.npz.

    python -m oracle.make_golden_snf
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/benchmarking"


def reference_modules():
    for name in ("matplotlib", "matplotlib.pyplot", "mpl_toolkits", "mpl_toolkits.mplot3d", "librosa"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.modules["mpl_toolkits.mplot3d"].Axes3D = object
    sys.path.insert(0, REF)
    import CRPUtils
    import SimilarityFusion
    return CRPUtils, SimilarityFusion


def synth_pair_songs(seed, lengths):
    """Small songs with both feature kinds: chroma from the bench generator, mfcc a smoothed random walk."""
    sys.path.insert(0, ROOT)
    from acoss_b200 import synth
    rng = np.random.default_rng(seed)
    songs = []
    for k, L in enumerate(lengths):
        s = synth.make_song_set("covers80", seed=seed + k, n_songs=1, fixed_length=int(L))[0][0]
        fr = rng.normal(size=(int(L), 13)).cumsum(axis=0).astype(np.float32) * 0.1 + rng.normal(size=(int(L), 13)).astype(np.float32)
        s["mfcc"] = np.ascontiguousarray(fr.T)
        songs.append(s)
    return songs


def main():
    CRPUtils, SF = reference_modules()
    from oracle import dp
    lib = dp.ref_lib()
    assert lib is not None, "oracle/_ref must be built (python -m oracle.build)"
    out = {}
    cases = [(11, (40, 52), 9, 0.095), (12, (64, 47), 9, 0.095), (13, (90, 75), 5, 0.1)]
    out["cases"] = np.array(cases, dtype=object).astype(str) if False else np.array([(c[0], c[1][0], c[1][1], c[2]) for c in cases], np.int64)
    out["kappas"] = np.array([c[3] for c in cases])
    for t, (seed, lengths, m, kappa) in enumerate(cases):
        songs = synth_pair_songs(seed, lengths)
        Si, Sj = songs
        ci, cj = Si["chroma"].astype(np.float64), Sj["chroma"].astype(np.float64)
        mi, mj = Si["mfcc"].astype(np.float64), Sj["mfcc"].astype(np.float64)
        Ws = []
        oti = CRPUtils.get_oti(Si["gchroma"], Sj["gchroma"])
        C1 = np.roll(ci, oti, axis=0)
        csm = CRPUtils.sliding_csm(CRPUtils.get_csm(C1.T, cj.T), m)
        M, N = csm.shape
        K = int(kappa * (M + N))
        ssma = CRPUtils.sliding_csm(CRPUtils.get_csm(C1.T, C1.T), m)
        ssmb = CRPUtils.sliding_csm(CRPUtils.get_csm(cj.T, cj.T), m)
        Ws.append(SF.get_WCSMSSM(ssma, ssmb, csm, K))
        csm2 = CRPUtils.sliding_csm(CRPUtils.get_csm(mi.T, mj.T), m)
        ssma2 = CRPUtils.sliding_csm(CRPUtils.get_csm(mi.T, mi.T), m)
        ssmb2 = CRPUtils.sliding_csm(CRPUtils.get_csm(mj.T, mj.T), m)
        Ws.append(SF.get_WCSMSSM(ssma2, ssmb2, csm2, K))
        F = SF.snf_ws(Ws, K=K, niters=3, reg_diag=True, verbose_times=False)
        block = F[0:M, M::]
        B = CRPUtils.csm_to_binary_mutual(-block, kappa).astype(np.uint8)
        D = np.zeros(M * N, np.float32)
        q = lib.qmax(B.flatten(), D, M, N)
        d = lib.dmax(B.flatten(), D, M, N)
        out["c%d_W0_rowsum" % t] = Ws[0].sum(1)
        out["c%d_W1_rowsum" % t] = Ws[1].sum(1)
        out["c%d_fused" % t] = block
        out["c%d_crp" % t] = np.packbits(B, axis=1)
        out["c%d_scores" % t] = np.array([q, d, oti, M, N, K], np.float64)
    path = os.path.join(ROOT, "tests", "golden", "snf_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
