"""CPU: the SNF oracle (oracle/snf_oracle.py) against golden vectors produced by the reference's own SimilarityFusion.py
(oracle/make_golden_snf.py)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

from oracle import snf_oracle  # noqa: E402
from oracle.make_golden_snf import synth_pair_songs  # noqa: E402


def test_snf_oracle_matches_reference_golden():
    G = np.load(os.path.join(ROOT, "tests", "golden", "snf_golden.npz"))
    for t, (seed, La, Lb, m) in enumerate(G["cases"]):
        kappa = float(G["kappas"][t])
        Si, Sj = synth_pair_songs(int(seed), (int(La), int(Lb)))
        r = snf_oracle.early_snf_pair(Si["chroma"].astype(np.float64), Sj["chroma"].astype(np.float64), Si["gchroma"], Sj["gchroma"],
                                      Si["mfcc"].astype(np.float64), Sj["mfcc"].astype(np.float64), m=int(m), kappa=kappa, niters=3)
        q, d, oti, M, N, K = G["c%d_scores" % t]
        assert (r["oti"], r["M"], r["N"], r["K"]) == (int(oti), int(M), int(N), int(K))
        assert np.allclose(r["fused"], G["c%d_fused" % t], rtol=1e-10, atol=0)
        crp = np.unpackbits(G["c%d_crp" % t], axis=1)[:, :int(N)]
        assert np.array_equal(r["crp"], crp)
        assert (r["qmax_raw"], r["dmax_raw"]) == (q, d)
