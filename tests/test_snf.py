"""SNF front end (row F3): acoss_snf_pairs through the C ABI against the golden vectors produced by the reference's own
SimilarityFusion.py (tests/golden/snf_golden.npz).  CPU: the kernel sources under the SIMT emulator (smallest case);
GPU: every golden case, a larger batch against the oracle, and the EarlySNF plugin class.
Bars: fused cross block within 1e-10 relative (float64 exp and summation orders differ from numpy / scipy in the last
bits); CRP bits and Qmax / Dmax identical on these cases (their neighbour gaps are far above that noise)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "emu"))

from acoss_b200 import _native  # noqa: E402
from oracle import snf_oracle  # noqa: E402
from oracle.make_golden_snf import synth_pair_songs  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden", "snf_golden.npz")


def _check_golden(ctx, cases):
    G = np.load(GOLD)
    for t in cases:
        seed, La, Lb, m = [int(x) for x in G["cases"][t]]
        kappa = float(G["kappas"][t])
        songs = synth_pair_songs(seed, (La, Lb))
        ctx.upload_songs(songs)
        ctx.upload_aux(np.concatenate([s["mfcc"].T for s in songs], axis=0))
        r = ctx.snf_pairs(np.array([[0, 1]], np.int32), m=m, kappa=kappa, niters=3, want_fused=True, want_crp=True)
        q, d, oti, M, N, K = G["c%d_scores" % t]
        assert r["oti"][0] == int(oti) and tuple(r["shape"][0]) == (int(M), int(N))
        ref = -G["c%d_fused" % t]
        assert np.abs(r["fused"][0] - ref).max() <= 1e-10 * np.abs(ref).max()
        crp = np.unpackbits(G["c%d_crp" % t], axis=1)[:, :int(N)]
        assert np.array_equal(r["crp"][0], crp)
        assert (r["qmax_raw"][0], r["dmax_raw"][0]) == (q, d)


def test_emu_snf_golden_smallest_case():
    from build_emu import build
    ctx = _native.Context(0, build())
    try:
        _check_golden(ctx, [0])
        with pytest.raises(_native.AcossError):   # songs too short for the neighbour counts
            ctx.snf_pairs(np.array([[0, 1]], np.int32), m=30, kappa=0.095)
    finally:
        ctx.close()


@pytest.mark.gpu
def test_gpu_snf_golden_cases():
    ctx = _native.Context(0)
    try:
        _check_golden(ctx, [0, 1, 2])
    finally:
        ctx.close()


@pytest.mark.gpu
def test_gpu_snf_batch_vs_oracle_and_plugin(tmp_path, monkeypatch):
    """A batch of pairs of bench-like lengths against the oracle, through the EarlySNF plugin class."""
    from acoss_b200.early_snf import EarlySNF
    monkeypatch.chdir(tmp_path)
    songs = synth_pair_songs(31, (150, 260, 210, 330, 180))
    labels = ["a", "a", "b", "b", "c"]
    alg = EarlySNF(songs=[dict(s) for s in songs], labels=labels, shortname="snf", do_memmaps=True)
    idxs = np.array([(0, 1), (2, 3), (1, 4), (3, 0), (4, 2)], np.int32)
    sims = alg.similarity(idxs)
    assert sorted(sims) == sorted(alg.similarity_types)
    ctx, slot = alg._ctx["chroma"]
    raw = ctx.snf_pairs(slot[idxs].astype(np.int32), m=9, kappa=0.095, niters=3, want_fused=True, want_crp=True)
    for p, (i, j) in enumerate(idxs):
        Si, Sj = songs[i], songs[j]
        o = snf_oracle.early_snf_pair(Si["chroma"].astype(np.float64), Sj["chroma"].astype(np.float64), Si["gchroma"], Sj["gchroma"],
                                      Si["mfcc"].astype(np.float64), Sj["mfcc"].astype(np.float64), m=9, kappa=0.095, niters=3)
        ref = -o["fused"]
        assert np.abs(raw["fused"][p] - ref).max() <= 1e-10 * np.abs(ref).max(), p
        assert np.array_equal(raw["crp"][p], o["crp"]), (p, int((raw["crp"][p] != o["crp"]).sum()))
        assert sims["snf_qmax"][p] == o["qmax_raw"] / (o["M"] + o["N"]) and sims["snf_dmax"][p] == o["dmax_raw"] / (o["M"] + o["N"])
        assert alg.Ds["snf_qmax"][i, j] == np.float32(sims["snf_qmax"][p])
    alg.close()
