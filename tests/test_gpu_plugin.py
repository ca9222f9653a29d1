"""GPU tests of the drop-in plugin surface: Serra09.all_pairwise / do_batch / load_batches /
getEvalStatistics produce the reference's artefacts with the reference's numbers."""
import os

import numpy as np
import pytest

from acoss_b200 import hdf5io, synth
from acoss_b200.serra09 import Serra09
from oracle import crp_oracle

pytestmark = pytest.mark.gpu


def test_all_pairwise_matches_oracle_and_writes_artefacts(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    songs, labels = synth.make_song_set("covers80", seed=80, n_songs=24)
    alg = Serra09(songs=songs, labels=labels, shortname="t24", do_memmaps=True, do_mfcc=False, cachedir="cache")
    alg.all_pairwise(parallel=1, n_cores=8, symmetric=True)
    ref = crp_oracle.all_pairs(songs, m=9, kappa=0.095)   # CoverAlgorithm.py:166-183 restated
    for key in ("chroma_qmax", "chroma_dmax"):
        assert np.array_equal(np.asarray(alg.Ds[key]), ref[key]), key
        assert os.path.getsize("cache/Serra09_t24_%s_dmat" % key) == 24 * 24 * 4
        raw = np.fromfile("cache/Serra09_t24_%s_dmat" % key, dtype=np.float32).reshape(24, 24)
        assert np.array_equal(raw, ref[key])
    saved = hdf5io.load("cache/Serra09_t24_Ds.h5")
    assert np.array_equal(saved["chroma_qmax"], ref["chroma_qmax"])
    assert os.path.exists("cache/Serra09_t24_clique_info.txt")
    from scipy.io import loadmat
    mat = loadmat("cache/Serra09_t24_Ds.mat")
    assert mat["chroma_qmax"].dtype == np.float32 and np.array_equal(mat["chroma_dmax"], ref["chroma_dmax"])
    # the score matrices all_pairwise left on the device are the host ones, bit for bit, and getEvalStatistics ranks
    # them there (no N x N upload) with the same result as through the host matrix
    assert alg._Ds_dev_valid
    for key in ("chroma_qmax", "chroma_dmax"):
        assert np.array_equal(alg._Ds_dev[key].cpu().numpy(), ref[key]), key
    MR, MRR, MDR, MAP, tops = alg.getEvalStatistics("chroma_qmax")
    alg._Ds_dev_valid = False
    host = alg.getEvalStatistics("chroma_qmax")
    alg._Ds_dev_valid = True
    assert (MR, MRR, MDR, MAP) == host[:4] and np.array_equal(tops, host[4])
    assert 0 < MAP <= 1 and MR >= 1
    # covers are recognisable in the synthetic set: the mate of most songs ranks first
    assert tops[0] >= 12
    alg.cleanup_memmap()
    assert not os.path.exists("cache/Serra09_t24_chroma_qmax_dmat")
    alg.close()


def test_batch_mode_equals_all_pairwise(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    songs, labels = synth.make_song_set("covers80", seed=81, n_songs=16, length_range=(100, 220))
    full = Serra09(songs=songs, labels=labels, shortname="full16", do_memmaps=True, do_mfcc=False, cachedir="cache")
    full.all_pairwise(symmetric=True)
    Dfull = np.array(full.Ds["chroma_qmax"])
    full.close()
    batch = Serra09(songs=songs, labels=labels, shortname="b16", do_memmaps=False, do_mfcc=False, cachedir="cache")
    w = 8
    for idx in range(3):               # (16/8) * (16/8 + 1) / 2 blocks, like runcovers80.sh
        batch.do_batch(w, idx, wsub=4)
    # resume: a second call finds every sub-block completed and recomputes nothing
    n_before = batch.kernel_launches()
    batch.do_batch(w, 0, wsub=4)
    assert batch.kernel_launches() == n_before
    agg = Serra09(songs=songs, labels=labels, shortname="agg16", do_memmaps=True, do_mfcc=False, cachedir="cache")
    agg.load_batches("cache/Serra09_b16_")
    D = np.array(agg.Ds["chroma_qmax"])
    off = ~np.eye(16, dtype=bool)
    # batch mode scores (row >= col) pairs, i.e. (j, i) where all_pairwise scored (i, j): equal by symmetry
    assert np.allclose(D[off], Dfull[off], rtol=0, atol=0)
    assert np.all(np.diag(D) > 0)      # self pairs, counted twice (CoverAlgorithm.py:314-315)
    batch.close()
    agg.close()


def test_mfcc_similarity_types(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    rng = np.random.default_rng(2)
    songs, labels = synth.make_song_set("covers80", seed=82, n_songs=6, length_range=(80, 150))
    for s in songs:
        L = s["chroma"].shape[1]
        s["mfcc"] = (rng.normal(size=(13, L)).cumsum(axis=1) * 0.2 + rng.normal(size=(13, L))).astype(np.float32)
    alg = Serra09(songs=songs, labels=labels, shortname="m6", do_memmaps=False, do_mfcc=True)
    idxs = np.array([(0, 1), (2, 5), (4, 3)], np.int32)
    sims = alg.similarity(idxs)
    assert sorted(sims) == ["chroma_dmax", "chroma_qmax", "mfcc_dmax", "mfcc_qmax"]
    for p, (i, j) in enumerate(idxs):
        r = crp_oracle.serra09_pair(songs[i]["mfcc"].astype(np.float64), songs[j]["mfcc"].astype(np.float64),
                                    None, None, 9, 0.095, do_oti=False)
        assert sims["mfcc_qmax"][p] == r["qmax"] and sims["mfcc_dmax"][p] == r["dmax"]
    alg.close()


def test_raw_features_are_aggregated_on_the_device(tmp_path, monkeypatch):
    """Songs given as full-rate features (crema (n, 12) float64 + mfcc_htk (13, n) float32, the
    fields of the reference's per-song files, extractors.py:43-54): load_features_batch (GPU) must
    give exactly what load_features (numpy, Serra09.py:99-114 restated) gives, and the scores
    computed from them must match."""
    monkeypatch.chdir(tmp_path)
    rng = np.random.default_rng(5)
    raw = []
    for s in range(6):
        n = int(rng.integers(4000, 9000))
        base = np.abs(rng.standard_normal((n // 50 + 2, 12))).repeat(50, axis=0)[:n]
        crema = (base + 0.05 * rng.random((n, 12))).astype(np.float64)
        mfcc = (rng.standard_normal((13, n)) * 20).astype(np.float32)
        mfcc[3, 7] = np.nan
        raw.append({"crema": crema, "mfcc_htk": mfcc, "label": "c%d" % (s // 2)})
    a = Serra09(songs=raw, shortname="raw", do_memmaps=False, do_mfcc=True, cachedir="cache")
    b = Serra09(songs=raw, shortname="raw2", do_memmaps=False, do_mfcc=True, cachedir="cache")
    b.load_features_batch(np.arange(6))
    for i in range(6):
        fa, fb = a.load_features(i), b.all_feats[i]
        assert np.array_equal(fa["gchroma"], fb["gchroma"])
        assert fa["chroma"].dtype == fb["chroma"].dtype and np.array_equal(fa["chroma"], fb["chroma"])
        assert fa["mfcc"].dtype == fb["mfcc"].dtype and np.array_equal(fa["mfcc"], fb["mfcc"])
    idx = np.array([(0, 1), (2, 3), (0, 5)])
    sa, sb = a.similarity(idx), b.similarity(idx)
    for k in sa:
        assert np.array_equal(sa[k], sb[k]), k
    a.close()
    b.close()


def test_chen_fusion_plugin_vs_oracle(tmp_path, monkeypatch):
    """ChenFusion.similarity (ChenFusion.py:57-67): blocked OTI + Euclidean CSM of the one-step "stack" + row-only kNN +
    qmax / dmax on zeroed tables, un-normalised; then normalize_by_length (ChenFusion.py:69-79)."""
    from acoss_b200.chen_fusion import ChenFusion
    from oracle import crp_oracle, dp
    songs, labels = synth.make_song_set("covers80", seed=71, n_songs=10, length_range=(100, 420))
    monkeypatch.chdir(tmp_path)
    alg = ChenFusion(songs=songs, labels=labels, shortname="chen", do_memmaps=True)
    alg.all_pairwise(symmetric=True)
    n = len(songs)
    for i in range(n):
        for j in range(i + 1, n):
            _, _, S = crp_oracle.crp_pair(songs[i]["chroma"].astype(np.float64), songs[j]["chroma"].astype(np.float64),
                                          songs[i]["gchroma"], songs[j]["gchroma"], 1, 0.095, want_S=True)
            B = crp_oracle.knn_mask(S, 0.095)
            assert alg.Ds["qmax"][i, j] == np.float32(dp.qmax_only(B)) and alg.Ds["dmax"][i, j] == np.float32(dp.dmax_clean(B))
            assert alg.Ds["qmax"][j, i] == alg.Ds["qmax"][i, j]
    q = np.array(alg.Ds["qmax"])
    alg.normalize_by_length()
    L = np.array([s["chroma"].shape[1] for s in songs], np.float64)
    with np.errstate(divide="ignore"):
        assert np.array_equal(np.asarray(alg.Ds["qmax"]), (np.sqrt(L)[None, :] / q.astype(np.float64)).astype(np.float32))
    with pytest.raises(NotImplementedError):
        alg.do_late_fusion()
    with pytest.raises(NotImplementedError):
        ChenFusion(songs=songs, labels=labels, tau=2, do_memmaps=False)
    alg.close()
