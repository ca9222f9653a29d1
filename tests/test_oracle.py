"""CPU tests: the oracle restatement against the golden vectors produced by the
reference itself (oracle/make_golden.py), and against the compiled reference C
file when it is available (oracle/_ref)."""
import numpy as np
import pytest

from conftest import golden_case_songs, golden_cases
from oracle import crp_oracle, dp


def test_golden_pairs_match_oracle(golden):
    G = golden
    n_checked = 0
    for case in golden_cases(G):
        tag, m, kappa = case["tag"], case["m"], case["kappa"]
        songs = golden_case_songs(G, tag)
        for p in range(case["npairs"]):
            key = "%s_p%d" % (tag, p)
            i, j = G[key + "_ij"]
            Si, Sj = songs[i], songs[j]
            oti, B = crp_oracle.crp_pair(Si["chroma"].astype(np.float64), Sj["chroma"].astype(np.float64),
                                         Si["gchroma"], Sj["gchroma"], m, kappa)
            assert oti == int(G[key + "_oti"])
            M, N = G[key + "_shape"]
            assert B.shape == (M, N)
            assert np.array_equal(crp_oracle.pack_bits(B), G[key + "_crp"])
            q, d = dp.qmax_dmax_serra09(B)
            dcl = dp.dmax_clean(B)
            sw = dp.sw_score(B)
            ref = G[key + "_scores"]
            assert q == ref[0] and d == ref[1] and dcl == ref[2]
            assert sw == pytest.approx(ref[3], rel=1e-6, abs=1e-6)
            n_checked += 1
    assert n_checked >= 30


def test_golden_bare_dp_match_oracle(golden):
    G = golden
    for t, (M, N) in enumerate(G["dp_shapes"]):
        B = crp_oracle.unpack_bits(G["dp_%d_crp" % t], N)
        assert B.shape == (M, N)
        ref = G["dp_%d_scores" % t]
        q, d = dp.qmax_dmax_serra09(B)
        assert (q, d) == (ref[0], ref[1])
        assert dp.dmax_clean(B) == ref[2]
        assert dp.sw_score(B) == pytest.approx(ref[3], rel=1e-6, abs=1e-6)


def test_oracle_c_matches_compiled_reference():
    ref = dp.ref_lib()
    if ref is None:
        pytest.skip("reference C file neither buildable nor prebuilt here")
    rng = np.random.default_rng(7)
    for trial in range(40):
        M, N = rng.integers(1, 90, size=2)
        B = (rng.uniform(size=(M, N)) < rng.uniform(0.03, 0.5)).astype(np.uint8)
        assert dp.qmax_dmax_serra09(B) == dp.qmax_dmax_serra09(B, ref)
        assert dp.dmax_clean(B) == dp.dmax_clean(B, ref)
        assert dp.sw_score(B) == pytest.approx(dp.sw_score(B, ref), rel=1e-6, abs=1e-6)


def test_round_half_even_neighbour_count():
    # SURVEY.md section 7 item 3 (CRPUtils.py:191)
    exp = {100: 10, 142: 13, 300: 28, 500: 48, 5: 0, 6: 1}
    for n, k in exp.items():
        assert crp_oracle.n_neighbours(0.095, n) == k


def test_pack_roundtrip():
    rng = np.random.default_rng(0)
    B = (rng.uniform(size=(13, 77)) < 0.3).astype(np.uint8)
    assert np.array_equal(crp_oracle.unpack_bits(crp_oracle.pack_bits(B), 77), B)


def test_explicit_embedding_equals_diag_window():
    # CRPUtils.py:8-22 vs :24-45 (SURVEY.md 8a row A5'): same distances
    rng = np.random.default_rng(3)
    X = rng.uniform(size=(40, 12))
    Y = rng.uniform(size=(55, 12))
    m = 9
    def embed(Z):
        Mz = Z.shape[0] - m + 1
        return np.concatenate([Z[k:k + Mz] for k in range(m)], axis=1)
    a = crp_oracle.cross_dist(embed(X), embed(Y))
    b = crp_oracle.diag_window(crp_oracle.cross_dist(X, Y), m)
    assert np.allclose(a, b, rtol=1e-10, atol=1e-10)
