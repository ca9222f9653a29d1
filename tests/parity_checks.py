"""Parity checks shared by the GPU tests (tests/test_gpu_parity.py, through the real C ABI on a
B200) and the CPU emulator tests (tests/test_emu_kernels.py, the same kernel sources stepped by
tests/emu/cuda_emu.h).  Every function takes a live `_native.Context` and compares it with the
oracle (oracle/) or the golden vectors generated from the reference (tests/golden/).

Bars (BASELINE.json north_star): OTI and CRP bits exact; Qmax / Dmax maxima exact (their
penalties are dyadic); Smith-Waterman within 1e-5 relative (its -0.7 is not dyadic).
"""
from __future__ import annotations

import numpy as np

from acoss_b200 import _native, synth
from oracle import crp_oracle, dp

SW_RTOL = 1e-5
TIE_MARGIN = 1e-9  # relative gap between k-th and (k+1)-th neighbour below which a pair is "tied"


def sw_close(a, b):
    return abs(float(a) - float(b)) <= SW_RTOL * max(1.0, abs(float(b)))


def golden_cases(G):
    return [dict(zip(G["cases"].dtype.names, row)) for row in G["cases"].tolist()]


def golden_songs(G, tag):
    n = len(G["%s_lengths" % tag])
    return [{"chroma": G["%s_chroma_%d" % (tag, s)], "gchroma": G["%s_gchroma_%d" % (tag, s)]} for s in range(n)]


def check_golden_pairs(ctx, G, tags=None, flag_sets=(_native.FLAG_NO_SWAP, 0)):
    """Golden pairs produced by the reference itself: OTI, CRP bits, qmax, dmax (stale table), sw."""
    n = 0
    for case in golden_cases(G):
        tag, m, kappa = case["tag"], case["m"], case["kappa"]
        if tags is not None and tag not in tags:
            continue
        ctx.upload_songs(golden_songs(G, tag))
        pairs = np.array([G["%s_p%d_ij" % (tag, p)] for p in range(case["npairs"])])
        for flags in flag_sets:
            r = ctx.serra09_pairs(pairs, m=m, kappa=kappa, flags=flags, want_crp=True, want_sw=True)
            for p in range(case["npairs"]):
                key = "%s_p%d" % (tag, p)
                M, N = G[key + "_shape"]
                ref = crp_oracle.unpack_bits(G[key + "_crp"], N)
                sc = G[key + "_scores"]
                where = "case %s pair %d flags %d" % (tag, p, flags)
                assert tuple(r["shape"][p]) == (M, N), where
                assert r["oti"][p] == int(G[key + "_oti"]), where
                assert np.array_equal(r["crp"][p], ref), where + ": %d CRP bits differ" % int(
                    (r["crp"][p] != ref).sum())
                assert r["qmax_raw"][p] == sc[0], where
                assert r["dmax_raw"][p] == sc[1], where
                assert sw_close(r["sw_raw"][p], sc[3]), where
                n += 1
            # clean-table dmax (ChenFusion.py:66)
            rc = ctx.serra09_pairs(pairs, m=m, kappa=kappa, flags=flags | _native.FLAG_DMAX_CLEAN)
            for p in range(case["npairs"]):
                assert rc["dmax_raw"][p] == G["%s_p%d_scores" % (tag, p)][2], "clean dmax case %s pair %d" % (tag, p)
    return n


def check_golden_dp(ctx, G):
    """Bare DP known answers (degenerate shapes included) through acoss_dp_packed and the
    pySeqAlign-style scalar calls."""
    plots, refs = [], []
    for t, (M, N) in enumerate(G["dp_shapes"]):
        plots.append(crp_oracle.unpack_bits(G["dp_%d_crp" % t], N))
        refs.append(G["dp_%d_scores" % t])
    stale = ctx.dp_packed(plots, flags=0, want_sw=True)
    clean = ctx.dp_packed(plots, flags=_native.FLAG_DMAX_CLEAN, want_sw=True)
    for t, ref in enumerate(refs):
        where = "dp plot %d shape %s" % (t, plots[t].shape)
        assert stale[t, 0] == ref[0] and stale[t, 1] == ref[1], where
        assert clean[t, 0] == ref[0] and clean[t, 1] == ref[2], where
        assert sw_close(stale[t, 2], ref[3]), where
    # scalar drop-ins (argument order of pySeqAlign.pyx:7-23)
    for t in (4, 7, 9):
        B = plots[t]
        M, N = B.shape
        assert ctx.qmax(B.flatten(), M, N) == refs[t][0]
        assert ctx.dmax(B.flatten(), M, N) == refs[t][2]
        assert ctx.qmax_dmax(B.flatten(), M, N) == (refs[t][0], refs[t][1])
        assert sw_close(ctx.swconstrained(B.flatten(), M, N), refs[t][3])
    return len(plots)


def oracle_pair(songs, i, j, m, kappa, want_sw=False):
    """The oracle's answer for one pair + its tie margins."""
    Si, Sj = songs[i], songs[j]
    oti, B, S = crp_oracle.crp_pair(Si["chroma"].astype(np.float64), Sj["chroma"].astype(np.float64),
                                    Si["gchroma"], Sj["gchroma"], m, kappa, want_S=True)
    q, d = dp.qmax_dmax_serra09(B)
    out = {"oti": oti, "crp": B, "qmax_raw": q, "dmax_raw": d, "margin": min(crp_oracle.tie_margins(S, kappa))}
    if want_sw:
        out["sw_raw"] = dp.sw_score(B)
    return out


def check_against_oracle(ctx, songs, pairs, m=9, kappa=0.095, flags=0, want_sw=False, want_crp=True):
    """Seeded synthetic songs: compare every pair with the oracle.  Returns (#pairs, #tied pairs
    excluded from the bit-exact CRP set)."""
    pairs = np.asarray(pairs, dtype=np.int32).reshape(-1, 2)
    ctx.upload_songs(songs)
    r = ctx.serra09_pairs(pairs, m=m, kappa=kappa, flags=flags, want_crp=want_crp, want_sw=want_sw)
    tied = 0
    for p, (i, j) in enumerate(pairs):
        o = oracle_pair(songs, int(i), int(j), m, kappa, want_sw)
        where = "pair %d = (%d, %d) shape %s flags %d" % (p, i, j, o["crp"].shape, flags)
        assert r["oti"][p] == o["oti"], where
        if not (o["margin"] > TIE_MARGIN):
            tied += 1
            continue
        if want_crp:
            assert np.array_equal(r["crp"][p], o["crp"]), where + ": %d CRP bits differ" % int(
                (r["crp"][p] != o["crp"]).sum())
        assert r["qmax_raw"][p] == o["qmax_raw"], where + " qmax %r vs %r" % (r["qmax_raw"][p], o["qmax_raw"])
        assert r["dmax_raw"][p] == o["dmax_raw"], where + " dmax %r vs %r" % (r["dmax_raw"][p], o["dmax_raw"])
        if want_sw:
            assert sw_close(r["sw_raw"][p], o["sw_raw"]), where
    return len(pairs), tied


def check_options_invariant(ctx, songs, pairs, m=9, kappa=0.095):
    """The tuning knobs select different kernels / scratch placements; results must not move."""
    ctx.upload_songs(songs)
    base = ctx.serra09_pairs(pairs, m=m, kappa=kappa, want_crp=True, want_sw=True, flags=_native.FLAG_NO_SWAP)
    for opts in ({"force_wide_sweep": 1}, {"force_wide_sweep": 1, "tiled_wide": 0}, {"force_gscratch": 1}, {"dp_max_cpl": 16},
                 {"chunk_pairs": 3}, {"force_wide_sweep": 1, "force_gscratch": 1, "dp_max_cpl": 8, "chunk_pairs": 2}):
        for k, v in opts.items():
            ctx.set_option(k, v)
        try:
            r = ctx.serra09_pairs(pairs, m=m, kappa=kappa, want_crp=True, want_sw=True, flags=_native.FLAG_NO_SWAP)
        finally:
            for k, v in {"force_wide_sweep": 0, "force_gscratch": 0, "dp_max_cpl": 4, "chunk_pairs": 1 << 18,
                         "tiled_wide": 1}.items():
                ctx.set_option(k, v)
        for key in ("qmax_raw", "dmax_raw", "sw_raw", "oti"):
            assert np.array_equal(r[key], base[key]), "%s changed under %s" % (key, opts)
        for a, b in zip(r["crp"], base["crp"]):
            assert np.array_equal(a, b), "CRP changed under %s" % (opts,)
    return base


def check_dense(ctx, shapes, seed=5, kappa=0.095):
    """EarlySNF tail (EarlySNF.py:86-90): dense matrix -> mutual kNN mask -> qmax/dmax (+ sw on the
    non-mutual mask as EarlySNF_Old.py:200-201 does)."""
    rng = np.random.default_rng(seed)
    mats = [synth.dense_affinity_matrix(rng, M, N) for (M, N) in shapes]
    for dt in (np.float32, np.float64):
        ms = [a.astype(dt) for a in mats]
        out, crps = ctx.align_dense(ms, kappa=kappa, want_sw=True, want_crp=True)
        out_nm, crps_nm = ctx.align_dense(ms, kappa=kappa, flags=_native.FLAG_NON_MUTUAL | _native.FLAG_DMAX_CLEAN,
                                          want_sw=True, want_crp=True)
        for b, a in enumerate(ms):
            a64 = a.astype(np.float64)
            B = crp_oracle.mutual_mask(a64, kappa)
            assert np.array_equal(crps[b], B), "dense %d mutual mask" % b
            q, d = dp.qmax_dmax_serra09(B)
            assert out[b, 0] == q and out[b, 1] == d, "dense %d scores" % b
            assert sw_close(out[b, 2], dp.sw_score(B))
            Bn = crp_oracle.knn_mask(a64, kappa)
            assert np.array_equal(crps_nm[b], Bn), "dense %d row mask" % b
            assert out_nm[b, 0] == dp.qmax_only(Bn) and out_nm[b, 1] == dp.dmax_clean(Bn)
            assert sw_close(out_nm[b, 2], dp.sw_score(Bn))
    return len(mats)


def check_degenerate(ctx):
    """Songs shorter than the window, plots too small for the recurrences: zeros, never an error
    (SequenceAlignment.c:78-80,117-119,151-153)."""
    rng = np.random.default_rng(9)
    lens = [3, 9, 10, 11, 12, 40]
    songs = []
    for L in lens:
        fr = rng.uniform(0.01, 1.0, size=(L, 12)).astype(np.float32)
        songs.append({"chroma": np.ascontiguousarray(fr.T), "gchroma": synth.global_chroma_f64(fr)})
    ctx.upload_songs(songs)
    pairs = np.array([(i, j) for i in range(len(lens)) for j in range(len(lens)) if i != j], np.int32)
    r = ctx.serra09_pairs(pairs, m=9, kappa=0.095, want_crp=True, want_sw=True)
    for p, (i, j) in enumerate(pairs):
        M, N = lens[i] - 8, lens[j] - 8
        assert tuple(r["shape"][p]) == (M, N)
        if M <= 0 or N <= 0:
            assert r["qmax_raw"][p] == 0 and r["dmax_raw"][p] == 0 and r["sw_raw"][p] == 0
            continue
        o = oracle_pair(songs, int(i), int(j), 9, 0.095, True)
        assert np.array_equal(r["crp"][p], o["crp"])
        assert r["qmax_raw"][p] == o["qmax_raw"] and r["dmax_raw"][p] == o["dmax_raw"]
        assert sw_close(r["sw_raw"][p], o["sw_raw"])
    return len(pairs)


def check_symmetry(ctx, songs, n_pairs=200, seed=3, m=9, kappa=0.095):
    """Size-independent property used at full song lengths: scoring (j, i) rolls the other song
    by the complementary shift, so the CRP is the transpose and Qmax/Dmax are unchanged
    (both recurrences are transpose-invariant)."""
    rng = np.random.default_rng(seed)
    n = len(songs)
    ij = rng.integers(0, n, size=(n_pairs, 2)).astype(np.int32)
    ij = ij[ij[:, 0] != ij[:, 1]]
    ctx.upload_songs(songs)
    a = ctx.serra09_pairs(ij, m=m, kappa=kappa)
    b = ctx.serra09_pairs(ij[:, ::-1].copy(), m=m, kappa=kappa)
    comp = (12 - a["oti"]) % 12
    same = comp == b["oti"]  # a tie in the OTI agreement can break this; skip those pairs
    assert same.mean() > 0.95
    assert np.array_equal(a["qmax_raw"][same], b["qmax_raw"][same])
    assert np.array_equal(a["dmax_raw"][same], b["dmax_raw"][same])
    return int(same.sum())


def check_ties_golden(ctx, G, pairs_per_case=None):
    """Songs with exactly repeated frames (tests/golden/ties_golden.npz, reference-produced by
    oracle/make_golden_ties.py).  Where equal distances straddle the k-th neighbour the reference keeps exactly k
    (argpartition's pick among the tied cells follows the rounding noise of its cumsum differences), the product keeps
    the cells strictly below the separator: DESIGN.md section 2, deviation 1.  Asserted: OTI; every cell whose row and
    column are both untied; whole plots and scores of the pairs without a tied line.  Counted and returned: the cells
    and scores that differ on tied lines."""
    m, kappa = int(G["m"]), float(G["kappa"])
    st = {"pairs": 0, "pairs_tied": 0, "cells": 0, "cells_on_tied_lines": 0, "bits_missing": 0, "bits_extra": 0,
          "ones_ref": 0, "qmax_rel": [], "dmax_rel": [], "pairs_score_equal": 0}
    for tag, n, npairs in G["cases"].tolist():
        if pairs_per_case is not None:
            npairs = min(npairs, pairs_per_case)
        songs = [{"chroma": G["%s_chroma_%d" % (tag, s)], "gchroma": G["%s_gchroma_%d" % (tag, s)]} for s in range(n)]
        ctx.upload_songs(songs)
        pairs = np.array([G["%s_p%d_ij" % (tag, p)] for p in range(npairs)])
        r = ctx.serra09_pairs(pairs, m=m, kappa=kappa, flags=_native.FLAG_NO_SWAP, want_crp=True)
        for p in range(npairs):
            key = "%s_p%d" % (tag, p)
            M, N = G[key + "_shape"]
            ref = crp_oracle.unpack_bits(G[key + "_crp"], N)
            rows = np.unpackbits(G[key + "_tied_rows"])[:M].astype(bool)
            cols = np.unpackbits(G[key + "_tied_cols"])[:N].astype(bool)
            sc = G[key + "_scores"]
            got = r["crp"][p]
            where = "ties case %s pair %d" % (tag, p)
            assert tuple(r["shape"][p]) == (M, N), where
            assert r["oti"][p] == int(G[key + "_oti"]), where
            clean = ~rows[:, None] & ~cols[None, :]
            assert np.array_equal(got[clean], ref[clean]), where + ": %d bits differ off the tied lines" % int(
                (got[clean] != ref[clean]).sum())
            st["pairs"] += 1
            st["cells"] += int(M) * int(N)
            st["ones_ref"] += int(ref.sum())
            if rows.any() or cols.any():
                st["pairs_tied"] += 1
                st["cells_on_tied_lines"] += int((~clean).sum())
                st["bits_missing"] += int((ref & ~got & 1).sum())
                st["bits_extra"] += int((got & ~ref & 1).sum())
                st["qmax_rel"].append(abs(float(r["qmax_raw"][p]) - sc[0]) / max(sc[0], 1.0))
                st["dmax_rel"].append(abs(float(r["dmax_raw"][p]) - sc[1]) / max(sc[1], 1.0))
                st["pairs_score_equal"] += int(r["qmax_raw"][p] == sc[0] and r["dmax_raw"][p] == sc[1])
            else:
                assert np.array_equal(got, ref), where
                assert r["qmax_raw"][p] == sc[0] and r["dmax_raw"][p] == sc[1], where
    return st
