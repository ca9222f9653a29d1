"""CPU-only: the product's kernel SOURCES compiled with g++ against the SIMT emulator
(tests/emu/cuda_emu.h) and checked against the golden vectors / the oracle on small cases.
This validates indexing and synchronisation logic where no GPU exists; the parity tests
proper are tests/test_gpu_parity.py (real library, real B200)."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "emu"))

import parity_checks as pc  # noqa: E402
from acoss_b200 import _native, synth  # noqa: E402


@pytest.fixture(scope="module")
def ctx():
    from build_emu import build
    c = _native.Context(0, build())
    yield c
    c.close()


def test_emu_golden_pairs_small(ctx, golden):
    assert pc.check_golden_pairs(ctx, golden, tags=("c", "e"), flag_sets=(_native.FLAG_NO_SWAP, 0)) >= 12


def test_emu_golden_dp(ctx, golden):
    assert pc.check_golden_dp(ctx, golden) == 12


def test_emu_options_invariant(ctx):
    songs, _ = synth.make_song_set("covers80", seed=21, n_songs=4, length_range=(30, 60))
    pairs = np.array([(0, 1), (1, 2), (2, 3), (3, 0), (0, 2)], np.int32)
    base = pc.check_options_invariant(ctx, songs, pairs, m=5, kappa=0.1)
    # and the base itself is right
    for p, (i, j) in enumerate(pairs):
        o = pc.oracle_pair(songs, int(i), int(j), 5, 0.1, True)
        assert np.array_equal(base["crp"][p], o["crp"])
        assert base["qmax_raw"][p] == o["qmax_raw"] and base["dmax_raw"][p] == o["dmax_raw"]


def test_emu_tiled_long_song_sweep(ctx):
    """kernels_sweep1t.cuh under emulation: a column song that needs two tiles (1024 frame columns each, m - 1
    overlap), T matrix + row selection + tiled CRP emission, against the oracle; then the one-CTA general sweep
    (tiled_wide = 0) must give the same bits."""
    songs, _ = synth.make_song_set("covers80", seed=51, n_songs=2, length_range=(40, 70))
    songs += synth.make_song_set("covers80", seed=52, n_songs=1, fixed_length=1450)[0]
    pairs = np.array([(0, 2), (2, 1)], np.int32)
    n, tied = pc.check_against_oracle(ctx, songs, pairs, m=9, kappa=0.095, flags=_native.FLAG_NO_SWAP)
    assert n == 2 and tied == 0
    pc.check_against_oracle(ctx, songs, pairs[:1], m=20, kappa=0.095, flags=0)
    ctx.set_option("force_wide_sweep", 1)
    try:
        a = ctx.serra09_pairs(pairs, m=9, kappa=0.095, want_crp=True)
        ctx.set_option("tiled_wide", 0)
        b = ctx.serra09_pairs(pairs, m=9, kappa=0.095, want_crp=True)
    finally:
        ctx.set_option("tiled_wide", 1)
        ctx.set_option("force_wide_sweep", 0)
    for x, y in zip(a["crp"], b["crp"]):
        assert np.array_equal(x, y)
    assert np.array_equal(a["qmax_raw"], b["qmax_raw"]) and np.array_equal(a["dmax_raw"], b["dmax_raw"])


def test_emu_dense(ctx):
    assert pc.check_dense(ctx, [(20, 37), (45, 33), (5, 4)]) == 3


def test_emu_degenerate(ctx):
    assert pc.check_degenerate(ctx) == 30


def test_emu_dp_multiblock(ctx):
    # plots wider than one column block: hand-over through the boundary scratch
    rng = np.random.default_rng(4)
    plots = [(rng.uniform(size=(M, N)) < dens).astype(np.uint8)
             for (M, N, dens) in [(30, 200, 0.2), (25, 131, 0.4), (40, 129, 0.08), (12, 300, 0.3)]]
    from oracle import dp
    ctx.set_option("dp_max_cpl", 4)
    try:
        out = ctx.dp_packed(plots, want_sw=True)
    finally:
        ctx.set_option("dp_max_cpl", 4)
    for b, B in enumerate(plots):
        q, d = dp.qmax_dmax_serra09(B)
        assert (out[b, 0], out[b, 1]) == (q, d)
        assert pc.sw_close(out[b, 2], dp.sw_score(B))


def test_emu_integer_key_sweeps(ctx):
    """gen 4 / gen 7 kernel sources under emulation: several warps per row (halo columns), two diagonals per thread
    (the 512-column class), inflated safety gap (float64 fallback path), other windows."""
    songs, _ = synth.make_song_set("covers80", seed=33, n_songs=5, length_range=(60, 200))
    songs += synth.make_song_set("covers80", seed=34, n_songs=2, length_range=(500, 560))[0]
    pairs = np.array([(0, 1), (1, 2), (2, 3), (3, 4), (4, 0), (5, 6), (6, 2)], np.int32)
    try:
        for gen in (4, 7):
            ctx.set_option("sweep_gen", gen)
            # the default generation on every pair and both gaps; its predecessor (cross-checked at full size on the
            # GPU, test_sweep_generations_agree_and_exact_path) on the pairs that reach their widest CTAs
            sub = pairs if gen == 7 else pairs[[0, 3, 5, 6]]
            for gap in ((2, 4096) if gen == 7 else (4096,)):
                ctx.set_option("gap_units", gap)
                n, tied = pc.check_against_oracle(ctx, songs, sub, m=9, kappa=0.095, flags=0)
                assert n == len(sub) and tied == 0
        ctx.set_option("gap_units", 2)
        pc.check_against_oracle(ctx, songs, pairs[:4], m=5, kappa=0.1, flags=_native.FLAG_NO_SWAP)
    finally:
        ctx.set_option("sweep_gen", 7)
        ctx.set_option("gap_units", 2)


def test_emu_feature_aggregation(ctx):
    """k_feat_aggregate / k_feat_gchroma under emulation against numpy (median, mean, global chroma)."""
    from oracle import features_oracle as fo
    rng = np.random.default_rng(3)
    for dtype in (np.float32, np.float64):
        songs = []
        for n in (1, 7, 39, 40, 41, 80, 333):
            a = rng.random((n, 12)).astype(dtype) ** 2
            a[rng.random((n, 12)) < 0.05] = 0
            songs.append(a)
        agg, g = ctx.aggregate_songs(songs, fac=40, aggregate="median", want_gchroma=True)
        for s, (a, got) in enumerate(zip(songs, agg)):
            assert np.array_equal(got, fo.sync(a.T, 40, np.median)), (dtype, s)
            assert np.array_equal(g[s], fo.global_chroma(a)), (dtype, s)
    songs = []
    for n in (3, 40, 47, 300):
        a = (rng.standard_normal((13, n)) * 30).astype(np.float32)
        a[0, 0] = np.nan
        songs.append(a)
    agg = ctx.aggregate_songs(songs, fac=40, aggregate="mean", layout="bins_x_frames", zero_nonfinite=True)
    for a, got in zip(songs, agg):
        m = np.array(a)
        m[np.isnan(m)] = 0
        assert np.array_equal(got, fo.sync(m, 40, np.mean))


def test_emu_dp16_packed_lanes(ctx):
    """k_dp16 (Qmax + Dmax on packed 16-bit lanes) under emulation: ragged lanes / column blocks, dense plots where the
    Dmax bonus terms matter, all three lane widths, stale and clean Dmax tables -- against the oracle and against the
    int32 kernel (dp_packed = 0)."""
    from oracle import dp
    rng = np.random.default_rng(78)
    shapes = [(12, c) for c in (4, 5, 7, 33, 127, 129, 131, 257)] + [(r, 130) for r in (3, 4, 5, 7, 10)] + [(40, 140), (5, 515)]
    plots = [(rng.uniform(size=(M, N)) < dens).astype(np.uint8) for (M, N) in shapes for dens in (0.08, 0.5, 0.9)]
    for cpl in (4, 8, 16):
        ctx.set_option("dp_max_cpl", cpl)
        try:
            out = ctx.dp_packed(plots, want_sw=False)
            outc = ctx.dp_packed(plots, flags=_native.FLAG_DMAX_CLEAN, want_sw=False)
            ctx.set_option("dp_packed", 0)
            ref = ctx.dp_packed(plots, want_sw=False)
        finally:
            ctx.set_option("dp_max_cpl", 4)
            ctx.set_option("dp_packed", 1)
        assert np.array_equal(out[:, :2], ref[:, :2]), cpl
        for b, B in enumerate(plots):
            q, d = dp.qmax_dmax_serra09(B)
            assert (out[b, 0], out[b, 1]) == (q, d), (cpl, B.shape, b % 3)
            assert outc[b, 1] == dp.dmax_clean(B), (cpl, B.shape, b % 3)


def test_emu_bound_score_matrices(ctx):
    """acoss_bind_score_matrices under emulation ("device" memory is host memory there): the scores of a call land at
    [song of slot a][song of slot b] as raw / (M + N) rounded to float32, other cells stay untouched."""
    songs, _ = synth.make_song_set("covers80", seed=23, n_songs=5, length_range=(30, 60))
    ctx.upload_songs(songs)
    n = 9
    song_of_slot = np.array([7, 2, 5, 0, 3], np.int32)
    Dq = np.full((n, n), -1.0, np.float32)
    Dd = np.full((n, n), -1.0, np.float32)
    pairs = np.array([(0, 1), (1, 2), (3, 4), (4, 0)], np.int32)
    ctx.bind_score_matrices(Dq.ctypes.data, Dd.ctypes.data, n, song_of_slot)
    try:
        r = ctx.serra09_pairs(pairs, m=5, kappa=0.1)
    finally:
        ctx.bind_score_matrices(0, 0, 0)
    want_q = np.full((n, n), -1.0, np.float32)
    want_d = np.full((n, n), -1.0, np.float32)
    for p, (a, b) in enumerate(pairs):
        mn = float(r["shape"][p].sum())
        want_q[song_of_slot[a], song_of_slot[b]] = np.float32(np.float64(r["qmax_raw"][p]) / mn)
        want_d[song_of_slot[a], song_of_slot[b]] = np.float32(np.float64(r["dmax_raw"][p]) / mn)
    assert np.array_equal(Dq, want_q) and np.array_equal(Dd, want_d)
    again = ctx.serra09_pairs(pairs, m=5, kappa=0.1)   # unbound: nothing is written any more
    assert np.array_equal(again["qmax_raw"], r["qmax_raw"]) and np.array_equal(Dq, want_q)
