"""GPU parity tests: the real libacoss_b200.so through its C ABI on a B200, against the golden
vectors generated from the reference and against the oracle on seeded synthetic inputs."""
import numpy as np
import pytest

import parity_checks as pc
from acoss_b200 import _native, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = _native.Context(0)  # raises if the CUDA library or the GPU is missing: no fallback
    yield c
    c.close()


def test_golden_pairs(ctx, golden):
    assert pc.check_golden_pairs(ctx, golden) >= 60


def test_golden_dp(ctx, golden):
    assert pc.check_golden_dp(ctx, golden) == 12


def test_degenerate(ctx):
    assert pc.check_degenerate(ctx) == 30


def test_dense_tail(ctx):
    assert pc.check_dense(ctx, [(20, 37), (45, 33), (5, 4), (300, 412), (700, 150), (1300, 90)]) == 6


def test_covers80_shaped_sample_vs_oracle(ctx):
    # covers80-shaped set (BASELINE configs[0]); a 60-pair sample keeps the oracle to ~20 s
    songs, _ = synth.make_song_set("covers80", seed=80)
    rng = np.random.default_rng(1)
    pairs = rng.integers(0, len(songs), size=(60, 2)).astype(np.int32)
    pairs = pairs[pairs[:, 0] != pairs[:, 1]]
    n, tied = pc.check_against_oracle(ctx, songs, pairs, flags=_native.FLAG_NO_SWAP, want_sw=True)
    assert tied <= 2
    n2, _ = pc.check_against_oracle(ctx, songs, pairs[:20], flags=0, want_sw=False)


def test_options_invariant(ctx):
    songs, _ = synth.make_song_set("covers80", seed=22, n_songs=8, length_range=(120, 700))
    pairs = np.array([(i, j) for i in range(8) for j in range(8) if i != j], np.int32)[:24]
    pc.check_options_invariant(ctx, songs, pairs)


def test_long_songs_vs_oracle(ctx):
    # wide sweeps (several column groups), global-scratch ring, multi-block DP, m = 20
    songs, _ = synth.make_song_set("covers80", seed=33, n_songs=4, length_range=(900, 2300))
    pairs = np.array([(0, 1), (1, 0), (2, 3), (0, 3)], np.int32)
    pc.check_against_oracle(ctx, songs, pairs, m=20, kappa=0.095, flags=_native.FLAG_NO_SWAP, want_sw=True)
    pc.check_against_oracle(ctx, songs, pairs[:2], m=20, kappa=0.095, flags=0)


def test_stress_config_4096_frames(ctx):
    """BASELINE configs[4]: 4096-frame songs, m = 20, kappa = 0.095 (M = N = 4077, k = 387).  One pair against the
    oracle bit for bit (the oracle's Python diagonal loop takes seconds at this size); the rest of a mixed
    1024..4096 batch against size-independent properties: (i, j) and (j, i) give transposed CRPs and equal
    scores, and the batched call equals the one-pair calls (variable-length batching)."""
    songs, _ = synth.make_song_set("covers80", seed=4096, n_songs=2, fixed_length=4096)
    songs += synth.make_song_set("covers80", seed=4097, n_songs=4, length_range=(1024, 4096))[0]
    n, tied = pc.check_against_oracle(ctx, songs, np.array([(0, 1)], np.int32), m=20, kappa=0.095,
                                      flags=_native.FLAG_NO_SWAP)
    assert n == 1 and tied == 0
    pairs = np.array([(0, 1), (1, 0), (2, 3), (3, 2), (4, 5), (5, 4), (0, 4), (2, 1)], np.int32)
    r = ctx.serra09_pairs(pairs, m=20, kappa=0.095, flags=_native.FLAG_NO_SWAP, want_crp=True)
    assert tuple(r["shape"][0]) == (4077, 4077)
    for a, b in ((0, 1), (2, 3), (4, 5)):
        # OTI(i, j) and OTI(j, i) are complementary rotations, so the two plots are transposes of each other
        assert (12 - r["oti"][a]) % 12 == r["oti"][b]
        assert np.array_equal(r["crp"][a], r["crp"][b].T)
        assert r["qmax_raw"][a] == r["qmax_raw"][b]
    for p in (2, 6, 7):
        one = ctx.serra09_pairs(pairs[p:p + 1], m=20, kappa=0.095, flags=_native.FLAG_NO_SWAP, want_crp=True)
        assert np.array_equal(one["crp"][0], r["crp"][p])
        assert one["qmax_raw"][0] == r["qmax_raw"][p] and one["dmax_raw"][0] == r["dmax_raw"][p]
    dens = r["crp"][0].mean()
    assert 0.02 < dens < 0.095


def test_dp_ragged_shapes_vs_oracle(ctx):
    """k_dp on plots whose last lane / last column block is ragged, dense enough that the Dmax bonus terms next
    to the plot's right edge matter (the steady rows take their maxima over the padded columns too)."""
    from oracle import dp
    rng = np.random.default_rng(77)
    shapes = [(40, c) for c in (4, 5, 6, 7, 9, 31, 33, 125, 126, 127, 129, 130, 131, 133, 255, 257, 383, 385)]
    shapes += [(r, 130) for r in (4, 5, 6, 7, 8, 9, 10)] + [(300, 381), (129, 129), (6, 515)]
    plots = [(rng.uniform(size=(M, N)) < dens).astype(np.uint8) for (M, N) in shapes for dens in (0.08, 0.5, 0.9)]
    for cpl in (4, 8, 16):
        ctx.set_option("dp_max_cpl", cpl)
        try:
            out = ctx.dp_packed(plots, want_sw=False)
            outc = ctx.dp_packed(plots, flags=_native.FLAG_DMAX_CLEAN, want_sw=False)
        finally:
            ctx.set_option("dp_max_cpl", 4)
        for b, B in enumerate(plots):
            q, d = dp.qmax_dmax_serra09(B)
            assert (out[b, 0], out[b, 1]) == (q, d), (cpl, B.shape, b % 3)
            assert outc[b, 1] == dp.dmax_clean(B), (cpl, B.shape, b % 3)


def test_mfcc_path_no_oti(ctx):
    # Serra09.py:177-184: 13 bins, no OTI / roll
    rng = np.random.default_rng(13)
    songs = []
    for L in (180, 240, 310):
        fr = rng.normal(size=(L, 13)).cumsum(axis=0).astype(np.float32) * 0.1 + rng.normal(size=(L, 13)).astype(np.float32)
        songs.append({"chroma": np.ascontiguousarray(fr.T), "gchroma": None})
    from oracle import crp_oracle, dp
    ctx.upload_songs(songs)
    pairs = np.array([(0, 1), (1, 2), (2, 0)], np.int32)
    r = ctx.serra09_pairs(pairs, flags=_native.FLAG_NO_OTI | _native.FLAG_NO_SWAP, want_crp=True)
    for p, (i, j) in enumerate(pairs):
        _, B = crp_oracle.crp_pair(songs[i]["chroma"].astype(np.float64), songs[j]["chroma"].astype(np.float64),
                                   None, None, 9, 0.095, do_oti=False)
        assert np.array_equal(r["crp"][p], B)
        assert (r["qmax_raw"][p], r["dmax_raw"][p]) == dp.qmax_dmax_serra09(B)


def test_full_size_symmetry_property(ctx):
    # whatisacover-shaped lengths (BASELINE configs[1]); no oracle at this size
    songs, _ = synth.make_song_set("whatisacover", seed=10000, n_songs=400)
    assert pc.check_symmetry(ctx, songs, n_pairs=2000) > 1500


def test_float64_input(ctx):
    songs, _ = synth.make_song_set("covers80", seed=5, n_songs=4, length_range=(100, 200))
    rng = np.random.default_rng(0)
    for s in songs:  # values that are NOT float32-representable (crema is float64)
        s["chroma"] = s["chroma"].astype(np.float64) + rng.uniform(0, 1e-9, size=s["chroma"].shape)
        s["gchroma"] = synth.global_chroma_f64(s["chroma"].T)
    pairs = np.array([(0, 1), (2, 3), (1, 3)], np.int32)
    pc.check_against_oracle(ctx, songs, pairs, flags=_native.FLAG_NO_SWAP)


def test_errors(ctx):
    songs, _ = synth.make_song_set("covers80", seed=5, n_songs=2, length_range=(50, 60))
    ctx.upload_songs(songs)
    with pytest.raises(_native.AcossError):
        ctx.serra09_pairs(np.array([[0, 7]], np.int32))
    with pytest.raises(_native.AcossError):
        ctx.serra09_pairs(np.array([[0, 1]], np.int32), kappa=0.0)


def test_sweep_generations_agree_and_exact_path(ctx):
    """The integer-key sweeps (gen 4: SIMT, gen 7: FP64 tensor cores + chained selection) against the
    float64-selection sweep (gen 2) and the oracle; then with the safety gap G inflated so that
    many rows take the float64 fallback (flagged separators, band patches) instead of the
    integer compare: the CRP must not move."""
    songs, _ = synth.make_song_set("covers80", seed=44, n_songs=10, length_range=(130, 900))
    rng = np.random.default_rng(2)
    pairs = rng.integers(0, len(songs), size=(40, 2)).astype(np.int32)
    pairs = pairs[pairs[:, 0] != pairs[:, 1]][:24]
    ctx.upload_songs(songs)
    try:
        ctx.set_option("sweep_gen", 2)
        base = ctx.serra09_pairs(pairs, want_crp=True)
        for gen in (4, 7):
            for gap in (2, 256, 1 << 14, 1 << 18):
                ctx.set_option("sweep_gen", gen)
                ctx.set_option("gap_units", gap)
                r = ctx.serra09_pairs(pairs, want_crp=True)
                where = "gen %d gap %d" % (gen, gap)
                assert np.array_equal(r["oti"], base["oti"]), where
                for a, b in zip(r["crp"], base["crp"]):
                    assert np.array_equal(a, b), where + ": %d CRP bits differ" % int((a != b).sum())
                assert np.array_equal(r["qmax_raw"], base["qmax_raw"]) and np.array_equal(r["dmax_raw"], base["dmax_raw"]), where
    finally:
        ctx.set_option("sweep_gen", 7)
        ctx.set_option("gap_units", 2)
    # and the default configuration against the oracle on the same pairs
    n, tied = pc.check_against_oracle(ctx, songs, pairs[:10], flags=_native.FLAG_NO_SWAP)
    assert tied <= 1


def test_generations_agree_at_full_size(ctx):
    # whatisacover-shaped lengths, 3000 pairs: the three sweep generations must give identical maxima
    songs, _ = synth.make_song_set("whatisacover", seed=10000, n_songs=300)
    rng = np.random.default_rng(7)
    pairs = rng.integers(0, len(songs), size=(3200, 2)).astype(np.int32)
    pairs = pairs[pairs[:, 0] != pairs[:, 1]][:3000]
    ctx.upload_songs(songs)
    out = {}
    try:
        for gen in (2, 4, 7):
            ctx.set_option("sweep_gen", gen)
            out[gen] = ctx.serra09_pairs(pairs)
    finally:
        ctx.set_option("sweep_gen", 7)
    for gen in (4, 7):
        assert np.array_equal(out[gen]["qmax_raw"], out[2]["qmax_raw"]), gen
        assert np.array_equal(out[gen]["dmax_raw"], out[2]["dmax_raw"]), gen
