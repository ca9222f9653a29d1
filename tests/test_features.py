"""Feature preparation on the device against the numpy restatement of
Serra09.load_features: bit-exact for float32 (hpcp, mfcc_htk) and float64 (crema)."""
import numpy as np
import pytest

from oracle import features_oracle as fo


def _songs(rng, dtype, n_list, bins=12):
    out = []
    for n in n_list:
        a = rng.random((n, bins)).astype(dtype) ** 2
        a[rng.random((n, bins)) < 0.05] = 0  # repeated values: ties inside a block
        out.append(a)
    return out


def test_sync_closed_form_equals_restated_librosa_slicing():
    """librosa.util.sync's slicing (fix_frames + index_to_slice, restated in oracle/features_oracle.py) against the
    closed form the kernels are tested with, on the edge cases: n a multiple of fac, n < fac, n = fac, n = fac + 1."""
    rng = np.random.default_rng(5)
    for n in (1, 7, 39, 40, 41, 79, 80, 81, 120, 333, 4000):
        a = rng.random((12, n))
        for agg in (np.median, np.mean):
            lit = fo.sync_slices(a, np.arange(0, n, 40), agg)
            assert lit.shape[1] == -(-n // 40)
            assert np.array_equal(lit, fo.sync(a, 40, agg)), (n, agg.__name__)
    assert list(fo.fix_frames(np.arange(0, 80, 40), 0, 80)) == [0, 40, 80]
    assert list(fo.fix_frames(np.arange(0, 30, 40), 0, 30)) == [0, 30]
    assert list(fo.fix_frames(np.arange(0, 81, 40), 0, 81)) == [0, 40, 80, 81]


def test_oracle_sync_matches_numpy_semantics():
    rng = np.random.default_rng(0)
    a = rng.random((12, 95)).astype(np.float32)
    s = fo.sync(a, 40, np.median)
    assert s.shape == (12, 3) and s.dtype == np.float32
    assert np.array_equal(s[:, 0], np.median(a[:, 0:40], axis=1))
    assert np.array_equal(s[:, 2], np.median(a[:, 80:95], axis=1))
    g = fo.global_chroma(a.T)
    assert g.max() == 1.0
    with pytest.raises(IOError):
        fo.global_chroma(np.zeros((5, 13)))


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_chroma_median_and_gchroma_bit_exact(dtype):
    from acoss_b200 import _native
    ctx = _native.Context(0)
    rng = np.random.default_rng(3)
    songs = _songs(rng, dtype, [1, 7, 39, 40, 41, 80, 1234, 17999, 18000])
    agg, g = ctx.aggregate_songs(songs, fac=40, aggregate="median", want_gchroma=True)
    for s, (a, got) in enumerate(zip(songs, agg)):
        ref = fo.sync(a.T, 40, np.median)
        assert got.dtype == dtype and got.shape == ref.shape
        assert np.array_equal(got, ref), "song %d" % s
        assert np.array_equal(g[s], fo.global_chroma(a)), "gchroma %d" % s
    # other block sizes, odd sizes
    for fac in (1, 2, 5, 33, 64):
        agg = ctx.aggregate_songs(songs[:7], fac=fac)
        for a, got in zip(songs, agg):
            assert np.array_equal(got, fo.sync(a.T, fac, np.median)), fac


@pytest.mark.gpu
def test_mfcc_mean_bit_exact_and_nonfinite():
    from acoss_b200 import _native
    ctx = _native.Context(0)
    rng = np.random.default_rng(4)
    songs = []
    for n in (3, 40, 47, 999, 5000):
        a = (rng.standard_normal((13, n)) * 30).astype(np.float32)
        a[0, 0] = np.nan
        if n > 45:
            a[5, 44] = np.inf
            a[6, 45] = -np.inf
        songs.append(a)
    agg = ctx.aggregate_songs(songs, fac=40, aggregate="mean", layout="bins_x_frames", zero_nonfinite=True)
    for a, got in zip(songs, agg):
        m = np.array(a)
        m[np.isnan(m)] = 0
        m[np.isinf(m)] = 0
        assert np.array_equal(got, fo.sync(m, 40, np.mean))


@pytest.mark.gpu
def test_aggregate_errors():
    from acoss_b200 import _native
    ctx = _native.Context(0)
    with pytest.raises(_native.AcossError):
        ctx.aggregate_songs([np.zeros((10, 12), np.float32)], fac=65)
    with pytest.raises(_native.AcossError):
        ctx.aggregate_songs([np.zeros((12, 10), np.float32)], layout="bins_x_frames", want_gchroma=True)
