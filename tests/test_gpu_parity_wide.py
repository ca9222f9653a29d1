"""The wide parity gate of SURVEY.md 8(d): CRP bits / OTI / Qmax / Dmax of the CUDA path (through the C ABI) against the
oracle on (i) 10 000 uniformly drawn pairs of the bench workload (whatisacover-shaped, BASELINE configs[1]), (ii) all
12 720 pairs of the covers80-shaped set (configs[0]), (iii) every kernel-dispatch bucket at m = 9, (iv) the non-mutual /
clean-table flags through acoss_serra09_pairs, (v) a benchmark-shaped sample with cliques of 13 (configs[2]) and (vi) the
13-bin mfcc half of Serra09.similarity at bench lengths.  The oracle runs on all host cores (tests/wide_oracle.py)."""
import itertools

import numpy as np
import pytest

import wide_oracle as wo
from acoss_b200 import _native

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = _native.Context(0)
    yield c
    c.close()


def _run(ctx, spec, pairs, flag_sets, do_oti=True, nonmutual=False, max_tied=0):
    songs = wo.build_songs(spec)
    pool = wo.OraclePool(spec)
    try:
        rows = pool.run(pairs, do_oti=do_oti, nonmutual=nonmutual)
    finally:
        pool.close()
    total_tied = total_diff = 0
    for flags in flag_sets:
        n, tied, tied_diff = wo.compare(ctx, songs, pairs, rows, flags=flags)
        assert n == len(pairs)
        total_tied = max(total_tied, tied)
        total_diff = max(total_diff, tied_diff)
    print("\n[wide parity] %s: %d pairs x %d flag sets bit-exact; %d tied pairs (relative neighbour gap < 1e-9) were "
          "only counted, %d of them differ" % (spec[0:2], len(pairs) - total_tied, len(flag_sets), total_tied, total_diff))
    # pairs too close to call (long songs have thousands of rows and columns, each with its own k-th / (k+1)-th gap):
    # at most 1 % of the pairs, and at most 0.2 % may actually differ from the oracle's choice
    assert total_tied <= max(max_tied, len(pairs) // 100)
    assert total_diff <= max(max_tied, len(pairs) // 500)
    return total_tied


def test_whatisacover_10000_pairs(ctx):
    spec = ("set", "whatisacover", {"seed": 10000})
    rng = np.random.default_rng(123)
    ij = rng.integers(0, 10000, size=(10400, 2)).astype(np.int32)
    ij = ij[ij[:, 0] != ij[:, 1]][:10000]
    _run(ctx, spec, ij, (0,), max_tied=2)


def test_covers80_all_12720_pairs(ctx):
    spec = ("set", "covers80", {"seed": 80})
    ij = np.array(list(itertools.combinations(range(160), 2)), np.int32)
    assert len(ij) == 12720
    _run(ctx, spec, ij, (0,), max_tied=2)


def test_dispatch_buckets_m9(ctx):
    """>= 200 pairs per column-song length class: the eight capacities of the default sweep (128, 256, ..., 1024
    columns), the integer-key SIMT kernels (1016..1320), the float64 kernel (1321..1408) and the tiled path (> 1408);
    with and without the orientation swap so that both the separator sweep and the CRP sweep see every class."""
    ranges = [(60, 119, 8), (120, 247, 8), (248, 375, 8), (376, 503, 8), (504, 631, 8), (632, 759, 8), (760, 887, 8),
              (888, 1015, 8), (1016, 1320, 8), (1321, 1408, 6), (1409, 1700, 4)]
    spec = ("ranges", 900, ranges)
    starts = np.cumsum([0] + [c for _, _, c in ranges])
    rng = np.random.default_rng(5)
    pairs = []
    for b, (_, _, cnt) in enumerate(ranges):
        own = np.arange(starts[b], starts[b + 1])
        for t in range(200):
            i = int(rng.choice(own))
            j = int(rng.integers(0, starts[-1]))
            while j == i:
                j = int(rng.integers(0, starts[-1]))
            pairs.append((i, j) if t % 2 else (j, i))
    pairs = np.array(pairs, np.int32)
    _run(ctx, spec, pairs, (0, _native.FLAG_NO_SWAP), max_tied=2)


def test_nonmutual_and_clean_flags_through_serra09_pairs(ctx):
    spec = ("set", "covers80", {"seed": 81, "n_songs": 60, "length_range": (150, 900)})
    rng = np.random.default_rng(9)
    ij = rng.integers(0, 60, size=(340, 2)).astype(np.int32)
    ij = ij[ij[:, 0] != ij[:, 1]][:300]
    # ChenFusion / Tempogram style: row kNN only (CRPUtils.py:169-199) and a zeroed Dmax table (ChenFusion.py:66)
    nm = _native.FLAG_NON_MUTUAL | _native.FLAG_DMAX_CLEAN
    _run(ctx, spec, ij, (nm, nm | _native.FLAG_NO_SWAP), nonmutual=True, max_tied=2)
    _run(ctx, spec, ij, (_native.FLAG_DMAX_CLEAN,), max_tied=2)


def test_benchmark_shaped_sample_cliques_of_13(ctx):
    spec = ("set", "benchmark", {"seed": 15000, "n_songs": 650})  # 50 cliques of 13
    from acoss_b200 import synth
    _, labels = synth.make_song_set("benchmark", seed=15000, n_songs=650)
    labels = np.array(labels)
    rng = np.random.default_rng(3)
    within = []
    for lab in np.unique(labels)[:40]:
        idx = np.flatnonzero(labels == lab)
        for a, b in itertools.combinations(idx, 2):
            within.append((a, b))
    within = np.array(within, np.int32)
    within = within[rng.permutation(len(within))[:700]]
    rnd = rng.integers(0, 650, size=(340, 2)).astype(np.int32)
    rnd = rnd[rnd[:, 0] != rnd[:, 1]][:300]
    _run(ctx, spec, np.concatenate([within, rnd]), (0,), max_tied=2)


def test_mfcc_half_at_bench_lengths(ctx):
    spec = ("mfcc", 77, 120, 450.0)
    rng = np.random.default_rng(4)
    ij = rng.integers(0, 120, size=(560, 2)).astype(np.int32)
    ij = ij[ij[:, 0] != ij[:, 1]][:500]
    _run(ctx, spec, ij, (_native.FLAG_NO_OTI, _native.FLAG_NO_OTI | _native.FLAG_NO_SWAP), do_oti=False, max_tied=2)
