"""SURVEY.md 8f row F2 -- getEvalStatistics (CoverAlgorithm.py:330-418).
CPU: the oracle against the golden numbers the reference's own function produced, the host
bookkeeping, and the kernel source under the SIMT emulator.  GPU (-m gpu): the real library
through the C ABI against the oracle, ties, large cliques, the plugin method and its CSV."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "emu"))

from acoss_b200 import _native  # noqa: E402
from oracle import eval_oracle as eo  # noqa: E402

EVAL_GOLDEN = os.path.join(HERE, "golden", "eval_golden.npz")


def golden_cases():
    G = np.load(EVAL_GOLDEN, allow_pickle=False)
    for tag in G["tags"].tolist():
        sizes = G["%s_clique_sizes" % tag]
        songs = G["%s_clique_songs" % tag]
        cliques, at = [], 0
        for K in sizes:
            cliques.append([int(s) for s in songs[at:at + K]])
            at += K
        yield tag, G["%s_D" % tag], cliques, G["%s_stats" % tag], G["%s_tops" % tag]


def same_float(a, b):
    return (np.isnan(a) and np.isnan(b)) or a == b


def test_oracle_matches_reference_golden():
    n = 0
    for tag, D, cliques, stats, tops in golden_cases():
        for fast in (False, True):
            with np.errstate(all="ignore"):
                import warnings
                with warnings.catch_warnings():
                    warnings.simplefilter("ignore")
                    MR, MRR, MDR, MAP, t = eo.eval_statistics(D, cliques, fast=fast)
            for got, want in zip((MR, MRR, MDR, MAP), stats):
                assert same_float(got, want), (tag, fast, got, want)   # bit-exact float64
            assert np.array_equal(t, tops), tag
        n += 1
    assert n == 6


def test_eval_layout():
    cliques = [[4], [0, 7, 2], [5, 1], [3], [6, 8, 9, 10]]
    order, Ks, cstart, csize, offs, rows, total = _native.eval_layout(cliques)
    assert list(Ks) == [4, 3, 2, 1, 1]
    assert list(order[:9]) == [6, 8, 9, 10, 0, 7, 2, 5, 1] and sorted(order[9:]) == [3, 4]
    assert rows == 9 and total == 4 * 3 + 3 * 2 + 2 * 1
    assert list(cstart[:rows]) == [0] * 4 + [4] * 3 + [7] * 2
    assert list(csize[:rows]) == [4] * 4 + [3] * 3 + [2] * 2
    assert list(offs[:rows]) == [0, 3, 6, 9, 12, 14, 16, 18, 19]
    order, Ks, cstart, csize, offs, rows, total = _native.eval_layout([[0], [1], [2]])
    assert rows == 0 and total == 0


def check_ctx_against_oracle(ctx, D, cliques, fast=True):
    iranks, Ks, rows = ctx.eval_ranks(D, cliques)
    _, _, per_row, oKs = (eo.eval_ranks_fast if fast else eo.eval_ranks)(D, cliques)
    assert np.array_equal(Ks, oKs)
    assert rows == len(per_row)
    for i in range(rows):
        assert np.array_equal(iranks[i], per_row[i]), i
    return rows


def random_case(rng, sizes, levels=None):
    n = int(np.sum(sizes))
    perm = rng.permutation(n)
    cliques, at = [], 0
    for K in sizes:
        cliques.append([int(x) for x in perm[at:at + K]])
        at += K
    if levels:   # few distinct scores: every row is full of ties
        D = rng.integers(0, levels, size=(n, n)).astype(np.float32) / np.float32(levels)
    else:
        D = rng.random((n, n)).astype(np.float32)
    D = np.triu(D, 1)
    return (D + D.T).astype(np.float32), cliques


def test_emu_eval_kernel():
    from build_emu import build
    ctx = _native.Context(0, build())
    try:
        for tag, D, cliques, stats, tops in golden_cases():
            check_ctx_against_oracle(ctx, D, cliques, fast=False)
        rng = np.random.default_rng(5)
        # ties (stable order decides), a clique larger than one member pass (16), odd sizes
        assert check_ctx_against_oracle(ctx, *random_case(rng, [3, 2, 2, 1, 1], levels=4)) == 7
        assert check_ctx_against_oracle(ctx, *random_case(rng, [37, 18, 17, 2, 1, 1, 1])) == 74
        assert check_ctx_against_oracle(ctx, *random_case(rng, [20, 5, 1], levels=7)) == 25
        # the one-compare-per-member kernel for large cliques (default: binary search + histogram)
        ctx.set_option("eval_direct", 1)
        try:
            assert check_ctx_against_oracle(ctx, *random_case(rng, [33, 13, 6, 1], levels=9)) == 52
        finally:
            ctx.set_option("eval_direct", 0)
        iranks, Ks, rows = ctx.eval_ranks(np.zeros((3, 3), np.float32), [[0], [1], [2]])
        assert rows == 0 and iranks == []
        with pytest.raises(_native.AcossError):
            ctx.lib.acoss_eval_ranks  # symbol exists
            bad = np.array([0, 0, 1], np.int32)   # not a permutation
            one = np.array([0], np.int32)
            two = np.array([2], np.int32)
            off = np.zeros(1, np.int64)
            out = np.zeros(4, np.int32)
            rc = ctx.lib.acoss_eval_ranks(ctx._h, np.zeros((3, 3), np.float32).ctypes.data, 3,
                                          _native._ptr(bad, _native._c_i32p), _native._ptr(one, _native._c_i32p),
                                          _native._ptr(two, _native._c_i32p), 1, _native._ptr(off, _native._c_i64p),
                                          _native._ptr(out, _native._c_i32p), 0)
            ctx._check(rc, "acoss_eval_ranks")
    finally:
        ctx.close()


# ------------------------------------------------------------------ GPU
@pytest.fixture(scope="module")
def gpu_ctx():
    c = _native.Context(0)
    yield c
    c.close()


@pytest.mark.gpu
def test_gpu_eval_golden_and_random(gpu_ctx):
    for tag, D, cliques, stats, tops in golden_cases():
        check_ctx_against_oracle(gpu_ctx, D, cliques, fast=False)
    rng = np.random.default_rng(6)
    assert check_ctx_against_oracle(gpu_ctx, *random_case(rng, [13] * 100 + [1] * 200)) == 1300
    assert check_ctx_against_oracle(gpu_ctx, *random_case(rng, [2] * 700 + [1] * 3)) == 1400
    assert check_ctx_against_oracle(gpu_ctx, *random_case(rng, [300, 40, 33, 17, 16, 2, 1], levels=50)) == 408
    assert check_ctx_against_oracle(gpu_ctx, *random_case(rng, [5] * 50 + [1] * 7, levels=3)) == 250
    gpu_ctx.set_option("eval_direct", 1)
    try:
        assert check_ctx_against_oracle(gpu_ctx, *random_case(rng, [13] * 60 + [40, 7, 1], levels=30)) == 827
    finally:
        gpu_ctx.set_option("eval_direct", 0)
    assert gpu_ctx.last_eval_ms() > 0


@pytest.mark.gpu
def test_gpu_eval_full_size_rows(gpu_ctx):
    """Benchmark-subset shape (15 000 songs: 1000 cliques of 13 + 2000 singletons):
    every rank of a sample of rows against numpy's stable argsort, and the whole set against
    size-independent properties (ranks strictly ascending, within 1..n-1)."""
    rng = np.random.default_rng(15000)
    sizes = [13] * 1000 + [1] * 2000
    n = 15000
    perm = rng.permutation(n)
    cliques, at = [], 0
    for K in sizes:
        cliques.append([int(x) for x in perm[at:at + K]])
        at += K
    D = rng.random((n, n), dtype=np.float32)
    D = np.round(D * 4096) / np.float32(4096)     # ~3.7 equal scores per value per row: ties everywhere
    iu = np.triu_indices(n, 1)
    D.T[iu] = D[iu]
    iranks, Ks, rows = gpu_ctx.eval_ranks(D, cliques)
    assert rows == 13000
    order = _native.eval_layout(cliques)[0]
    allr = np.stack(iranks)
    assert allr.shape == (13000, 12)
    assert np.all(np.diff(allr, axis=1) > 0) and allr.min() >= 1 and allr.max() <= n - 1
    for p in rng.choice(rows, size=40, replace=False):
        i = order[p]
        row = D[i, order].copy()
        row[p] = -np.inf
        o = np.argsort(-row, kind="stable")
        start = (p // 13) * 13
        want = (np.nonzero((o >= start) & (o < start + 13))[0] + 1)[:-1]
        assert np.array_equal(iranks[p], want), p


@pytest.mark.gpu
def test_gpu_plugin_eval_statistics(tmp_path, monkeypatch):
    """CoverAlgorithm.getEvalStatistics through the plugin class: numbers equal to the reference's
    own function (golden) bit for bit, CSV row in the reference's format (CoverAlgorithm.py:404-417)."""
    from acoss_b200.cover_algorithm import CoverAlgorithm
    monkeypatch.chdir(tmp_path)
    for tag, D, cliques, stats, tops in golden_cases():
        n = D.shape[0]
        alg = CoverAlgorithm(name="golden", shortname=tag, do_memmaps=False, songs=[None] * n)
        alg.Ds = {"main": D}
        alg.cliques = {"clique_%i" % i: set(c) for i, c in enumerate(cliques)}
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            MR, MRR, MDR, MAP, t = alg.getEvalStatistics("main")
        for got, want in zip((MR, MRR, MDR, MAP), stats):
            assert same_float(got, want), (tag, got, want)
        assert np.array_equal(t, tops)
        lines = open("results_%s.csv" % tag).read().splitlines()
        assert lines[0] == "name, MR, MRR, MDR, MAP,Top-1,Top-10,Top-100,Top-1000"
        assert lines[1].startswith("golden_main,")
