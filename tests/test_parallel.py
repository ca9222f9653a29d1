"""CPU tests of the multi-GPU plumbing: tile sharding logic, and a world_size-2 gloo run of the
song broadcast + score gather (the compute is replaced by a deterministic stand-in)."""
import os
import subprocess
import sys

import numpy as np

from acoss_b200 import parallel

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_tiles_partition_all_pairs_exactly_once():
    n, T = 37, 8
    tiles = parallel.tile_list(n, T)
    seen = set()
    for t in tiles:
        for i, j in parallel.tile_pairs(n, T, t):
            assert i < j and (i, j) not in seen
            seen.add((int(i), int(j)))
    assert len(seen) == n * (n - 1) // 2


def test_tile_costs_and_balance():
    rng = np.random.default_rng(0)
    lengths = rng.integers(150, 1500, size=300)
    T, m = 16, 9
    tiles = parallel.tile_list(300, T)
    costs = parallel.tile_costs(lengths, T, tiles, m)
    w = np.maximum(lengths - m + 1, 0).astype(np.float64)
    for t in (0, 5, len(tiles) - 1):
        pairs = parallel.tile_pairs(300, T, tiles[t])
        assert np.isclose(costs[t], (w[pairs[:, 0]] * w[pairs[:, 1]]).sum())
    for world in (2, 4, 8):
        shares, load = parallel.assign_tiles(costs, world)
        assert sorted(np.concatenate(shares).tolist()) == list(range(len(tiles)))
        assert load.max() / load.mean() < 1.05
    allp = np.concatenate([parallel.shard_pairs(300, lengths, 4, r, T=T) for r in range(4)])
    assert len(allp) == 300 * 299 // 2 and len({(int(a), int(b)) for a, b in allp}) == len(allp)


WORKER = r'''
import os, sys
import numpy as np
sys.path.insert(0, %r)
sys.path.insert(0, os.path.join(%r, "tests", "emu"))
from acoss_b200 import parallel, synth, _native
from acoss_b200.serra09 import Serra09
import build_emu
# TEST INFRASTRUCTURE: the CPU SIMT-emulator build of the same kernel sources stands in for the GPU library
_native.DEFAULT_LIB = build_emu.build()
rank, world, local, dev = parallel.init_distributed(backend="gloo")
out_dir = sys.argv[1]
songs, labels = synth.make_song_set("covers80", seed=3, n_songs=9, length_range=(24, 44))
os.chdir(out_dir)
alg = Serra09(songs=songs, labels=labels, shortname="dist", do_memmaps=(rank == 0), do_mfcc=False, cachedir="cache_dist")
info = parallel.all_pairwise_distributed(alg, T=4)
if rank == 0:
    np.savez(os.path.join(out_dir, "dist.npz"), q=np.asarray(alg.Ds["chroma_qmax"]), d=np.asarray(alg.Ds["chroma_dmax"]),
             imb=info["lpt_imbalance"], n_local=info["pairs_local"])
alg.close()
import torch.distributed as dist
dist.destroy_process_group()
'''


def test_two_rank_gloo_all_pairwise_distributed(tmp_path):
    """world_size 2 over gloo: every rank packs its half of the songs, the arena is all-gathered, the tiles are
    scored by both ranks (through the emulator build of the kernels) and the float32 score tiles gathered on rank 0
    must give the matrix the single-process all_pairwise gives."""
    sys.path.insert(0, os.path.join(ROOT, "tests", "emu"))
    import build_emu
    from acoss_b200 import _native, synth
    from acoss_b200.serra09 import Serra09
    lib = build_emu.build()
    script = tmp_path / "worker.py"
    script.write_text(WORKER % (ROOT, ROOT))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", str(script), str(tmp_path)]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-3000:]
    z = np.load(str(tmp_path / "dist.npz"))
    # single process, same emulator library
    songs, labels = synth.make_song_set("covers80", seed=3, n_songs=9, length_range=(24, 44))
    old, cwd = _native.DEFAULT_LIB, os.getcwd()
    _native.DEFAULT_LIB = lib
    os.chdir(str(tmp_path))
    try:
        alg = Serra09(songs=songs, labels=labels, shortname="single", do_memmaps=True, do_mfcc=False, cachedir="cache_single")
        alg.all_pairwise(symmetric=True)
        q, d = np.asarray(alg.Ds["chroma_qmax"]).copy(), np.asarray(alg.Ds["chroma_dmax"]).copy()
        alg.close()
    finally:
        _native.DEFAULT_LIB = old
        os.chdir(cwd)
    assert q.any() and np.array_equal(z["q"], q) and np.array_equal(z["d"], d)
    assert np.array_equal(q, q.T) and not np.diag(q).any()
    assert 0 < int(z["n_local"]) < 36 and float(z["imb"]) >= 1.0
    assert os.path.exists(str(tmp_path / "cache_dist" / "Serra09_dist_Ds.h5"))


def test_place_tiles_matches_fancy_indexing():
    n, T, world = 23, 5, 3
    rng = np.random.default_rng(1)
    lengths = rng.integers(30, 90, size=n)
    tiles, shares, load = parallel.shard_tiles(n, lengths, world, T=T)
    ref = np.zeros((n, n), np.float32)
    per_rank = []
    for r in range(world):
        pr = [parallel.tile_pairs(n, T, tiles[t]) for t in shares[r]]
        pr = np.concatenate(pr) if pr else np.zeros((0, 2), np.int32)
        v = (pr[:, 0] * 100 + pr[:, 1]).astype(np.float32)
        ref[pr[:, 0], pr[:, 1]] = v
        per_rank.append(v[None, :])
        assert sum(parallel.tile_pair_count(n, T, tiles[t]) for t in shares[r]) == len(pr)
    Ds = {"k": np.zeros((n, n), np.float32)}
    parallel.place_tiles(Ds, ["k"], n, T, tiles, shares, per_rank)
    assert np.array_equal(Ds["k"], ref)
