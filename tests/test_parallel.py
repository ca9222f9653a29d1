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
from acoss_b200 import parallel, synth, _native
rank, world, local, dev = parallel.init_distributed(backend="gloo")
packed = None
if rank == 0:
    songs, _ = synth.make_song_set("covers80", seed=3, n_songs=10, length_range=(20, 40))
    packed = _native.pack_songs(songs)
packed = parallel.broadcast_songs(packed, src=0, device=dev)
frames, offs, lens, g = packed
mine = parallel.shard_pairs(len(lens), lens, world, rank, T=4, m=9)
# stand-in score: a deterministic function of the shared data, so a wrong broadcast shows up
score = np.array([frames[offs[i]:offs[i] + lens[i]].sum() * 1e-3 + g[j].sum() for i, j in mine])
idxs, sc = parallel.gather_scores(mine, {"chroma_qmax": score, "chroma_dmax": 2 * score}, dst=0, device=dev)
if rank == 0:
    np.savez(sys.argv[1], idxs=idxs, q=sc["chroma_qmax"], d=sc["chroma_dmax"], frames=frames, offs=offs, lens=lens, g=g)
import torch.distributed as dist
dist.barrier()
dist.destroy_process_group()
'''


def test_two_rank_gloo_broadcast_and_gather(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % ROOT)
    out = str(tmp_path / "out.npz")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", str(script), out]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    z = np.load(out)
    n = len(z["lens"])
    got = {(int(i), int(j)): (q, d) for (i, j), q, d in zip(z["idxs"], z["q"], z["d"])}
    assert len(got) == n * (n - 1) // 2
    for (i, j), (q, d) in got.items():
        ref = z["frames"][z["offs"][i]:z["offs"][i] + z["lens"][i]].sum() * 1e-3 + z["g"][j].sum()
        assert np.isclose(q, ref) and np.isclose(d, 2 * ref)
