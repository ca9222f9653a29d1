"""2-GPU test of the multi-GPU all-pairs path over NCCL: parallel.all_pairwise_distributed on 200 songs under torchrun
(songs all-gathered device to device, tiles dealt by cost, float32 score tiles gathered on rank 0) must give the matrix
the single-GPU CoverAlgorithm.all_pairwise gives, bit for bit.  Skipped on a box with fewer than two GPUs."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
import numpy as np
sys.path.insert(0, %r)
from acoss_b200 import parallel, synth
from acoss_b200.serra09 import Serra09
rank, world, local, dev = parallel.init_distributed()
out_dir = sys.argv[1]
os.chdir(out_dir)
songs, labels = synth.make_song_set("benchmark", seed=15000, n_songs=200)
alg = Serra09(songs=songs, labels=labels, shortname="dist", do_memmaps=(rank == 0), do_mfcc=False, cachedir="cache_dist",
              device=local)
info = parallel.all_pairwise_distributed(alg, T=16)
if rank == 0:
    np.savez(os.path.join(out_dir, "dist.npz"), q=np.asarray(alg.Ds["chroma_qmax"]), d=np.asarray(alg.Ds["chroma_dmax"]),
             imb=info["lpt_imbalance"], n_local=info["pairs_local"])
alg.close()
import torch.distributed as dist
dist.destroy_process_group()
'''


def test_two_gpu_all_pairwise_distributed_matches_single_gpu(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from acoss_b200 import synth
    from acoss_b200.serra09 import Serra09
    script = tmp_path / "worker.py"
    script.write_text(WORKER % ROOT)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29541", str(script), str(tmp_path)]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-3000:]
    z = np.load(str(tmp_path / "dist.npz"))
    songs, labels = synth.make_song_set("benchmark", seed=15000, n_songs=200)
    cwd = os.getcwd()
    os.chdir(str(tmp_path))
    try:
        alg = Serra09(songs=songs, labels=labels, shortname="single", do_memmaps=True, do_mfcc=False, cachedir="cache_single")
        alg.all_pairwise(symmetric=True)
        q, d = np.asarray(alg.Ds["chroma_qmax"]).copy(), np.asarray(alg.Ds["chroma_dmax"]).copy()
        alg.close()
    finally:
        os.chdir(cwd)
    assert q.any() and np.array_equal(z["q"], q) and np.array_equal(z["d"], d)
    assert 0 < int(z["n_local"]) < 19900 and 1.0 <= float(z["imb"]) < 1.05
