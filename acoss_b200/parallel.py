"""Multi-GPU plumbing for the all-pairs run: one process per GPU (torchrun), songs shared once,
the upper-triangular pair matrix cut into T x T song tiles that are dealt to ranks by estimated
cost, score tiles gathered back on rank 0.  There is no per-pair communication and no
data-path collective: pairs are independent units.

torch.distributed is used only as the transport (NCCL over NVLink/NVSwitch on the GPU box, gloo
in the CPU tests); the compute goes through the C ABI on each rank's own device.
"""
from __future__ import annotations

import os

import numpy as np


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def init_distributed(backend=None):
    """Initialise torch.distributed from the torchrun environment.  Returns (rank, world, local_rank, device)."""
    import torch
    import torch.distributed as dist
    rank, world, local = dist_env()
    use_cuda = torch.cuda.is_available()
    if use_cuda:
        torch.cuda.set_device(local)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        dist.init_process_group(backend=backend or ("nccl" if use_cuda else "gloo"), rank=rank, world_size=world)
    device = torch.device("cuda", local) if use_cuda else torch.device("cpu")
    return rank, world, local, device


def _bcast_array(arr, meta_src, src, device):
    """Broadcast a numpy array from `src` (other ranks pass arr=None)."""
    import torch
    import torch.distributed as dist
    rank = dist.get_rank()
    meta = [None]
    if rank == src:
        meta = [(arr.shape, str(arr.dtype))]
    dist.broadcast_object_list(meta, src=src)
    shape, dtype = meta[0]
    if rank == src:
        t = torch.from_numpy(np.ascontiguousarray(arr)).to(device)
    else:
        t = torch.empty(shape, dtype=getattr(torch, dtype), device=device)
    dist.broadcast(t, src=src)
    return t.cpu().numpy()


def broadcast_songs(packed, src=0, device=None):
    """Share the packed song arena (frames, offsets, lengths, gchroma) from rank `src` with every
    rank -- the one feature collective of a run.  Single-process: returns `packed` unchanged."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return packed
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    rank = dist.get_rank()
    has_g = [packed[3] is not None] if rank == src else [None]
    dist.broadcast_object_list(has_g, src=src)
    out = []
    for k in range(4):
        if k == 3 and not has_g[0]:
            out.append(None)
            continue
        out.append(_bcast_array(packed[k] if rank == src else None, None, src, device))
    return tuple(out)


# ---------------------------------------------------------------------------- tiling
def tile_list(n_songs, T):
    """Upper-triangular (incl. diagonal) T x T song tiles of the pair matrix, row-major."""
    nt = (n_songs + T - 1) // T
    return [(a, b) for a in range(nt) for b in range(a, nt)]


def tile_pairs(n_songs, T, tile):
    """The (i < j) pairs inside one tile, in CoverAlgorithm.all_pairwise order (row-major)."""
    a, b = tile
    i0, i1 = a * T, min((a + 1) * T, n_songs)
    j0, j1 = b * T, min((b + 1) * T, n_songs)
    I, J = np.meshgrid(np.arange(i0, i1), np.arange(j0, j1), indexing="ij")
    keep = I < J
    return np.stack([I[keep], J[keep]], axis=1).astype(np.int32)


def tile_costs(lengths, T, tiles, m=9):
    """Estimated work of each tile: sum over its pairs of M*N (cells of the cross-recurrence plot)."""
    w = np.maximum(np.asarray(lengths, dtype=np.float64) - m + 1, 0.0)
    n = len(w)
    nt = (n + T - 1) // T
    s = np.array([w[a * T:(a + 1) * T].sum() for a in range(nt)])
    s2 = np.array([(w[a * T:(a + 1) * T] ** 2).sum() for a in range(nt)])
    costs = []
    for a, b in tiles:
        if a != b:
            costs.append(s[a] * s[b])
        else:
            costs.append(0.5 * (s[a] ** 2 - s2[a]))
    return np.array(costs)


def assign_tiles(costs, world):
    """Longest-processing-time greedy: tiles by decreasing cost to the least-loaded rank.
    Returns a list of tile-index arrays, one per rank (deterministic on every rank)."""
    order = np.argsort(-costs, kind="stable")
    load = np.zeros(world)
    owner = np.zeros(len(costs), dtype=np.int64)
    for t in order:
        r = int(np.argmin(load))
        owner[t] = r
        load[r] += costs[t]
    return [np.nonzero(owner == r)[0] for r in range(world)], load


def shard_pairs(n_songs, lengths, world, rank, T=128, m=9):
    """This rank's share of the all-pairs list: concatenated pairs of its tiles."""
    tiles = tile_list(n_songs, T)
    shares, _ = assign_tiles(tile_costs(lengths, T, tiles, m), world)
    mine = [tile_pairs(n_songs, T, tiles[t]) for t in shares[rank]]
    if not mine:
        return np.zeros((0, 2), np.int32)
    return np.concatenate(mine, axis=0)


def gather_scores(idxs, scores, dst=0, device=None):
    """Gather per-rank (idxs (P,2) int32, scores {key: float64 (P,)}) on rank `dst`.
    Returns (idxs_all, scores_all) on dst, (None, None) elsewhere.  Single-process: identity."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return idxs, scores
    world, rank = dist.get_world_size(), dist.get_rank()
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    keys = sorted(scores.keys())
    n_local = torch.tensor([idxs.shape[0]], dtype=torch.int64, device=device)
    counts = [torch.zeros(1, dtype=torch.int64, device=device) for _ in range(world)]
    dist.all_gather(counts, n_local)
    counts = [int(c.item()) for c in counts]
    nmax = max(max(counts), 1)
    # one padded float64 block per rank: [i, j, score_0, score_1, ...]
    block = np.zeros((nmax, 2 + len(keys)), dtype=np.float64)
    block[:idxs.shape[0], 0:2] = idxs
    for k, key in enumerate(keys):
        block[:idxs.shape[0], 2 + k] = scores[key]
    t = torch.from_numpy(block).to(device)
    outs = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(outs, t)  # NCCL has no gather-to-one on every version; all_gather is equivalent here
    if rank != dst:
        return None, None
    parts = [o.cpu().numpy()[:c] for o, c in zip(outs, counts)]
    allb = np.concatenate(parts, axis=0)
    idxs_all = allb[:, 0:2].astype(np.int32)
    scores_all = {key: allb[:, 2 + k].copy() for k, key in enumerate(keys)}
    return idxs_all, scores_all


def all_pairwise_distributed(alg, T=128, symmetric=True):
    """The multi-GPU form of CoverAlgorithm.all_pairwise (CoverAlgorithm.py:138-184): every rank
    scores the pairs of its tiles, rank 0 assembles Ds, symmetrises and saves '<prefix>_Ds.h5'."""
    import torch.distributed as dist
    from . import hdf5io
    rank, world, _ = dist_env()
    lengths = np.array([alg.load_features(i)['chroma'].shape[1] for i in range(alg.N)])
    mine = shard_pairs(alg.N, lengths, world, rank, T=T, m=alg.m)
    sims = alg.similarity(mine)
    idxs_all, sims_all = gather_scores(mine, sims)
    if rank == 0:
        for key, v in sims_all.items():
            if key in alg.Ds:
                alg.Ds[key][idxs_all[:, 0], idxs_all[:, 1]] = v
        alg.get_all_clique_ids()
        if symmetric:
            for key in alg.Ds:
                alg.Ds[key] += alg.Ds[key].T
        hdf5io.save("%s_Ds.h5" % alg.get_cacheprefix(), {k: np.asarray(v) for k, v in alg.Ds.items()})
    if world > 1 and dist.is_initialized():
        dist.barrier()
    return sims
