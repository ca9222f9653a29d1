"""Multi-GPU plumbing for the all-pairs run: one process per GPU (torchrun).

  * SONGS are all-gathered once: every rank loads / packs the features of its contiguous share of the songs,
    copies that shard to its device, and one all_gather over NCCL (NVLink / NVSwitch) gives every rank the whole
    arena in device memory; acoss_upload_songs takes the device pointer (device-to-device, no host bounce).
  * PAIRS: the upper-triangular pair matrix is cut into T x T song tiles that are dealt to ranks by estimated
    cost (sum of M * N), longest-processing-time greedy.  Pairs are independent: no data-path collective.
  * SCORES come back as float32 tiles: one gather of [n_types, n_pairs_of_rank] float32 to rank 0, which knows
    every rank's tile list and drops the tiles into the N x N matrices by slicing.

torch.distributed is used only as the transport (NCCL on the GPU box, gloo in the CPU tests); the compute goes
through the C ABI on each rank's own device (CoverAlgorithm.py:138-184 is the single-process
original).
"""
from __future__ import annotations

import os
import time

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
        kw = {}
        if use_cuda and (backend in (None, "nccl")):
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend=backend or ("nccl" if use_cuda else "gloo"), rank=rank, world_size=world, **kw)
    device = torch.device("cuda", local) if use_cuda else torch.device("cpu")
    return rank, world, local, device


def _is_dist():
    import torch.distributed as dist
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


def _default_device():
    import torch
    import torch.distributed as dist
    return torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")


def _allgather(t):
    """all_gather of equal-shaped tensors -> [world, *t.shape] (flat buffers: gloo and NCCL agree on those)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size()
    flat = t.contiguous().view(-1)
    out = torch.empty(world * flat.numel(), dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(out, flat)
    return out.view((world,) + tuple(t.shape))


def song_shares(n_songs, world):
    """Contiguous song-id ranges [s0, s1) per rank (the shard a rank loads and contributes to the all-gather)."""
    cuts = np.linspace(0, n_songs, world + 1).astype(np.int64)
    return [(int(cuts[r]), int(cuts[r + 1])) for r in range(world)]


# ---------------------------------------------------------------------------- songs: all-gather
class SharedArena:
    """The whole song arena on this rank's device after the all-gather.  frames: torch tensor [total_frames, bins]
    (float32 / float64) on `device`; gchroma: [n_songs, bins] float64 tensor or None; offsets / lengths: host arrays."""

    def __init__(self, frames, gchroma, offsets, lengths, seconds):
        self.frames, self.gchroma, self.offsets, self.lengths, self.seconds = frames, gchroma, offsets, lengths, seconds

    def upload_into(self, ctx):
        import torch
        if self.frames.is_cuda:
            torch.cuda.current_stream(self.frames.device).synchronize()   # the collective has finished
            ctx.upload_device(self.frames.data_ptr(), np.float32 if self.frames.dtype == torch.float32 else np.float64,
                              self.frames.shape[1], self.offsets, self.lengths,
                              None if self.gchroma is None else self.gchroma.data_ptr())
        else:  # gloo / CPU tests: the same call with host pointers
            ctx.upload(self.frames.numpy(), self.offsets, self.lengths, None if self.gchroma is None else self.gchroma.numpy())


def allgather_songs(shard, device=None):
    """shard = (frames [n_frames_local, bins] host array (pinned or not), lengths int32 [n_local], gchroma float64
    [n_local, bins] or None) of THIS rank's songs (ranks hold consecutive song-id ranges, rank order = song order).
    Returns a SharedArena.  One host->device copy of the local shard, one all_gather of lengths, one of frames, one of
    gchroma; nothing goes back to the host."""
    import torch
    import torch.distributed as dist
    t0 = time.perf_counter()
    frames, lengths, g = shard
    single = not _is_dist()
    if device is None:
        device = _default_device() if not single else (torch.device("cuda", torch.cuda.current_device())
                                                       if torch.cuda.is_available() else torch.device("cpu"))
    lengths = np.ascontiguousarray(lengths, dtype=np.int32)
    n_local = int(lengths.size)
    nf_local = int(lengths.astype(np.int64).sum())
    bins = int(frames.shape[1])
    tf = torch.from_numpy(np.ascontiguousarray(frames[:max(nf_local, 0)])).to(device, non_blocking=True)
    tg = None if g is None else torch.from_numpy(np.ascontiguousarray(g, dtype=np.float64)).to(device, non_blocking=True)
    if single:
        offsets = np.zeros(n_local, np.int64)
        if n_local > 1:
            offsets[1:] = np.cumsum(lengths[:-1].astype(np.int64))
        return SharedArena(tf, tg, offsets, lengths, time.perf_counter() - t0)
    world = dist.get_world_size()
    meta = torch.tensor([n_local, nf_local, 0 if g is None else 1], dtype=torch.int64, device=device)
    metas = _allgather(meta).cpu().numpy()
    ns, nf, has_g = metas[:, 0], metas[:, 1], bool(metas[:, 2].min())
    max_s, max_f = int(ns.max()), int(nf.max())
    # lengths
    tl = torch.zeros(max_s, dtype=torch.int32, device=device)
    tl[:n_local] = torch.from_numpy(lengths).to(device)
    all_l = _allgather(tl).cpu().numpy()
    lengths_all = np.concatenate([all_l[r, :ns[r]] for r in range(world)]).astype(np.int32)
    # frames: equal-sized (padded) shards, compacted on the device
    pad = torch.zeros((max_f, bins), dtype=tf.dtype, device=device)
    pad[:nf_local] = tf
    all_f = _allgather(pad)
    arena = torch.cat([all_f[r, :nf[r]] for r in range(world)], dim=0).contiguous()
    garena = None
    if has_g:
        pg = torch.zeros((max_s, bins), dtype=torch.float64, device=device)
        pg[:n_local] = tg
        all_g = _allgather(pg)
        garena = torch.cat([all_g[r, :ns[r]] for r in range(world)], dim=0).contiguous()
    offsets = np.zeros(lengths_all.size, np.int64)
    if lengths_all.size > 1:
        offsets[1:] = np.cumsum(lengths_all[:-1].astype(np.int64))
    if arena.is_cuda:
        torch.cuda.synchronize(device)
    return SharedArena(arena, garena, offsets, lengths_all, time.perf_counter() - t0)


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


def tiles_pairs(n_songs, T, tiles):
    """tile_pairs for a list of tiles, concatenated in list order; full off-diagonal tiles are enumerated in one
    vectorised step (a per-step cost of the distributed pass: the reference's itertools.combinations)."""
    if not len(tiles):
        return np.zeros((0, 2), np.int32)
    gi, gj = np.divmod(np.arange(T * T, dtype=np.int32), T)   # row-major (i, j) inside a full tile
    out = [None] * len(tiles)
    full = [k for k, (a, b) in enumerate(tiles) if a != b and (a + 1) * T <= n_songs and (b + 1) * T <= n_songs]
    if full:
        ab = np.array([tiles[k] for k in full], dtype=np.int32)
        I = (ab[:, 0:1] * T + gi[None, :])
        J = (ab[:, 1:2] * T + gj[None, :])
        blk = np.stack([I, J], axis=2)                          # (n_full, T*T, 2)
        for q, k in enumerate(full):
            out[k] = blk[q]
    for k in range(len(tiles)):
        if out[k] is None:
            out[k] = tile_pairs(n_songs, T, tiles[k])
    return np.concatenate(out, axis=0).astype(np.int32, copy=False)


def tile_costs(lengths, T, tiles, m=9):
    """Estimated work of each tile: sum over its pairs of M*N (cells of the cross-recurrence plot)."""
    w = np.maximum(np.asarray(lengths, dtype=np.float64) - m + 1, 0.0)
    n = len(w)
    nt = (n + T - 1) // T
    pad = np.zeros(nt * T)
    pad[:n] = w
    s = pad.reshape(nt, T).sum(axis=1)
    s2 = (pad.reshape(nt, T) ** 2).sum(axis=1)
    if not len(tiles):
        return np.zeros(0)
    ab = np.asarray(tiles, dtype=np.int64).reshape(-1, 2)
    a, b = ab[:, 0], ab[:, 1]
    return np.where(a != b, s[a] * s[b], 0.5 * (s[a] ** 2 - s2[a]))


def assign_tiles(costs, world):
    """Longest-processing-time greedy: tiles by decreasing cost to the least-loaded rank (ties: lowest rank).
    Returns (list of tile-index arrays, one per rank, each in ascending tile order; per-rank load).  Deterministic
    on every rank."""
    import heapq
    order = np.argsort(-costs, kind="stable")
    heap = [(0.0, r) for r in range(world)]
    owner = np.zeros(len(costs), dtype=np.int64)
    cl = costs.tolist()
    for t in order.tolist():
        load_r, r = heapq.heappop(heap)
        owner[t] = r
        heapq.heappush(heap, (load_r + cl[t], r))
    load = np.zeros(world)
    for load_r, r in heap:
        load[r] = load_r
    return [np.nonzero(owner == r)[0] for r in range(world)], load


def shard_tiles(n_songs, lengths, world, T=128, m=9, tiles=None):
    """(tiles, per-rank tile-index arrays, per-rank estimated load) of an all-pairs run."""
    tiles = tile_list(n_songs, T) if tiles is None else tiles
    shares, load = assign_tiles(tile_costs(lengths, T, tiles, m), world)
    return tiles, shares, load


def shard_pairs(n_songs, lengths, world, rank, T=128, m=9, tiles=None):
    """This rank's share of the all-pairs list: concatenated pairs of its tiles."""
    tiles, shares, _ = shard_tiles(n_songs, lengths, world, T, m, tiles)
    return tiles_pairs(n_songs, T, [tiles[t] for t in shares[rank]])


def tile_pair_count(n_songs, T, tile):
    a, b = tile
    ni = min((a + 1) * T, n_songs) - a * T
    nj = min((b + 1) * T, n_songs) - b * T
    return ni * nj if a != b else ni * (ni - 1) // 2


# ---------------------------------------------------------------------------- scores: float32 tiles to rank 0
def gather_score_tiles(local, dst=0, device=None, counts=None, timing=None):
    """local: float32 array [n_types, n_local_pairs] of this rank (pairs in the order of its tile list).
    Returns on `dst` a list of per-rank [n_types, n_pairs_of_rank] arrays, None elsewhere.  Single process: [local].
    counts: every rank's pair count when the caller knows them (the tile deal is deterministic), which saves a
    collective; timing: dict that receives the seconds spent waiting for / receiving the other ranks' buffers."""
    import torch
    import torch.distributed as dist
    local = np.ascontiguousarray(local, dtype=np.float32)
    if not _is_dist():
        return [local]
    world, rank = dist.get_world_size(), dist.get_rank()
    if device is None:
        device = _default_device()
    K = local.shape[0]
    if counts is None:
        n = torch.tensor([local.shape[1]], dtype=torch.int64, device=device)
        counts = _allgather(n).view(-1).cpu().numpy()
    counts = np.asarray(counts, dtype=np.int64)
    assert int(counts[rank]) == local.shape[1], "rank %d: %d scores for %d pairs" % (rank, local.shape[1], counts[rank])
    nmax = int(counts.max())
    t0 = time.perf_counter()
    buf = torch.zeros((K, max(nmax, 1)), dtype=torch.float32, device=device)
    buf[:, :local.shape[1]] = torch.from_numpy(local).to(device)
    outs = [torch.empty_like(buf) for _ in range(world)] if rank == dst else None
    dist.gather(buf, gather_list=outs, dst=dst)
    if rank != dst:
        return None
    if torch.device(device).type == 'cuda':
        torch.cuda.synchronize(device)
    t1 = time.perf_counter()
    host = torch.stack(outs).cpu().numpy()          # one device-to-host copy for all ranks
    if timing is not None:
        timing["gather_wait_s"] = t1 - t0
        timing["gather_d2h_s"] = time.perf_counter() - t1
    return [host[r][:, :int(c)] for r, c in enumerate(counts)]


def place_tiles(Ds, keys, n_songs, T, tiles, shares, per_rank_scores):
    """Drop every rank's score tiles into the N x N matrices Ds[key] (upper triangle; slicing, no fancy indexing)."""
    for r, sc in enumerate(per_rank_scores):
        off = 0
        for t in shares[r]:
            a, b = tiles[t]
            i0, i1 = a * T, min((a + 1) * T, n_songs)
            j0, j1 = b * T, min((b + 1) * T, n_songs)
            if a != b:
                cnt = (i1 - i0) * (j1 - j0)
                for k, key in enumerate(keys):
                    Ds[key][i0:i1, j0:j1] = sc[k, off:off + cnt].reshape(i1 - i0, j1 - j0)
            else:
                n = i1 - i0
                cnt = n * (n - 1) // 2
                iu = np.triu_indices(n, 1)   # row-major (i < j): the order of tile_pairs
                for k, key in enumerate(keys):
                    blk = np.zeros((n, n), np.float32)
                    blk[iu] = sc[k, off:off + cnt]
                    Ds[key][i0:i1, i0:i1] = blk
            off += cnt
        assert off == sc.shape[1], "rank %d: %d scores for %d tile pairs" % (r, sc.shape[1], off)


def score_tiles(alg, tile_ids, T, tiles=None, lengths=None, place=True):
    """One distributed pass over a subset of the tile list (songs must already be resident: share_songs_allgather /
    preload): the tiles are dealt to the ranks by estimated cost, every rank scores the pairs of its tiles, the float32
    score tiles are gathered on rank 0 and (place=True) dropped into alg.Ds there.  Returns a dict of figures."""
    rank, world, _ = dist_env()
    lengths = alg.arena_lengths() if lengths is None else lengths
    tiles = tile_list(alg.N, T) if tiles is None else tiles
    sub = [tiles[int(t)] for t in tile_ids]
    t0 = time.perf_counter()
    shares, load = assign_tiles(tile_costs(lengths, T, sub, alg.m), world)
    mine = tiles_pairs(alg.N, T, [sub[t] for t in shares[rank]])
    t1 = time.perf_counter()
    alg._defer_store = True   # rank 0 places whole tiles below; no per-pair scatter inside similarity()
    try:
        sims = alg.similarity(mine) if len(mine) else {k: np.zeros(0) for k in alg.similarity_types}
    finally:
        alg._defer_store = False
    t_score = time.perf_counter() - t1
    keys = sorted(sims.keys())
    t2 = time.perf_counter()
    local = np.stack([sims[k].astype(np.float32) for k in keys]) if keys else np.zeros((0, len(mine)), np.float32)
    timing = {}
    counts = [sum(tile_pair_count(alg.N, T, sub[t]) for t in shares[r]) for r in range(world)]
    gathered = gather_score_tiles(local, counts=counts, timing=timing)
    t3 = time.perf_counter()
    if rank == 0 and place and len(alg.Ds):
        have = [i for i, k in enumerate(keys) if k in alg.Ds]
        place_tiles(alg.Ds, [keys[i] for i in have], alg.N, T, sub, shares, [g[have] for g in gathered])
    t4 = time.perf_counter()
    return {"pairs_local": int(len(mine)), "pairs_total": int(sum(tile_pair_count(alg.N, T, t) for t in sub)),
            "deal_s": t1 - t0, "score_s": t_score, "gather_place_s": t4 - t2, "gather_s": t3 - t2, "place_s": t4 - t3,
            "gather_wait_s": timing.get("gather_wait_s", 0.0), "gather_d2h_s": timing.get("gather_d2h_s", 0.0),
            "lpt_imbalance": float(load.max() / max(load.mean(), 1e-30)),
            "tiles": len(sub), "keys": keys, "pairs_mine": mine}


def all_pairwise_distributed(alg, T=128, symmetric=True, save=True):
    """The multi-GPU form of CoverAlgorithm.all_pairwise (CoverAlgorithm.py:138-184): every rank loads its share of
    the songs, the arena is all-gathered device-to-device, every rank scores the pairs of its tiles, rank 0 receives
    float32 score tiles, assembles Ds, symmetrises (`Ds += Ds.T`) and saves '<prefix>_Ds.h5'.  Construct `alg` with
    do_memmaps=True on rank 0 only.  Returns a dict of timings / balance figures (every rank)."""
    import torch.distributed as dist
    from . import hdf5io
    if not symmetric:
        raise ValueError("all_pairwise_distributed scores the upper triangle and mirrors it (symmetric=True); the "
                         "reference's ordered-pair mode is single-process only")
    rank, world, _ = dist_env()
    if rank != 0 and alg.do_memmaps:
        raise ValueError("construct the plugin with do_memmaps=True on rank 0 only: every rank would truncate the same "
                         "'<prefix>_<type>_dmat' files")
    t0 = time.perf_counter()
    s0, s1 = song_shares(alg.N, world)[rank]
    kinds = alg.share_songs_allgather(range(s0, s1))        # features of MY songs -> all-gather -> device arenas
    t_share = time.perf_counter() - t0
    tiles = tile_list(alg.N, T)
    info = score_tiles(alg, range(len(tiles)), T, tiles)
    if rank == 0:
        alg.get_all_clique_ids()
        for key in alg.Ds:
            alg.Ds[key] += alg.Ds[key].T
        if save:
            hdf5io.save("%s_Ds.h5" % alg.get_cacheprefix(), {k: np.asarray(v) for k, v in alg.Ds.items()})
    if _is_dist():
        dist.barrier()
    info.pop("pairs_mine")
    info.update({"kinds": kinds, "share_songs_s": t_share, "total_s": time.perf_counter() - t0})
    return info
