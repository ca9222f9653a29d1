"""Drop-in mirror of the reference's Serra09 plugin (benchmarking/Serra09.py:73-196) whose
per-pair arithmetic runs on the GPU through the C ABI (acoss_b200._native).

Same constructor keywords, `load_features(i)` returns the same per-song dict
({'gchroma', 'chroma', 'mfcc', 'ssms'}), `similarity(idxs)` returns the same dict of float64
score arrays and writes `self.Ds[key][i][j]`.  The reference's per-pair Python loop
(Serra09.py:161) disappears: the whole (P, 2) index array goes to the device in one call.

Out of scope: the 'ssms_scatter_*' similarity types need kymatio's
Scattering2D; they are not produced (the keys are omitted, documented in DESIGN.md).
"""
from __future__ import annotations

import numpy as np

from . import _native
from .cover_algorithm import CoverAlgorithm


def global_chroma(chroma):
    """Global chroma of a (frames, bins) array (benchmarking/Serra09.py:24-28)."""
    if chroma.shape[1] not in [12, 24, 36]:
        raise IOError("Wrong axis for the input chroma array. Expected shape '(frame_size, bin_size)'")
    global_chroma = np.sum(chroma, axis=0)
    return global_chroma / np.max(global_chroma)


def sync_aggregate(data, fac, aggregate):
    """librosa.util.sync(data, np.arange(0, n, fac), aggregate=...) for a (d, n) array
    (Serra09.py:104,110): frames [fac*k, min(fac*(k+1), n)) are aggregated along the time axis."""
    d, n = data.shape
    starts = np.arange(0, n, fac)
    out = np.empty((d, len(starts)), dtype=data.dtype, order='F')
    for k, s in enumerate(starts):
        out[:, k] = aggregate(data[:, s:min(s + fac, n)], axis=1)
    return out


class Serra09(CoverAlgorithm):
    """
    Attributes
    ----------
    Same as CoverAlgorithm, plus
    chroma_type: string
        Type of chroma to use (key into features)
    downsample_fac: int
        The factor by which to downsample the HPCPs with median aggregation
    m: int
        Number of frames to stack in a delay embedding
    kappa: float
        Fraction of nearest neighbours kept per row / column of the cross-recurrence plot
    """

    def __init__(self, datapath="features_covers80", chroma_type='crema', shortname='Covers80', oti=True,
                 kappa=0.095, m=9, downsample_fac=40, do_memmaps=True, cachedir="cache", songs=None, labels=None,
                 device=0, do_mfcc=True, name="Serra09"):
        self.oti = oti
        self.m = m
        self.chroma_type = chroma_type
        self.kappa = kappa
        self.downsample_fac = downsample_fac
        self.device = device
        self.do_mfcc = do_mfcc
        self._ctx = {}        # 'chroma' / 'mfcc' -> (Context, song index -> arena slot)
        self._feat_ctx = None # context used by load_features_batch
        self._packed = {}     # kind -> (song ids, packed host arena): what load_features cached, ready to upload
        self.last_stage_ms = None
        similarity_types = ["chroma_qmax", "chroma_dmax"]
        if do_mfcc:
            similarity_types += ["mfcc_qmax", "mfcc_dmax"]
        CoverAlgorithm.__init__(self, name, datapath=datapath, shortname=shortname, cachedir=cachedir,
                                similarity_types=similarity_types, do_memmaps=do_memmaps, songs=songs, labels=labels)

    # ------------------------------------------------------------------ features
    def load_features(self, i):
        """Per-song features as Serra09.load_features computes them (Serra09.py:96-156, chroma and
        mfcc parts).  Already-prepared songs (a 'chroma' of shape (bins, L) plus 'gchroma', as the
        synthetic generator makes) pass through unchanged."""
        if i in self.all_feats:
            return self.all_feats[i]
        feats = CoverAlgorithm.load_features(self, i)
        if 'gchroma' in feats and 'chroma' in feats:
            ret = {'gchroma': feats['gchroma'], 'chroma': feats['chroma'], 'mfcc': feats.get('mfcc'), 'ssms': None}
        else:
            chroma = feats[self.chroma_type]                      # (n_frames, 12)
            gchroma = global_chroma(chroma)                        # Serra09.py:101-102
            chroma = sync_aggregate(chroma.T, self.downsample_fac, np.median)   # Serra09.py:104
            mfcc = None
            if self.do_mfcc and 'mfcc_htk' in feats:
                mfcc = np.array(feats['mfcc_htk'])                 # (13, n_frames)
                mfcc[np.isnan(mfcc)] = 0                           # Serra09.py:107-109
                mfcc[np.isinf(mfcc)] = 0
                mfcc = sync_aggregate(mfcc, self.downsample_fac, np.mean)       # Serra09.py:110
                N = min(chroma.shape[1], mfcc.shape[1])            # Serra09.py:111-114
                chroma = chroma[:, 0:N]
                mfcc = mfcc[:, 0:N]
            ret = {'gchroma': gchroma, 'chroma': chroma, 'mfcc': mfcc, 'ssms': None}
        self.all_feats[i] = ret
        return ret

    def load_features_batch(self, ids, batch=256):
        """load_features for many songs at once, the x40 aggregation and the global chroma done on
        the device (acoss_aggregate_songs).  Same results as load_features, bit
        for bit; songs that are already prepared or cached are skipped.  Feature files are read one
        batch at a time and only the fields this plugin uses are kept; a batch is split by dtype so that
        float32 songs are aggregated in float32 as numpy would; factors beyond the kernel's block size
        take the host path."""
        todo = [int(i) for i in np.unique(np.asarray(ids, dtype=np.int64)) if int(i) not in self.all_feats]
        if not todo:
            return
        if self.downsample_fac > 64:   # kFeatMaxFac of csrc/kernels_feat.cuh
            for i in todo:
                self.load_features(i)
            return
        for b0 in range(0, len(todo), batch):
            part = []
            for i in todo[b0:b0 + batch]:
                feats = CoverAlgorithm.load_features(self, i)
                if 'gchroma' in feats and 'chroma' in feats:
                    self.load_features(i)
                    continue
                chroma = np.asarray(feats[self.chroma_type])
                if chroma.shape[1] not in [12, 24, 36]:   # global_chroma's check, Serra09.py:25-26
                    raise IOError("Wrong axis for the input chroma array. Expected shape '(frame_size, bin_size)'")
                mfcc = np.asarray(feats['mfcc_htk']) if (self.do_mfcc and 'mfcc_htk' in feats) else None
                part.append((i, chroma, mfcc))
            if not part:
                continue
            if self._feat_ctx is None:
                self._feat_ctx = _native.Context(self.device)
            ctx = self._feat_ctx
            agg, g, mf = [None] * len(part), [None] * len(part), [None] * len(part)
            for dt in sorted({c.dtype for _, c, _ in part}, key=str):
                sel = [k for k, (_, c, _) in enumerate(part) if c.dtype == dt]
                a, gg = ctx.aggregate_songs([part[k][1] for k in sel], fac=self.downsample_fac, aggregate="median",
                                            want_gchroma=True)
                for q, k in enumerate(sel):
                    agg[k], g[k] = a[q], gg[q].copy()
            for dt in sorted({m.dtype for _, _, m in part if m is not None}, key=str):
                sel = [k for k, (_, _, m) in enumerate(part) if m is not None and m.dtype == dt]
                res = ctx.aggregate_songs([part[k][2] for k in sel], fac=self.downsample_fac, aggregate="mean",
                                          layout="bins_x_frames", zero_nonfinite=True)
                for q, k in enumerate(sel):
                    mf[k] = res[q]
            for k, (i, _, _) in enumerate(part):
                chroma, mfcc = np.asfortranarray(agg[k]), mf[k]
                if mfcc is not None:
                    N = min(chroma.shape[1], mfcc.shape[1])      # Serra09.py:111-114
                    chroma = chroma[:, 0:N]
                    mfcc = np.asfortranarray(mfcc)[:, 0:N]
                self.all_feats[i] = {'gchroma': g[k], 'chroma': chroma, 'mfcc': mfcc, 'ssms': None}

    # ------------------------------------------------------------------ device arena
    def _context_for(self, kind, song_ids):
        """Make sure every song in song_ids is in the device arena of `kind`; returns
        (Context, slot lookup array).  The arena is rebuilt (union of old and new songs) when a
        call needs songs that are not resident; all_pairwise uploads everything once."""
        entry = self._ctx.get(kind)
        need = np.unique(np.asarray(song_ids, dtype=np.int64))
        if entry is not None:
            ctx, slot = entry
            if np.all(slot[need] >= 0):
                return ctx, slot
            need = np.union1d(need, np.nonzero(slot >= 0)[0])
        else:
            ctx = _native.Context(self.device)
        cached = self._packed.get(kind)
        if cached is not None and np.all(np.isin(need, cached[0])):
            need, packed = cached   # the cached host arena already covers these songs
        else:
            self.load_features_batch(need)
            songs = []
            for i in need:
                f = self.load_features(int(i))
                if kind == 'chroma':
                    songs.append({'chroma': np.asarray(f['chroma']), 'gchroma': f['gchroma']})
                else:
                    songs.append({'chroma': np.asarray(f['mfcc']), 'gchroma': None})
            packed = _native.pack_songs(songs, pinned=True)   # host arena, page-locked when a GPU is present
            self._packed[kind] = (need, packed)
        ctx.upload(*packed)
        self._upload_gen = getattr(self, '_upload_gen', 0) + 1   # (EarlySNF re-attaches its second arena after an upload)
        slot = -np.ones(self.N, dtype=np.int64)
        slot[need] = np.arange(len(need))
        self._ctx[kind] = (ctx, slot)
        return ctx, slot

    # ------------------------------------------------------------------ multi-GPU arena (acoss_b200.parallel)
    def pack_shard(self, song_ids, kind):
        """Host arena (pinned when a GPU is present) of the given songs only: (frames, lengths, gchroma or None)."""
        song_ids = [int(i) for i in song_ids]
        self.load_features_batch(song_ids)
        songs = []
        for i in song_ids:
            f = self.load_features(i)
            if kind == 'chroma':
                songs.append({'chroma': np.asarray(f['chroma']), 'gchroma': f['gchroma']})
            else:
                songs.append({'chroma': np.asarray(f['mfcc']), 'gchroma': None})
        if not songs:
            bins = 12 if kind == 'chroma' else 13
            return np.zeros((0, bins), np.float32), np.zeros(0, np.int32), (np.zeros((0, bins)) if kind == 'chroma' else None)
        frames, _, lengths, g = _native.pack_songs(songs, pinned=True)
        return frames, lengths, g

    def share_songs_allgather(self, my_song_ids):
        """Every rank contributes the features of its own songs; one all_gather per feature kind leaves the whole
        arena in this rank's device memory, which the C ABI adopts device-to-device.  Returns the kinds shared."""
        from . import parallel
        my_song_ids = list(my_song_ids)
        kinds = ['chroma']
        if self.do_mfcc and (not my_song_ids or self.load_features(my_song_ids[0])['mfcc'] is not None):
            kinds.append('mfcc')
        if not hasattr(self, '_shards'):
            self._shards = {}   # kind -> (song ids, packed pinned host shard): packed once, re-uploaded every call
        for kind in kinds:
            cached = self._shards.get(kind)
            if cached is None or cached[0] != my_song_ids:
                cached = (my_song_ids, self.pack_shard(my_song_ids, kind))
                self._shards[kind] = cached
            arena = parallel.allgather_songs(cached[1])
            entry = self._ctx.get(kind)
            ctx = entry[0] if entry is not None else _native.Context(self.device)
            arena.upload_into(ctx)
            self._ctx[kind] = (ctx, np.arange(self.N, dtype=np.int64))
            self._arena_lengths = np.asarray(arena.lengths)
        self._shared_kinds = kinds
        return kinds

    def arena_lengths(self):
        """Frames per song of the resident arena (after share_songs_allgather / preload)."""
        if getattr(self, '_arena_lengths', None) is not None:
            return self._arena_lengths
        return np.array([self.load_features(i)['chroma'].shape[1] for i in range(self.N)])

    def preload(self, song_ids=None):
        """Upload the features of song_ids (default: all songs) ahead of scoring."""
        ids = np.arange(self.N) if song_ids is None else song_ids
        self._context_for('chroma', ids)
        if self.do_mfcc and self.load_features(int(ids[0]))['mfcc'] is not None:
            self._context_for('mfcc', ids)

    # ------------------------------------------------------------------ scoring
    def similarity(self, idxs):
        """Scores for the (P, 2) song pairs in idxs (Serra09.py:158-196).  Returns
        {'chroma_qmax', 'chroma_dmax'[, 'mfcc_qmax', 'mfcc_dmax']: float64 (P,)} and stores them
        in self.Ds[key][i][j] as a side effect."""
        idxs = np.asarray(idxs).reshape(-1, 2)
        N = idxs.shape[0]
        similarities = {}
        if N == 0:
            return {s: np.zeros(0) for s in self.similarity_types}
        kinds = [('chroma', 0)]
        shared = getattr(self, '_shared_kinds', None)
        if (('mfcc' in shared) if shared is not None else
                (self.do_mfcc and self.load_features(int(idxs[0, 0]))['mfcc'] is not None)):
            kinds.append(('mfcc', _native.FLAG_NO_OTI))   # Serra09.py:177-184: no OTI for mfcc
        stage = {}
        for kind, flags in kinds:
            ctx, slot = self._context_for(kind, idxs.reshape(-1))
            pairs = slot[idxs].astype(np.int32)
            if getattr(self, '_Ds_dev', None) is not None and not self._Ds_dev_valid:
                self._bind_device_scores(kind, ctx, slot)
            r = ctx.serra09_pairs(pairs, m=self.m, kappa=self.kappa, flags=flags)
            MN = r['shape'].sum(axis=1).astype(np.float64)     # M + N (Serra09.py:174-175)
            MN[MN <= 0] = 1.0
            similarities['%s_qmax' % kind] = r['qmax_raw'].astype(np.float64) / MN
            similarities['%s_dmax' % kind] = r['dmax_raw'].astype(np.float64) / MN
            stage[kind] = ctx.last_stage_ms()
        self.last_stage_ms = stage
        if (self.do_memmaps or len(self.Ds) > 0) and not getattr(self, '_defer_store', False):
            for key in similarities:
                if key in self.Ds:
                    self.Ds[key][idxs[:, 0], idxs[:, 1]] = similarities[key]
        return similarities

    # ------------------------------------------------------------------ device-resident score matrices
    def device_scores_begin(self):
        """Keep the N x N score matrices on the device as well (acoss_bind_score_matrices): every similarity() call
        scatters its normalised scores into them next to returning them, all_pairwise symmetrises them there, and
        getEvalStatistics ranks them without the N x N matrix crossing PCIe.  Returns False (and does nothing) when
        there is no CUDA torch, the subclass scores differently (ChenFusion, EarlySNF) or the matrices would not fit."""
        if type(self).similarity is not Serra09.similarity:
            return False
        try:
            import torch
        except ImportError:
            return False
        if not torch.cuda.is_available() or self.N * self.N * 4 * len(self.similarity_types) > (16 << 30):
            return False
        dev = torch.device('cuda', self.device)
        self._Ds_dev = {k: torch.zeros((self.N, self.N), dtype=torch.float32, device=dev) for k in self.similarity_types}
        self._Ds_dev_bound = {}
        self._Ds_dev_valid = False
        return True

    def device_scores_end(self, symmetric):
        """all_pairwise is through: `Ds += Ds.T` on the device copies (CoverAlgorithm.py:180-182), mark them usable."""
        import torch
        for key, t in self._Ds_dev.items():
            if symmetric:
                t.copy_(t + t.t())
        torch.cuda.synchronize(self.device)
        for ctx, _ in self._ctx.values():
            ctx.bind_score_matrices(0, 0, 0)
        self._Ds_dev_bound = {}
        self._Ds_dev_valid = True

    def _bind_device_scores(self, kind, ctx, slot):
        gen = (id(ctx), getattr(self, '_upload_gen', 0))
        if self._Ds_dev_bound.get(kind) == gen:
            return
        used = np.nonzero(slot >= 0)[0]
        song_of_slot = -np.ones(int(slot[used].max()) + 1 if used.size else 1, dtype=np.int32)
        song_of_slot[slot[used]] = used
        q, d = self._Ds_dev.get('%s_qmax' % kind), self._Ds_dev.get('%s_dmax' % kind)
        ctx.bind_score_matrices(q.data_ptr() if q is not None else 0, d.data_ptr() if d is not None else 0, self.N,
                                song_of_slot)
        self._Ds_dev_bound[kind] = gen

    def drop_device_arena(self):
        """Forget what is resident on the device (the host-side feature cache stays): the next
        similarity() call uploads again.  bench.py uses it to put the H2D copy inside a timed step."""
        for kind, (ctx, slot) in list(self._ctx.items()):
            self._ctx[kind] = (ctx, -np.ones(self.N, dtype=np.int64))

    def host_arena_bytes(self):
        n = sum(sum(a.nbytes for a in packed if a is not None) for _, packed in self._packed.values())
        return n + sum(sum(a.nbytes for a in shard if a is not None) for _, shard in getattr(self, '_shards', {}).values())

    def kernel_launches(self):
        return sum(ctx.kernel_launches() for ctx, _ in self._ctx.values())

    def close(self):
        for ctx, _ in self._ctx.values():
            ctx.close()
        self._ctx = {}
        self._Ds_dev = None
        self._Ds_dev_valid = False
        if self._feat_ctx is not None:
            self._feat_ctx.close()
            self._feat_ctx = None


def main(argv=None):
    """Command line of benchmarking/Serra09.py:198-244 (same flags)."""
    import argparse
    parser = argparse.ArgumentParser(description="Benchmarking with Joan Serra's Cover id algorithm (GPU)",
                                     formatter_class=argparse.ArgumentDefaultsHelpFormatter)
    parser.add_argument("-d", '--datapath', type=str, action="store", default='features_covers80',
                        help="Path to data files")
    parser.add_argument("-s", "--shortname", type=str, action="store", default="covers80", help="Short name for dataset")
    parser.add_argument("-c", '--chroma_type', type=str, action="store", default='hpcp',
                        help="Type of chroma to use for experiments")
    parser.add_argument("-p", '--parallel', type=int, choices=(0, 1), action="store", default=0,
                        help="Accepted for compatibility; the GPU path ignores it")
    parser.add_argument("-n", '--n_cores', type=int, action="store", default=1, help="Accepted for compatibility")
    parser.add_argument("-r", '--range', type=str, action="store", default="", help="Range of pairs to do, '<w>-<idx>'")
    parser.add_argument("-f", '--features', type=int, action="store", default=-1, help="Batch index in which to compute features")
    parser.add_argument("-w", '--wsub', type=int, action="store", default=-1, help="Sub-block size in batch mode")
    parser.add_argument("-b", '--batch_path', type=str, action="store", default="", help="Prefix of batch files to aggregate")
    cmd_args = parser.parse_args(argv)
    do_memmaps = not (len(cmd_args.range) > 0 or cmd_args.features > -1)
    serra09 = Serra09(cmd_args.datapath, cmd_args.chroma_type, cmd_args.shortname, do_memmaps=do_memmaps)
    serra09.set_cache2dir("cache2")
    if len(cmd_args.batch_path) > 0:
        serra09.load_batches(cmd_args.batch_path)
        for similarity_type in serra09.Ds.keys():
            serra09.getEvalStatistics(similarity_type)
    elif cmd_args.features > -1:
        serra09.do_batch_features(50, cmd_args.features)
    elif len(cmd_args.range) > 0:
        [w, idx] = [int(s) for s in cmd_args.range.split("-")]
        serra09.do_batch(w, idx, cmd_args.wsub)
    else:
        serra09.all_pairwise(cmd_args.parallel, cmd_args.n_cores, symmetric=True)
        for similarity_type in serra09.Ds.keys():
            serra09.getEvalStatistics(similarity_type)
        serra09.cleanup_memmap()


if __name__ == '__main__':
    main()
