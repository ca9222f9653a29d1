"""Drop-in mirror of the reference's EarlySNF plugin (benchmarking/EarlySNF.py:12-97) on the CUDA path.

EarlySNF.similarity computes, per pair, the Serra09 scores of every feature kind AND the scores of the fused plot:
the (M+N) x (M+N) kernel matrices of the kinds (get_WCSMSSM) go through similarity network fusion (snf_ws, K =
int(kappa (M+N)), 3 iterations, regularised diagonal), minus the fused cross block is binarised (mutual kNN) and scored
with qmax / dmax (EarlySNF.py:44-57, 85-90).  All of that runs on the device (acoss_snf_pairs, csrc/kernels_snf.cuh).

Deviation, stated: the reference fuses chroma with the scattering-transform SSMs ('ssms', kymatio's Scattering2D on
librosa / skimage-resized self-similarity images: out of scope).  Here the second fused kind is
the 13-bin mfcc sequence, which EarlySNF.py:66-71 embeds and scores anyway; the 'ssms_scatter_*' similarity types are
not produced.
"""
from __future__ import annotations

import numpy as np

from .serra09 import Serra09


class EarlySNF(Serra09):
    def __init__(self, datapath="features_covers80", chroma_type='hpcp', shortname='benchmark', oti=True, kappa=0.095, m=9,
                 downsample_fac=40, do_memmaps=True, cachedir="cache", songs=None, labels=None, device=0, niters=3):
        self.niters = niters
        Serra09.__init__(self, datapath=datapath, chroma_type=chroma_type, shortname=shortname, oti=oti, kappa=kappa, m=m,
                         downsample_fac=downsample_fac, do_memmaps=False, cachedir=cachedir, songs=songs, labels=labels,
                         device=device, do_mfcc=True, name="EarlySNF")
        self.similarity_types = ["chroma_qmax", "chroma_dmax", "mfcc_qmax", "mfcc_dmax", "snf_qmax", "snf_dmax"]
        self.do_memmaps = do_memmaps
        self.Ds = {}
        if do_memmaps:
            import os
            if not os.path.exists(cachedir):
                os.makedirs(cachedir)
            for s in self.similarity_types:
                self.Ds[s] = np.memmap('%s_%s_dmat' % (self.get_cacheprefix(), s), shape=(self.N, self.N), mode='w+',
                                       dtype='float32')
        self._aux_gen = -1

    def similarity(self, idxs):
        idxs = np.asarray(idxs).reshape(-1, 2)
        if idxs.shape[0] == 0:
            return {s: np.zeros(0) for s in self.similarity_types}
        if self.load_features(int(idxs[0, 0]))['mfcc'] is None:
            raise ValueError("EarlySNF fuses the chroma and mfcc kinds: the songs carry no 'mfcc' / 'mfcc_htk' features")
        defer = getattr(self, '_defer_store', False)
        self._defer_store = True
        try:
            sims = Serra09.similarity(self, idxs)          # chroma_* and mfcc_* (EarlySNF.py:58-71)
        finally:
            self._defer_store = defer
        ctx, slot = self._ctx['chroma']
        _, slot2 = self._ctx['mfcc']
        if not np.array_equal(slot, slot2):
            raise RuntimeError("chroma and mfcc arenas hold different song sets")
        if self._aux_gen != self._upload_gen:              # second arena of the same songs beside the chroma arena
            ctx.upload_aux(self._packed['mfcc'][1][0])
            self._aux_gen = self._upload_gen
        r = ctx.snf_pairs(slot[idxs].astype(np.int32), m=self.m, kappa=self.kappa, niters=self.niters)
        MN = r['shape'].sum(axis=1).astype(np.float64)
        MN[MN <= 0] = 1.0
        sims['snf_qmax'] = r['qmax_raw'].astype(np.float64) / MN   # EarlySNF.py:89-90
        sims['snf_dmax'] = r['dmax_raw'].astype(np.float64) / MN
        if len(self.Ds) > 0 and not defer:
            for key in sims:
                if key in self.Ds:
                    self.Ds[key][idxs[:, 0], idxs[:, 1]] = sims[key]
        return sims
