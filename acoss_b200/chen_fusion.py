"""Drop-in mirror of the reference's ChenFusion plugin (benchmarking/ChenFusion.py:17-85) on the same CUDA path.

What the reference's similarity() computes per pair (ChenFusion.py:57-67): get_csm_blocked_oti (CRPUtils.py:138-166:
global OTI, every chroma block of song i rolled by it) -> Euclidean cross-similarity of the "stacked" features ->
csm_to_binary (row-wise kNN only, CRPUtils.py:169-199) -> qmax on a zeroed table, `D *= 0`, dmax on a zeroed table,
stored UN-normalised in Ds["qmax"] / Ds["dmax"].  The stacked features come from
`librosa.feature.stack_memory(chroma, self.tau, self.m)` (ChenFusion.py:52): positionally that is n_steps = tau,
delay = m, and with the class default tau = 1 a one-step "stack" is the chroma itself -- no delay embedding.  This
class therefore runs the Serra09 kernels with window 1, ACOSS_FLAG_NON_MUTUAL | ACOSS_FLAG_DMAX_CLEAN and no length
normalisation; tau > 1 (a genuine stack of tau frames m apart) is not implemented and raises.  The reference calls
`get_csm_euclidean` and `doSimilarityFusion`, neither of which exists in its CRPUtils.py / SimilarityFusion.py
(the script cannot run as published); Euclidean is `get_csm` (CRPUtils.py:67-84), and do_late_fusion raises here.
"""
from __future__ import annotations

import numpy as np

from . import _native
from .serra09 import Serra09


class ChenFusion(Serra09):
    def __init__(self, datapath="../features_covers80", chroma_type='hpcp', shortname='benchmark', oti=True, kappa=0.095,
                 tau=1, m=9, downsample_fac=40, do_memmaps=True, cachedir="cache", songs=None, labels=None, device=0):
        if tau != 1:
            raise NotImplementedError("ChenFusion with tau = %d: stack_memory(chroma, n_steps=tau, delay=m) stacks tau frames "
                                      "m apart; only the reference's default tau = 1 (no stacking) is implemented" % tau)
        self.tau = tau
        self.delay = m       # the reference's m ends up as stack_memory's `delay`, unused when tau = 1
        Serra09.__init__(self, datapath=datapath, chroma_type=chroma_type, shortname=shortname, oti=oti, kappa=kappa, m=1,
                         downsample_fac=downsample_fac, do_memmaps=False, cachedir=cachedir, songs=songs, labels=labels,
                         device=device, do_mfcc=False, name="ChenFusion")
        # similarity types of ChenFusion.py:41; the memmaps are (re)created under these names
        self.similarity_types = ["qmax", "dmax"]
        self.do_memmaps = do_memmaps
        self.Ds = {}
        if do_memmaps:
            import os
            if not os.path.exists(cachedir):
                os.makedirs(cachedir)
            for s in self.similarity_types:
                self.Ds[s] = np.memmap('%s_%s_dmat' % (self.get_cacheprefix(), s), shape=(self.N, self.N), mode='w+',
                                       dtype='float32')

    def similarity(self, idxs):
        """ChenFusion.py:57-67 for all pairs of idxs at once; returns {'qmax', 'dmax'} raw maxima (float64) and stores
        them in Ds as the reference does."""
        idxs = np.asarray(idxs).reshape(-1, 2)
        if idxs.shape[0] == 0:
            return {s: np.zeros(0) for s in self.similarity_types}
        ctx, slot = self._context_for('chroma', idxs.reshape(-1))
        flags = _native.FLAG_NON_MUTUAL | _native.FLAG_DMAX_CLEAN | (0 if self.oti else _native.FLAG_NO_OTI)
        r = ctx.serra09_pairs(slot[idxs].astype(np.int32), m=1, kappa=self.kappa, flags=flags)
        sims = {"qmax": r['qmax_raw'].astype(np.float64), "dmax": r['dmax_raw'].astype(np.float64)}
        if len(self.Ds) > 0 and not getattr(self, '_defer_store', False):
            for key in sims:
                self.Ds[key][idxs[:, 0], idxs[:, 1]] = sims[key]
        return sims

    def normalize_by_length(self):
        """ChenFusion.py:69-79: Ds[key][i, j] = sqrt(len(song j)) / Ds[key][i, j] (non-symmetric; zeros become inf, as
        in the reference)."""
        L = np.array([self.load_features(j)['chroma'].shape[1] for j in range(self.N)], dtype=np.float64)
        with np.errstate(divide='ignore'):
            for key in self.Ds:
                self.Ds[key][:, :] = (np.sqrt(L)[None, :] / np.asarray(self.Ds[key], dtype=np.float64)).astype(np.float32)

    def do_late_fusion(self):
        raise NotImplementedError("ChenFusion.do_late_fusion calls doSimilarityFusion, which SimilarityFusion.py does not define "
                                  "(benchmarking/ChenFusion.py:82); similarity-network fusion of score matrices is not part of "
                                  "this path")
