"""Drop-in mirror of the reference's plugin base class (benchmarking/CoverAlgorithm.py:12-418):
same constructor arguments, method names, pair enumeration, on-disk artefacts and evaluation
statistics.  Only the host-side driver lives here; subclasses score pairs on the GPU.

Differences that are deliberate:
  * `all_pairwise` hands the whole pair list to `similarity` in a few large calls instead of
    45 joblib chunks (CoverAlgorithm.py:169-173); `parallel` / `n_cores` are accepted and ignored.
  * per-song feature files may be deepdish-style `.h5` (read by acoss_b200.hdf5io, no HDF5
    library needed) or `.npz`; an in-memory `songs` list can replace `datapath` (synthetic runs).
  * `cleanup_memmap` actually removes the memmap files (the reference calls rmtree on a file
    and always prints 'Could not clean-up automatically.', CoverAlgorithm.py:319-328).
"""
from __future__ import annotations

import glob
import os
import time
import warnings

import numpy as np

from . import _native, hdf5io


class CoverAlgorithm(object):
    """
    Attributes
    ----------
    filepaths: list(string)
        List of paths to all files in the dataset
    cliques: {string: set}
        A dictionary of all cover cliques, where the cliques index into filepaths
    Ds: {string similarity type: ndarray(num files, num files)}
        A dictionary of pairwise similarity matrices, whose indices index into filepaths.
    """

    def __init__(self, name="Generic", datapath="features_benchmark", shortname="full", cachedir="cache",
                 cache2dir="cache2", similarity_types=["main"], do_memmaps=True, songs=None, labels=None):
        self.name = name
        self.shortname = shortname
        self.cachedir = cachedir
        self._songs = songs
        self._labels = labels
        if songs is not None:
            self.filepaths = ["mem://song_%06d" % i for i in range(len(songs))]
        else:
            self.filepaths = sorted(glob.glob("%s/*.h5" % datapath))
            if len(self.filepaths) == 0:
                self.filepaths = sorted(glob.glob("%s/*.npz" % datapath))
        self.cliques = {}
        self.all_feats = {}  # For caching loaded features
        self.N = len(self.filepaths)
        self.do_memmaps = do_memmaps
        self.similarity_types = similarity_types
        self.cache2dir = None
        self.Ds = {}
        if do_memmaps:
            if not os.path.exists(cachedir):
                os.makedirs(cachedir)
            for s in similarity_types:
                self.Ds[s] = np.memmap('%s_%s_dmat' % (self.get_cacheprefix(), s), shape=(self.N, self.N),
                                       mode='w+', dtype='float32')
        print("Initialized %s algorithm on %i songs in dataset %s" % (name, self.N, shortname))

    def set_cache2dir(self, cache2dir):
        self.cache2dir = cache2dir
        if not os.path.exists(cache2dir):
            os.mkdir(cache2dir)

    def get_cacheprefix(self):
        """Descriptive file prefix for cached features and distance matrices (CoverAlgorithm.py:59-64)."""
        return "%s/%s_%s" % (self.cachedir, self.name, self.shortname)

    # ------------------------------------------------------------------ features
    def _read_feature_file(self, i):
        if self._songs is not None:
            feats = dict(self._songs[i])
            if 'label' not in feats:
                feats['label'] = self._labels[i] if self._labels is not None else "song_%i" % i
            return feats
        path = self.filepaths[i]
        if path.endswith(".npz"):
            z = np.load(path, allow_pickle=False)
            feats = {k: z[k] for k in z.files}
            for k in ('label', 'track_id'):
                if k in feats:
                    feats[k] = str(feats[k])
            return feats
        return hdf5io.load(path)

    def load_features(self, i):
        """Load the fields of one song's feature file and record its cover clique as a side
        effect (CoverAlgorithm.py:66-90)."""
        feats = self._read_feature_file(i)
        if not feats['label'] in self.cliques:
            self.cliques[feats['label']] = set([])
        self.cliques[feats['label']].add(i)
        return feats

    def get_all_clique_ids(self, verbose=False):
        """Load all feature files to get clique information as a side effect; cached as
        '<prefix>_clique_info.txt' with lines "%i,%s" (CoverAlgorithm.py:92-114)."""
        filepath = "%s_clique_info.txt" % self.get_cacheprefix()
        if not os.path.exists(filepath):
            d = os.path.dirname(filepath)
            if d and not os.path.exists(d):
                os.makedirs(d)
            fout = open(filepath, "w")
            for i in range(len(self.filepaths)):
                feats = CoverAlgorithm.load_features(self, i)
                if verbose:
                    print(i)
                fout.write("%i,%s\n" % (i, feats['label']))
            fout.close()
        else:
            fin = open(filepath)
            for line in fin.readlines():
                i, label = line.split(",", 1)
                label = label.strip()
                if label not in self.cliques:
                    self.cliques[label] = set([])
                self.cliques[label].add(int(i))
            fin.close()

    def similarity(self, idxs):
        """Given (P, 2) song index pairs, score each and store the score in Ds[type][i, j] as a
        side effect (CoverAlgorithm.py:117-136).  Subclasses override."""
        (a, b) = idxs.shape
        for k in range(a):
            i = idxs[k][0]
            j = idxs[k][1]
            self.Ds["main"][i, j] = 0.0

    # ------------------------------------------------------------------ all pairs
    def all_pairs_list(self, symmetric=True):
        """Pair order of CoverAlgorithm.py:166-168: itertools.combinations / permutations of
        range(N), as an (P, 2) int32 array (built without the Python-level list)."""
        N = len(self.filepaths)
        if symmetric:
            I, J = np.triu_indices(N, 1)  # row-major upper triangle == combinations order
        else:
            I, J = np.nonzero(~np.eye(N, dtype=bool))  # row-major, i != j == permutations order
        return np.stack([I, J], axis=1).astype(np.int32)

    def all_pairwise(self, parallel=0, n_cores=12, symmetric=False, precomputed=False, chunk_pairs=1 << 22):
        """All pairwise comparisons (CoverAlgorithm.py:138-184): enumerate pairs, score, symmetrise
        (`Ds += Ds.T`), save '<prefix>_Ds.h5'."""
        tic = time.time()
        h5filename = "%s_Ds.h5" % self.get_cacheprefix()
        self._Ds_dev_valid = False
        if precomputed:
            self.Ds = hdf5io.load(h5filename)
            self.get_all_clique_ids()
        else:
            all_pairs = self.all_pairs_list(symmetric)
            # the score matrices are also kept on the device where the plugin can (Serra09.device_scores_begin), so
            # that getEvalStatistics ranks them there without another trip over PCIe
            on_device = bool(getattr(self, 'device_scores_begin', lambda: False)())
            for p0 in range(0, all_pairs.shape[0], chunk_pairs):
                self.similarity(all_pairs[p0:p0 + chunk_pairs])
            self.get_all_clique_ids()
            if symmetric:
                for similarity_type in self.Ds:
                    self.Ds[similarity_type] += self.Ds[similarity_type].T
            if on_device:
                self.device_scores_end(symmetric)
            d = os.path.dirname(h5filename)
            if d and not os.path.exists(d):
                os.makedirs(d)
            hdf5io.save(h5filename, {k: np.asarray(v) for k, v in self.Ds.items()})
            # MATLAB twin of the same dict ('<prefix>_Ds.mat', one float32 N x N variable per similarity
            # type): the reference's consumers under coverstats/ read .mat files, BASELINE names the layout
            try:
                from scipy.io import savemat
                savemat("%s_Ds.mat" % self.get_cacheprefix(), {k: np.asarray(v) for k, v in self.Ds.items()})
            except ImportError:  # scipy is optional on the host side
                pass
        print("Elapsed Time All Pairwise: %.3g" % (time.time() - tic))

    # ------------------------------------------------------------------ batch mode
    def do_batch_features(self, n_batches, idx):
        """CoverAlgorithm.py:186-201."""
        N = len(self.filepaths)
        w = int(np.ceil(N / n_batches))
        for i in np.arange(w) + idx * w:
            if i < N:
                self.load_features(i)

    def batch_block_idxs(self, w, idx, wsub, isub, jsub):
        """The (row, col) pairs of one sub-block, exactly as CoverAlgorithm.py:227-244 builds them:
        block idx of the lower-triangular (I >= J) block grid, sub-block (isub, jsub), row >= col."""
        N = len(self.filepaths)
        res = int(N / w)
        I, J = np.meshgrid(np.arange(res), np.arange(res))
        I, J = I.flatten(), J.flatten()
        I, J = I[I >= J], J[I >= J]
        i = I[idx]
        j = J[idx]
        pixi = np.arange(w)[isub * wsub:(isub + 1) * wsub]
        pixj = np.arange(w)[jsub * wsub:(jsub + 1) * wsub]
        I, J = np.meshgrid(pixi, pixj)
        idxs = np.array([I.flatten() + i * w, J.flatten() + j * w]).T
        idxs = idxs[idxs[:, 0] < N, :]
        idxs = idxs[idxs[:, 1] < N, :]
        idxs = idxs[idxs[:, 0] >= idxs[:, 1], :]
        return idxs

    def do_batch_subbatch(self, w, idx, wsub, isub, jsub):
        """Compute a sub-block of similarities (CoverAlgorithm.py:203-247)."""
        fout = "{}_{}_{}_{}_{}.h5".format(self.get_cacheprefix(), idx, wsub, isub, jsub)
        if os.path.exists(fout):
            similarities = hdf5io.load(fout)
            os.remove(fout)
        else:
            idxs = self.batch_block_idxs(w, idx, wsub, isub, jsub)
            similarities = self.similarity(idxs)
            similarities['idxs'] = idxs
        return similarities

    def do_batch(self, w, idx, wsub=-1):
        """Compute and save a block of similarities, resumable after every sub-block
        (CoverAlgorithm.py:249-295)."""
        similarities = {}
        blocks_completed = {}
        fout = "{}_{}.h5".format(self.get_cacheprefix(), idx)
        if os.path.exists(fout):
            try:
                res = hdf5io.load(fout)
                similarities = res['similarities']
                blocks_completed = {tuple(int(v) for v in b): True for b in res['blocks_completed']}
            except Exception:
                print("Error loading", fout, ": recomputing")
                similarities, blocks_completed = {}, {}
        if wsub == -1:
            wsub = w
        k = int(w / wsub)
        col_range = list(range(k))
        for i in range(k):
            for j in col_range:
                if not (i, j) in blocks_completed:
                    tic = time.time()
                    self.all_feats = {}
                    s = self.do_batch_subbatch(w, idx, wsub, i, j)
                    if len(similarities) == 0:
                        similarities = s
                    else:
                        for key in s:
                            similarities[key] = np.concatenate((similarities[key], s[key]))
                    blocks_completed[(i, j)] = True
                    d = os.path.dirname(fout)
                    if d and not os.path.exists(d):
                        os.makedirs(d)
                    # deepdish pickles the tuple-keyed dict; we store the keys as an (n, 2) array
                    hdf5io.save(fout, {'similarities': similarities,
                                       'blocks_completed': np.array(sorted(blocks_completed.keys()), dtype=np.int64).reshape(-1, 2)})
                    print("Elapsed Time Sub-Batch %i_%i_%i: %.3g" % (idx, i, j, time.time() - tic), flush=True)
            col_range = list(reversed(col_range))  # zigzag order (CoverAlgorithm.py:294-295)

    def load_batches(self, fileprefix):
        """Aggregate batch files into Ds (CoverAlgorithm.py:297-317; the diagonal is counted twice
        there too)."""
        files = glob.glob("{}*.h5".format(fileprefix))
        self._Ds_dev_valid = False
        for key in self.Ds.keys():
            self.Ds[key] = np.zeros_like(self.Ds[key])
        for f in files:
            loaded = hdf5io.load(f)
            if 'similarities' not in loaded:
                continue
            res = loaded['similarities']
            idxs = res['idxs']
            I = idxs[:, 0]
            J = idxs[:, 1]
            for key in self.Ds.keys():
                if key not in res:
                    continue
                self.Ds[key][I, J] += res[key]
                self.Ds[key][J, I] += res[key]
        self.get_all_clique_ids()

    def cleanup_memmap(self):
        """Remove the memmap files of the song-level similarity matrices."""
        for s in list(self.Ds):
            path = '%s_%s_dmat' % (self.get_cacheprefix(), s)
            try:
                if isinstance(self.Ds[s], np.memmap):
                    self.Ds[s]._mmap.close()
                if os.path.exists(path):
                    os.remove(path)
            except Exception:  # noqa
                print('Could not clean-up automatically.')

    # ------------------------------------------------------------------ evaluation
    def eval_ranks(self, similarity_type):
        """Step 1 + the sort and rank loop of getEvalStatistics (CoverAlgorithm.py:345-382) on the
        device (acoss_eval_ranks, csrc/kernels_eval.cuh): returns (ranks, AllMap) as the reference
        builds them, NaN where the reference leaves NaN."""
        N = self.Ds[similarity_type].shape[0]
        if getattr(self, '_eval_ctx', None) is None:
            self._eval_ctx = _native.Context(getattr(self, 'device', 0))
        cliques = [list(self.cliques[s]) for s in self.cliques]
        dev = (getattr(self, '_Ds_dev', None) or {}).get(similarity_type) if getattr(self, '_Ds_dev_valid', False) else None
        if dev is not None:
            # the matrix all_pairwise left on the device (ACOSS_FLAG_DEVICE_INPUT): no 4 N^2-byte upload
            iranks, Ks, rows = self._eval_ctx.eval_ranks(None, cliques, device_ptr=dev.data_ptr())
        else:
            D = np.ascontiguousarray(self.Ds[similarity_type], dtype=np.float32)
            iranks, Ks, rows = self._eval_ctx.eval_ranks(D, cliques)
        ranks = np.nan * np.ones(N)
        AllMap = np.nan * np.ones(N)
        for i in range(rows):
            ir = iranks[i]
            ranks[i] = ir[0]
            # P = [j / r for j, r in zip(range(1, K), iranks)]; np.mean(P)   (CoverAlgorithm.py:387-389)
            AllMap[i] = np.mean(np.arange(1, ir.size + 1, dtype=np.float64) / ir.astype(np.float64))
        return ranks, AllMap

    def getEvalStatistics(self, similarity_type, topsidx=[1, 10, 100, 1000]):
        """MR, MRR, MDR, MAP and Top-K for one similarity measure, plus a CSV row
        (CoverAlgorithm.py:330-418).  The argsort of the N x N matrix and the O(N^2) Python rank
        loop (:362-382) run on the device; the float64 statistics follow here from the integer
        ranks exactly as the reference computes them."""
        N = self.Ds[similarity_type].shape[0]
        ranks, AllMap = self.eval_ranks(similarity_type)
        MAP = np.nanmean(AllMap)
        ranks = ranks[np.isnan(ranks) == 0]
        MR = np.mean(ranks)
        MRR = 1.0 / N * (np.sum(1.0 / ranks))
        MDR = np.median(ranks)
        print("%s %s STATS\n-------------------------\nMR = %.3g\nMRR = %.3g\nMDR = %.3g\nMAP = %.3g" % (
            self.name, similarity_type, MR, MRR, MDR, MAP))
        tops = np.zeros(len(topsidx))
        for i in range(len(tops)):
            tops[i] = np.sum(ranks <= topsidx[i])
            print("Top-%i: %i" % (topsidx[i], tops[i]))
        resultsfile = "results_%s.csv" % self.shortname
        if not os.path.exists(resultsfile):
            fout = open(resultsfile, "w")
            fout.write("name, MR, MRR, MDR, MAP")
            for t in topsidx:
                fout.write(",Top-%i" % t)
            fout.write("\n")
            fout.close()
        fout = open(resultsfile, "a")
        fout.write("%s_%s," % (self.name, similarity_type))
        fout.write("%.3g, %.3g, %.3g, %.3g" % (MR, MRR, MDR, MAP))
        for t in tops:
            fout.write(", %.3g" % t)
        fout.write("\n")
        fout.close()
        return (MR, MRR, MDR, MAP, tops)
