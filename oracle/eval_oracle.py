"""oracle/eval_oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatement of CoverAlgorithm.getEvalStatistics (benchmarking/CoverAlgorithm.py:330-418),
SURVEY.md 8f row F2: clique-contiguous reordering (:345-357), -inf diagonal (:361), row-wise
descending argsort (:362), the rank loop (:367-390) and MR / MRR / MDR / MAP / Top-K (:391-400).
The loop is the reference's own, statement by statement (it is O(N^2) Python: small N only).
The only liberty: np.argsort(..., kind='stable') where the reference uses the default
introsort, whose order among EQUAL scores is implementation-defined; on tie-free rows both
give the same permutation.  Pinned against the reference's own function, executed from
/root/reference by oracle/make_golden_eval.py -> tests/golden/eval_golden.npz.
"""
from __future__ import annotations

from itertools import chain

import numpy as np


def eval_ranks(D, cliques):
    """Returns (ranks, AllMap, iranks_per_row, Ks): ranks / AllMap as the reference's arrays of
    the same names (NaN where untouched); iranks_per_row[i] = the list `iranks` of position i."""
    D = np.array(D, dtype=np.float32)
    N = D.shape[0]
    cliques = [list(c) for c in cliques]
    Ks = np.array([len(c) for c in cliques])
    idx = np.argsort(-Ks)
    Ks = Ks[idx]
    cliques = [cliques[i] for i in idx]
    idx = np.array(list(chain(*cliques)), dtype=int)
    D = D[idx, :]
    D = D[:, idx]
    np.fill_diagonal(D, -np.inf)
    idx = np.argsort(-D, 1, kind='stable')
    ranks = np.nan * np.ones(N)
    startidx = 0
    kidx = 0
    AllMap = np.nan * np.ones(N)
    per_row = []
    for i in range(N):
        if i >= startidx + Ks[kidx]:
            startidx += Ks[kidx]
            kidx += 1
            if Ks[kidx] < 2:
                break
        iranks = []
        for k in range(N):
            diff = idx[i, k] - startidx
            if diff >= 0 and diff < Ks[kidx]:
                iranks.append(k + 1)
        iranks = iranks[0:-1]
        if len(iranks) == 0:
            break
        per_row.append(np.array(iranks, dtype=np.int32))
        ranks[i] = iranks[0]
        P = np.array([float(j) / float(r) for (j, r) in zip(range(1, Ks[kidx]), iranks)])
        AllMap[i] = np.mean(P)
    return ranks, AllMap, per_row, Ks


def eval_ranks_fast(D, cliques):
    """Same result as eval_ranks with the inner O(N) Python loop vectorised (for N in the thousands)."""
    D = np.array(D, dtype=np.float32)
    N = D.shape[0]
    cliques = [list(c) for c in cliques]
    Ks = np.array([len(c) for c in cliques])
    idx = np.argsort(-Ks)
    Ks = Ks[idx]
    cliques = [cliques[i] for i in idx]
    idx = np.array(list(chain(*cliques)), dtype=int)
    ranks = np.nan * np.ones(N)
    AllMap = np.nan * np.ones(N)
    per_row = []
    start = 0
    for K in Ks:
        if K < 2:
            break
        for i in range(start, start + K):
            row = D[idx[i], idx].copy()
            row[i] = -np.inf
            order = np.argsort(-row, kind='stable')
            inclique = (order >= start) & (order < start + K)
            iranks = (np.nonzero(inclique)[0] + 1)[0:-1]
            per_row.append(iranks.astype(np.int32))
            ranks[i] = iranks[0]
            AllMap[i] = np.mean(np.arange(1, K, dtype=np.float64) / iranks.astype(np.float64))
        start += K
    return ranks, AllMap, per_row, Ks


def eval_statistics(D, cliques, topsidx=(1, 10, 100, 1000), fast=False):
    """(MR, MRR, MDR, MAP, tops) as CoverAlgorithm.py:391-400."""
    ranks, AllMap, _, _ = (eval_ranks_fast if fast else eval_ranks)(D, cliques)
    N = np.asarray(D).shape[0]
    MAP = np.nanmean(AllMap)
    ranks = ranks[np.isnan(ranks) == 0]
    MR = np.mean(ranks)
    MRR = 1.0 / N * (np.sum(1.0 / ranks))
    MDR = np.median(ranks)
    tops = np.zeros(len(topsidx))
    for i in range(len(tops)):
        tops[i] = np.sum(ranks <= topsidx[i])
    return MR, MRR, MDR, MAP, tops
