"""oracle/snf_oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

numpy restatement of the similarity-network-fusion front end of EarlySNF: per pair the
(M+N) x (M+N) kernel matrix [[W_SSM_A, W_CSM], [W_CSM^T, W_SSM_B]] of each feature kind, cross-diffusion
(benchmarking/SimilarityFusion.py:50-177, 202-277) and the fused cross block that EarlySNF hands to
csm_to_binary_mutual -> qmax / dmax (benchmarking/EarlySNF.py:44-57, 85-90).  Paths are relative to /root/reference/.
Pinned by tests/golden/snf_golden.npz, produced by oracle/make_golden_snf.py from the reference's own
SimilarityFusion.py imported with its plotting modules stubbed.  Only tests/ and bench.py's CPU leg may import this.

Deviation of the PRODUCT that this oracle follows on purpose: EarlySNF fuses the chroma kernel with the kernel of the
scattering-transform SSMs ('ssms', kymatio: out of scope); here the second feature kind is the 13-bin mfcc sequence
(EarlySNF.py:66-71 computes its CSM anyway), embedded like the chroma (no OTI).
"""
from __future__ import annotations

import numpy as np

from . import crp_oracle, dp


def get_W(D, K, Mu=0.5):
    """SimilarityFusion.py:50-71."""
    DSym = 0.5 * (D + D.T)
    np.fill_diagonal(DSym, 0)
    Neighbs = np.partition(DSym, K + 1, 1)[:, 0:K + 1]
    MeanDist = np.mean(Neighbs, 1) * float(K + 1) / float(K)
    Eps = MeanDist[:, None] + MeanDist[None, :] + DSym
    Eps = Eps / 3
    Denom = (2 * (Mu * Eps) ** 2)
    Denom[Denom == 0] = 1
    return np.exp(-DSym ** 2 / Denom)


def get_WCSM(CSMAB, k1, k2, Mu=0.5):
    """SimilarityFusion.py:74-89."""
    Neighbs1 = np.partition(CSMAB, k2, 1)[:, 0:k2]
    MeanDist1 = np.mean(Neighbs1, 1)
    Neighbs2 = np.partition(CSMAB, k1, 0)[0:k1, :]
    MeanDist2 = np.mean(Neighbs2, 0)
    Eps = MeanDist1[:, None] + MeanDist2[None, :] + CSMAB
    Eps /= 3
    return np.exp(-CSMAB ** 2 / (2 * (Mu * Eps) ** 2))


def get_WCSMSSM(SSMA, SSMB, CSMAB, K, Mu=0.5):
    """SimilarityFusion.py:91-132."""
    M, N = SSMA.shape[0], SSMB.shape[0]
    k1 = int(K * float(M) / (M + N))
    k2 = K - k1
    W = np.zeros((N + M, N + M))
    W[0:M, 0:M] = get_W(SSMA, k1, Mu)
    WC = get_WCSM(CSMAB, k1, k2, Mu)
    W[0:M, M::] = WC
    W[M::, 0:M] = WC.T
    W[M::, M::] = get_W(SSMB, k2, Mu)
    return W


def get_P(W, reg_diag=False):
    """SimilarityFusion.py:134-155."""
    if reg_diag:
        WNoDiag = np.array(W)
        np.fill_diagonal(WNoDiag, 0)
        RowSum = np.sum(WNoDiag, 1)
        RowSum[RowSum == 0] = 1
        return 0.5 * np.eye(W.shape[0]) + 0.5 * WNoDiag / RowSum[:, None]
    RowSum = np.sum(W, 1)
    RowSum[RowSum == 0] = 1
    return W / RowSum[:, None]


def get_S_dense(W, K):
    """SimilarityFusion.py:157-177 as a dense matrix: the K largest entries of every row (the element itself included),
    L1-normalised."""
    N = W.shape[0]
    J = np.argpartition(-W, K, 1)[:, 0:K]
    I = np.tile(np.arange(N)[:, None], (1, K))
    V = W[I, J]
    SNorm = np.sum(V, 1)
    SNorm[SNorm == 0] = 1
    S = np.zeros((N, N))
    S[I, J] = V / SNorm[:, None]
    return S


def snf_ws(Ws, K=5, niters=20, reg_diag=True):
    """SimilarityFusion.py:202-277 without the plotting / timing.  The reference rebinds `Pts = nextPts` at the end of
    the first iteration, after which both names are ONE list: from the second iteration on, feature kind i reads the
    matrices that kinds < i have already updated in this iteration (Gauss-Seidel), and `nextPts[i] *= 0` clears the
    array it is about to overwrite.  Restated with that aliasing."""
    Ps = [get_P(W, reg_diag) for W in Ws]
    Ss = [get_S_dense(W, K) for W in Ws]
    Pts = [np.array(P) for P in Ps]
    nextPts = [np.zeros(P.shape) for P in Pts]
    N = len(Pts)
    for it in range(niters):
        for i in range(N):
            nextPts[i] *= 0
            for k in range(N):
                if i == k:
                    continue
                nextPts[i] += Pts[k]
            nextPts[i] /= float(N - 1)
            A = Ss[i].dot(nextPts[i].T)
            nextPts[i] = Ss[i].dot(A.T)
            if reg_diag:
                PNoDiag = np.array(nextPts[i])
                np.fill_diagonal(PNoDiag, 0)
                RowSum = np.sum(PNoDiag, 1)
                RowSum[RowSum == 0] = 1
                nextPts[i] = 0.5 * np.eye(nextPts[i].shape[0]) + 0.5 * PNoDiag / RowSum[:, None]
        Pts = nextPts
    F = np.zeros(Pts[0].shape)
    for Pt in Pts:
        F += Pt
    return F / len(Pts)


def kernel_blocks(Xi, Xj, m):
    """The three sliding distance matrices of one feature kind (EarlySNF.py:48-56): (SSM_A, SSM_B, CSM) from (bins, L)
    arrays, get_csm + sliding_csm each (CRPUtils.py:67-84, 24-45)."""
    csm = crp_oracle.diag_window(crp_oracle.cross_dist(Xi.T, Xj.T), m)
    ssma = crp_oracle.diag_window(crp_oracle.cross_dist(Xi.T, Xi.T), m)
    ssmb = crp_oracle.diag_window(crp_oracle.cross_dist(Xj.T, Xj.T), m)
    return ssma, ssmb, csm


def early_snf_pair(chroma_i, chroma_j, g_i, g_j, mfcc_i, mfcc_j, m=9, kappa=0.095, niters=3):
    """EarlySNF.similarity for one pair (EarlySNF.py:44-57, 85-90) with (chroma, mfcc) as the two fused feature kinds.
    Returns dict: oti, fused (M, N) float64 cross block BEFORE the sign flip, crp (M, N) uint8, qmax_raw, dmax_raw."""
    shift = crp_oracle.oti(g_i, g_j)
    C1 = np.roll(chroma_i, shift, axis=0)
    ssma, ssmb, csm = kernel_blocks(C1, chroma_j, m)
    M, N = csm.shape
    K = int(kappa * (M + N))
    Ws = [get_WCSMSSM(ssma, ssmb, csm, K)]
    ssma, ssmb, csm = kernel_blocks(mfcc_i, mfcc_j, m)
    Ws.append(get_WCSMSSM(ssma, ssmb, csm, K))
    F = snf_ws(Ws, K=K, niters=niters, reg_diag=True)
    block = F[0:M, M::]
    B = crp_oracle.mutual_mask(-block, kappa)
    q, d = dp.qmax_dmax_serra09(B)
    return {"oti": shift, "fused": block, "crp": B, "qmax_raw": q, "dmax_raw": d, "M": M, "N": N, "K": K}
