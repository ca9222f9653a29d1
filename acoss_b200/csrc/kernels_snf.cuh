// kernels_snf.cuh -- similarity network fusion front end of EarlySNF, float64.
//
// Per pair and feature kind the reference builds the (M+N) x (M+N) kernel matrix
//        [ W(SSM_A)   W(CSM)  ]
//        [ W(CSM)^T   W(SSM_B)]                     (SimilarityFusion.py:91-132, get_WCSMSSM)
// from the windowed Euclidean distances of the two songs (EarlySNF.py:44-56: get_csm + sliding_csm, CRPUtils.py:24-45,
// 67-84), turns it into a row-normalised P with a regularised diagonal and a K-nearest-neighbour truncated S
// (SimilarityFusion.py:134-177), runs `niters` rounds of cross diffusion P_i <- S_i P_other S_i^T (:202-277) and hands
// minus the fused cross block to csm_to_binary_mutual -> qmax / dmax (EarlySNF.py:85-90).  Here:
//   k_snf_dist   n x n windowed distances of one feature kind (direct sum of squared differences; the roll by the OTI is
//                applied to song i's frames wherever they appear)
//   k_snf_meank  per row the mean of the k smallest distances of its SSM part and of its CSM part (np.partition + mean)
//   k_snf_w      D -> W in place (exp(-D^2 / (2 (mu eps)^2)), eps = (mean_r + mean_c + D) / 3)
//   k_snf_regp   row normalisation with the regularised diagonal (get_P(reg_diag=True)); also after every diffusion
//   k_snf_s      the K largest entries of every row of W, L1-normalised (get_S), stored [k][row]
//   k_snf_xst    T = X S^T        (CTA per row of X staged in shared memory, threads gather by S's indices)
//   k_snf_sy     Y = S T          (CTA per row of S, K rows of T accumulated, coalesced)
//   k_snf_fused  -(P_0 + P_1) / 2 restricted to the cross block, into the dense arena of the EarlySNF tail
// S P S^T is evaluated as S (P S^T): the same product, K (M+N)^2 multiply-adds per factor instead of the reference's
// sparse-times-dense calls; summation orders differ from scipy's, results agree to ~1e-15 relative.
#pragma once
#include "common.cuh"

namespace acoss {

constexpr int kSnfMaxN = 2048;     // M + N of one pair (a row is staged in 16 KB of shared memory)
constexpr int kSnfMaxK = 256;      // neighbours per row
constexpr int kSnfThreads = 256;

struct SnfDesc {
    int64_t fi, fj;        // first frame of song i / j in the arenas
    int32_t M, N, n, K, k1, k2, oti, pad;
    int64_t buf[5];        // n x n float64 matrices: 0, 1 = X_kind (D -> W -> scratch), 2, 3 = P, 4 = T   (doubles)
    int64_t md[2];         // per kind: 2 n doubles: [0, n) mean over the row's SSM part, [n, 2n) over its CSM part
    int64_t s[2];          // per kind: K n (index, value) entries, entry k of row r at k * n + r
    int64_t out;           // M x N fused cross block in the dense arena (doubles)
};

__device__ __forceinline__ int snf_block_sum(int v, int* s_red) {
    v = __reduce_add_sync(kFullMask, v);
    const int warp = threadIdx.x >> 5;
    __syncthreads();                       // s_red may still be read from the previous call
    if ((threadIdx.x & 31) == 0) s_red[warp] = v;
    __syncthreads();
    int t = 0;
#pragma unroll
    for (int w = 0; w < kSnfThreads / 32; ++w) t += s_red[w];
    return t;
}
__device__ __forceinline__ double snf_block_sum_d(double v, double* s_redd) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFullMask, v, o);
    const int warp = threadIdx.x >> 5;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_redd[warp] = v;
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < kSnfThreads / 32; ++w) t += s_redd[w];
    return t;
}

// k-th smallest (1-based) of len non-negative doubles given as bit patterns in shared memory: the largest P with
// #(keys < P) < k, found bit by bit.  All threads return the same value.
__device__ __forceinline__ unsigned long long snf_kth_smallest(const unsigned long long* keys, int len, int k, int* s_red) {
    unsigned long long prefix = 0ull;
    for (int bit = 62; bit >= 0; --bit) {
        const unsigned long long cand = prefix | (1ull << bit);
        int cnt = 0;
        for (int i = threadIdx.x; i < len; i += kSnfThreads) cnt += keys[i] < cand ? 1 : 0;
        cnt = snf_block_sum(cnt, s_red);
        if (cnt < k) prefix = cand;
    }
    return prefix;
}

// ---------------------------------------------------------------- distances
// grid (n_max, pairs): row a of the pair's n x n matrix of windowed distances of one feature kind, into mats + buf[kind].
__global__ void __launch_bounds__(kSnfThreads) k_snf_dist(const SnfDesc* __restrict__ descs, const double* __restrict__ frames,
                                                          double* __restrict__ mats, int bins, int m, int kind, int do_roll) {
    __shared__ double u[32 * kMaxBins];  // the m frames of this row's window (m <= 32)
    const SnfDesc d = descs[blockIdx.y];
    const int a = blockIdx.x;
    if (a >= d.n) return;
    const int rot = do_roll ? d.oti : 0;
    const bool a_i = a < d.M;
    const int64_t fa = a_i ? d.fi + a : d.fj + (a - d.M);
    for (int e = threadIdx.x; e < m * bins; e += kSnfThreads) {
        const int k = e / bins, b = e - k * bins;
        int src = b;
        if (a_i && rot) {   // np.roll(chroma_i, oti, axis=0)[b] = chroma_i[(b - oti) mod bins]
            src = b - rot;
            if (src < 0) src += bins;
        }
        u[e] = frames[(fa + k) * bins + src];
    }
    __syncthreads();
    double* row = mats + d.buf[kind] + (int64_t)a * d.n;
    for (int c = threadIdx.x; c < d.n; c += kSnfThreads) {
        const bool c_i = c < d.M;
        const double* v = frames + (c_i ? d.fi + c : d.fj + (c - d.M)) * bins;
        double acc = 0.0;
        for (int k = 0; k < m; ++k) {
            for (int b = 0; b < bins; ++b) {
                int src = b;
                if (c_i && rot) {
                    src = b - rot;
                    if (src < 0) src += bins;
                }
                const double df = u[k * bins + b] - v[k * bins + src];
                acc = fma(df, df, acc);
            }
        }
        row[c] = (c == a) ? 0.0 : sqrt(acc);
    }
}

// ---------------------------------------------------------------- mean of the k smallest
// grid (n_max, pairs): row r of D.  SSM part: the k' + 1 smallest of its own song's columns (the zero diagonal
// included), scaled (k' + 1) / k' (get_W, SimilarityFusion.py:61-63; k' = k1 for song i rows, k2 for song j rows);
// CSM part: the k'' smallest of the other song's columns (get_WCSM :84-87; k'' = k2 for song i rows -- Neighbs1 --
// and k1 for song j rows, which are the columns of the CSM -- Neighbs2).
__global__ void __launch_bounds__(kSnfThreads) k_snf_meank(const SnfDesc* __restrict__ descs, double* __restrict__ mats,
                                                           double* __restrict__ md_all, int kind) {
    __shared__ unsigned long long keys[kSnfMaxN];
    __shared__ int s_red[kSnfThreads / 32];
    __shared__ double s_redd[kSnfThreads / 32];
    const SnfDesc d = descs[blockIdx.y];
    const int r = blockIdx.x;
    if (r >= d.n) return;
    const double* row = mats + d.buf[kind] + (int64_t)r * d.n;
    for (int c = threadIdx.x; c < d.n; c += kSnfThreads) keys[c] = (unsigned long long)__double_as_longlong(row[c]);
    __syncthreads();
    const bool r_i = r < d.M;
    double* md = md_all + d.md[kind];
    for (int part = 0; part < 2; ++part) {
        // part 0: own song's columns; part 1: the other song's
        const bool own = part == 0;
        const int c0 = (r_i == own) ? 0 : d.M;
        const int len = (r_i == own) ? d.M : d.N;
        const int kq = own ? (r_i ? d.k1 : d.k2) : (r_i ? d.k2 : d.k1);
        const int k = own ? kq + 1 : kq;
        double mean = 0.0;
        if (k >= 1 && k <= len) {
            const unsigned long long kth = snf_kth_smallest(keys + c0, len, k, s_red);
            double part_sum = 0.0;
            int less = 0;
            for (int i = threadIdx.x; i < len; i += kSnfThreads)
                if (keys[c0 + i] < kth) {
                    part_sum += __longlong_as_double((long long)keys[c0 + i]);
                    ++less;
                }
            less = snf_block_sum(less, s_red);
            part_sum = snf_block_sum_d(part_sum, s_redd);
            const double total = part_sum + (double)(k - less) * __longlong_as_double((long long)kth);
            mean = total / (double)k;
            if (own) mean = mean * (double)(kq + 1) / (double)kq;
        }
        if (threadIdx.x == 0) md[part * d.n + r] = mean;
        __syncthreads();
    }
}

// ---------------------------------------------------------------- D -> W in place
__global__ void __launch_bounds__(kSnfThreads) k_snf_w(const SnfDesc* __restrict__ descs, double* __restrict__ mats,
                                                       const double* __restrict__ md_all, int kind) {
    const SnfDesc d = descs[blockIdx.y];
    const int r = blockIdx.x;
    if (r >= d.n) return;
    double* row = mats + d.buf[kind] + (int64_t)r * d.n;
    const double* md = md_all + d.md[kind];
    const bool r_i = r < d.M;
    const double own_r = md[r], oth_r = md[d.n + r];
    for (int c = threadIdx.x; c < d.n; c += kSnfThreads) {
        const bool c_i = c < d.M;
        const double D = row[c];
        double w;
        if (r_i == c_i) {   // SSM block (get_W): Eps = (MeanDist[r] + MeanDist[c] + D) / 3, zero denominators -> 1
            const double eps = ((own_r + md[c]) + D) / 3.0;
            double den = 2.0 * ((0.5 * eps) * (0.5 * eps));
            if (den == 0.0) den = 1.0;
            w = exp(-(D * D) / den);
        } else {            // CSM block (get_WCSM): Eps = (MeanDist1[row of song i] + MeanDist2[column of song j] + D) / 3
            const double eps = ((oth_r + md[d.n + c]) + D) / 3.0;
            w = exp(-(D * D) / (2.0 * ((0.5 * eps) * (0.5 * eps))));
        }
        row[c] = w;
    }
}

// ---------------------------------------------------------------- P = 0.5 I + 0.5 W_nodiag / rowsum
// (get_P(reg_diag=True), SimilarityFusion.py:143-150; the same lines close every diffusion step, :262-268).
// src and dst may be the same buffer (a row is read completely before it is written).
__global__ void __launch_bounds__(kSnfThreads) k_snf_regp(const SnfDesc* __restrict__ descs, double* __restrict__ mats, int src_id,
                                                          int dst_id) {
    __shared__ double s_redd[kSnfThreads / 32];
    const SnfDesc d = descs[blockIdx.y];
    const int r = blockIdx.x;
    if (r >= d.n) return;
    const double* src = mats + d.buf[src_id] + (int64_t)r * d.n;
    double* dst = mats + d.buf[dst_id] + (int64_t)r * d.n;
    double acc = 0.0;
    for (int c = threadIdx.x; c < d.n; c += kSnfThreads)
        if (c != r) acc += src[c];
    double sum = snf_block_sum_d(acc, s_redd);
    if (sum == 0.0) sum = 1.0;
    for (int c = threadIdx.x; c < d.n; c += kSnfThreads) dst[c] = (c == r) ? 0.5 : 0.5 * src[c] / sum;
}

// ---------------------------------------------------------------- S: the K largest of every row of W, normalised
__global__ void __launch_bounds__(kSnfThreads) k_snf_s(const SnfDesc* __restrict__ descs, const double* __restrict__ mats,
                                                       int32_t* __restrict__ sidx_all, double* __restrict__ sval_all, int kind) {
    __shared__ unsigned long long keys[kSnfMaxN];
    __shared__ int s_red[kSnfThreads / 32];
    __shared__ double s_sum;
    const SnfDesc d = descs[blockIdx.y];
    const int r = blockIdx.x;
    if (r >= d.n) return;
    const double* row = mats + d.buf[kind] + (int64_t)r * d.n;
    for (int c = threadIdx.x; c < d.n; c += kSnfThreads) keys[c] = (unsigned long long)__double_as_longlong(row[c]);
    __syncthreads();
    const int K = d.K;
    // K-th largest = (n - K + 1)-th smallest; W is in (0, 1]: positive doubles order like their bit patterns
    const unsigned long long thr = snf_kth_smallest(keys, d.n, d.n - K + 1, s_red);
    int gt = 0;
    for (int c = threadIdx.x; c < d.n; c += kSnfThreads) gt += keys[c] > thr ? 1 : 0;
    gt = snf_block_sum(gt, s_red);
    const int need_eq = K - gt;   // entries equal to the threshold that still belong (lowest columns first)
    int32_t* sidx = sidx_all + d.s[kind];
    double* sval = sval_all + d.s[kind];
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        int pos = 0, eq_taken = 0;
        double sum = 0.0;
        for (int c0 = 0; c0 < d.n; c0 += 32) {
            const int c = c0 + lane;
            const unsigned long long kv = c < d.n ? keys[c] : 0ull;
            const bool is_gt = c < d.n && kv > thr, is_eq = c < d.n && kv == thr;
            const unsigned eqm = __ballot_sync(kFullMask, is_eq);
            const int eq_before = eq_taken + __popc(eqm & ((1u << lane) - 1u));
            const bool take = is_gt || (is_eq && eq_before < need_eq);
            const unsigned tm = __ballot_sync(kFullMask, take);
            if (take) {
                const int k = pos + __popc(tm & ((1u << lane) - 1u));
                sidx[(int64_t)k * d.n + r] = c;
                sval[(int64_t)k * d.n + r] = __longlong_as_double((long long)kv);
                sum += __longlong_as_double((long long)kv);
            }
            pos += __popc(tm);
            eq_taken += __popc(eqm);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(kFullMask, sum, o);
        if (lane == 0) s_sum = (sum == 0.0) ? 1.0 : sum;
    }
    __syncthreads();
    const double norm = s_sum;
    for (int k = threadIdx.x; k < K; k += kSnfThreads) sval[(int64_t)k * d.n + r] /= norm;
}

// ---------------------------------------------------------------- T = X S^T
// grid (n_max, pairs): row c of X in shared memory; T[c][r] = sum_k X[c][idx_k(r)] * val_k(r).
__global__ void __launch_bounds__(kSnfThreads) k_snf_xst(const SnfDesc* __restrict__ descs, double* __restrict__ mats,
                                                         const int32_t* __restrict__ sidx_all, const double* __restrict__ sval_all,
                                                         int x_id, int kind, int t_id) {
    __shared__ double xrow[kSnfMaxN];
    const SnfDesc d = descs[blockIdx.y];
    const int c = blockIdx.x;
    if (c >= d.n) return;
    const double* X = mats + d.buf[x_id] + (int64_t)c * d.n;
    for (int j = threadIdx.x; j < d.n; j += kSnfThreads) xrow[j] = X[j];
    __syncthreads();
    const int32_t* sidx = sidx_all + d.s[kind];
    const double* sval = sval_all + d.s[kind];
    double* T = mats + d.buf[t_id] + (int64_t)c * d.n;
    for (int r = threadIdx.x; r < d.n; r += kSnfThreads) {
        double acc = 0.0;
        for (int k = 0; k < d.K; ++k) acc = fma(xrow[sidx[(int64_t)k * d.n + r]], sval[(int64_t)k * d.n + r], acc);
        T[r] = acc;
    }
}

// ---------------------------------------------------------------- Y = S T
// grid (n_max, pairs): row r of S in shared memory; Y[r][c] = sum_k val_k(r) * T[idx_k(r)][c].
__global__ void __launch_bounds__(kSnfThreads) k_snf_sy(const SnfDesc* __restrict__ descs, double* __restrict__ mats,
                                                        const int32_t* __restrict__ sidx_all, const double* __restrict__ sval_all,
                                                        int kind, int t_id, int y_id) {
    __shared__ int32_t idx[kSnfMaxK];
    __shared__ double val[kSnfMaxK];
    const SnfDesc d = descs[blockIdx.y];
    const int r = blockIdx.x;
    if (r >= d.n) return;
    for (int k = threadIdx.x; k < d.K; k += kSnfThreads) {
        idx[k] = sidx_all[d.s[kind] + (int64_t)k * d.n + r];
        val[k] = sval_all[d.s[kind] + (int64_t)k * d.n + r];
    }
    __syncthreads();
    const double* T = mats + d.buf[t_id];
    double* Y = mats + d.buf[y_id] + (int64_t)r * d.n;
    for (int c = threadIdx.x; c < d.n; c += kSnfThreads) {
        double acc = 0.0;
        for (int k = 0; k < d.K; ++k) acc = fma(val[k], T[(int64_t)idx[k] * d.n + c], acc);
        Y[c] = acc;
    }
}

// ---------------------------------------------------------------- minus the fused cross block
// fused_score (SimilarityFusion.py:179-186): (P_0 + P_1) / 2; EarlySNF.py:86: -csm[0:M, M::].
__global__ void __launch_bounds__(kSnfThreads) k_snf_fused(const SnfDesc* __restrict__ descs, const double* __restrict__ mats,
                                                           double* __restrict__ dense, int a_id, int b_id) {
    const SnfDesc d = descs[blockIdx.y];
    const int a = blockIdx.x;
    if (a >= d.M) return;
    const double* A = mats + d.buf[a_id] + (int64_t)a * d.n + d.M;
    const double* B = mats + d.buf[b_id] + (int64_t)a * d.n + d.M;
    double* out = dense + d.out + (int64_t)a * d.N;
    for (int b = threadIdx.x; b < d.N; b += kSnfThreads) out[b] = -((A[b] + B[b]) / 2.0);
}

}  // namespace acoss
