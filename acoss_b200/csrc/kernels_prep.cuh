// kernels_prep.cuh -- per-frame squared norms and the optimal transposition index.
#pragma once
#include "common.cuh"

namespace acoss {

// One thread per frame: norms[f] = sum_b frames[f][b]^2 in float64
// (the np.sum(X**2, 1) terms of get_csm, CRPUtils.py:82).
template <typename Tin>
__global__ void __launch_bounds__(256) k_norms(const Tin* __restrict__ frames, double* __restrict__ norms,
                                                int64_t total_frames, int bins) {
    int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= total_frames) return;
    const Tin* p = frames + f * bins;
    double acc = 0.0;
    for (int b = 0; b < bins; ++b) {
        double v = (double)p[b];
        acc = fma(v, v, acc);
    }
    norms[f] = acc;
}

// numpy's pairwise summation of a contiguous float64 vector of n < 128 elements
// (what np.sum does inside get_oti, CRPUtils.py:133): eight running partial sums
// combined as ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)), then the tail added in order.
// Reproducing the order makes near-tied agreements resolve to the same argmax.
__device__ __forceinline__ double numpy_sum_small(const double* a, int n) {
    if (n < 8) {
        double res = 0.0;
        for (int i = 0; i < n; ++i) res = __dadd_rn(res, a[i]);
        return res;
    }
    double r[8];
    for (int k = 0; k < 8; ++k) r[k] = a[k];
    int i = 8;
    for (; i < n - (n % 8); i += 8)
        for (int k = 0; k < 8; ++k) r[k] = __dadd_rn(r[k], a[i + k]);
    double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                           __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
    for (; i < n; ++i) res = __dadd_rn(res, a[i]);
    return res;
}

// One thread per pair: oti = argmax_s sum_k g_i[(k - s) mod bins] * g_j[k]
// (get_oti, CRPUtils.py:109-136; np.roll(C1, s)[k] = C1[(k - s) mod n]; np.argmax keeps
// the first maximum).  Products and sums are rounded separately (no FMA), like numpy.
__global__ void __launch_bounds__(128) k_oti(PairWork* __restrict__ pairs, int n_pairs,
                                             const double* __restrict__ gchroma, int bins, int no_oti) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pairs) return;
    if (no_oti) {
        pairs[p].oti = 0;
        return;
    }
    const double* gi = gchroma + (int64_t)pairs[p].song_i * bins;
    const double* gj = gchroma + (int64_t)pairs[p].song_j * bins;
    double a[kMaxBins], b[kMaxBins], prod[kMaxBins];
    for (int k = 0; k < bins; ++k) {
        a[k] = gi[k];
        b[k] = gj[k];
    }
    int best = 0;
    double best_v = 0.0;
    for (int s = 0; s < bins; ++s) {
        for (int k = 0; k < bins; ++k) {
            int src = k - s;
            if (src < 0) src += bins;
            prod[k] = __dmul_rn(a[src], b[k]);
        }
        double v = numpy_sum_small(prod, bins);
        if (s == 0 || v > best_v) {
            best_v = v;
            best = s;
        }
    }
    pairs[p].oti = best;
}

// Scores of a chunk into device-resident N x N score matrices (acoss_bind_score_matrices): what Serra09.similarity
// stores, raw / (M + N) (Serra09.py:174-175), computed like the host does (float64 division, rounded to the matrices'
// float32) and written at [row][col] = (song id of the pair's first song, of its second).  One thread per pair.
__global__ void __launch_bounds__(128) k_scatter_scores(const PairWork* __restrict__ pairs, int n_pairs,
                                                        const float* __restrict__ qraw, const float* __restrict__ draw,
                                                        const int32_t* __restrict__ song_of_slot, float* __restrict__ Dq,
                                                        float* __restrict__ Dd, int64_t n) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pairs) return;
    const PairWork w = pairs[p];
    const int64_t i = song_of_slot ? song_of_slot[w.song_i] : w.song_i;
    const int64_t j = song_of_slot ? song_of_slot[w.song_j] : w.song_j;
    if (i < 0 || j < 0 || i >= n || j >= n) return;
    double mn = (double)(w.M + w.N);
    if (mn <= 0.0) mn = 1.0;
    if (Dq) Dq[i * n + j] = (float)((double)qraw[p] / mn);
    if (Dd) Dd[i * n + j] = (float)((double)draw[p] / mn);
}

}  // namespace acoss
