// kernels_sweep.cuh -- the cross-similarity sweep: squared Euclidean cross-distance of two
// songs' frames (get_csm, CRPUtils.py:67-84), implicit delay embedding as a windowed sum along
// diagonals (sliding_csm, CRPUtils.py:24-45; the explicit embedding of sliding_window,
// CRPUtils.py:8-22, is never materialised), and either
//   MODE 0: per-row k-nearest-neighbour separators (csm_to_binary's argpartition,
//           CRPUtils.py:186-199) -- run once per orientation so that rows AND columns of the
//           pair get separators (csm_to_binary_mutual, CRPUtils.py:201-219), or
//   MODE 1: the mutual-kNN binary cross-recurrence plot, bit-packed.
//
// One CTA sweeps one (pair, orientation) top to bottom.  Thread t owns CPL = 4 consecutive
// columns (frames of the column song, held in registers for the whole sweep, pre-scaled by
// -2); the row song's frame is broadcast from a small shared-memory staging buffer.  Per cell:
//   C[a][b]  = max(0, |x_a|^2 + |y_b|^2 - 2 x_a.y_b)                  (12 DFMA)
//   P[a][b]  = P[a-1][b-1] + C[a][b]                                  (prefix along the diagonal)
//   T[a][b]  = P[a][b] - P[a-m][b-m] = sum_{k<m} C[a-k][b-k]          (= S[a-m+1][b-m+1]^2)
// P rows live in an (m+1)-deep ring (shared memory, or a per-CTA global scratch when the
// songs are too long).  The reference ranks sqrt(T); ranking T is order-equivalent.
// All arithmetic is float64: the reference's default features (crema) are float64 and the
// CRP must be bit-exact, which float32 distances cannot guarantee.
#pragma once
#include "common.cuh"

namespace acoss {

struct SweepArgs {
    const double* frames;   // [total_frames][BINS]
    const double* norms;    // [total_frames]
    const int64_t* offs;    // [n_songs] first frame of each song
    const PairWork* pairs;
    const uint32_t* items;  // MODE 0: pair*2 + orientation; MODE 1: pair
    int n_items;
    unsigned int* counter;  // work-queue head (zeroed before launch)
    double* seps;
    uint32_t* crp;
    int m;
    int R;                  // strip rows (MODE 0), multiple of the warp count
    int Wmax;               // row stride (doubles) of ring / strip, >= every item's padded width
    int use_gscratch;       // ring + strip in global scratch instead of shared memory
    double* gscratch;       // [gridDim.x][gscratch_stride]
    int64_t gscratch_stride;
    uint32_t flags;
};

constexpr int kSweepCPL = 4;
constexpr int kStageRows = 32;  // rows of the row-song staged per refill

__host__ __device__ inline size_t sweep_smem_fixed_bytes(int Wmax, int bins) {
    size_t bitrows = (size_t)2 * (Wmax / 32 + 2) * sizeof(uint32_t);
    bitrows = (bitrows + 15) & ~(size_t)15;
    size_t xs = (size_t)kStageRows * (bins + 1) * sizeof(double);
    return bitrows + xs;
}

// k-th smallest separator of row[0..n): returns s with exactly k elements <= s (s is the
// midpoint between the k-th and (k+1)-th smallest, so later comparisons have slack).
// Warp-cooperative, all lanes return the same value.  k <= 0 -> -inf, k >= n -> +inf.
// Exact float64 ties across the k-th position cannot be split by a threshold: the tied group
// is left out (the reference's own choice there is numpy-introselect-order, i.e. undefined).
template <typename T>
__device__ inline double warp_select_sep(const T* row, int n, int64_t stride, int k, double guess, int lane) {
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    if (k <= 0) return -INF;
    if (k >= n) return INF;
    double vmin = INF, vmax = -INF;
    for (int j = lane; j < n; j += 32) {
        double v = (double)row[j * stride];
        vmin = fmin(vmin, v);
        vmax = fmax(vmax, v);
    }
    vmin = warp_min(vmin);
    vmax = warp_max(vmax);
    double lo = vmin, hi = vmax, t = vmin;
    int clo, chi = n;
    {
        int c = 0;
        for (int j = lane; j < n; j += 32) c += ((double)row[j * stride] <= lo) ? 1 : 0;
        clo = warp_sum(c);
    }
    if (clo > k) return -INF;  // ties at the minimum straddle k
    bool found = (clo == k);
    if (!found) {
        const bool use_guess = (guess > lo) && (guess < hi);
        for (int iter = 0; iter < 400 && !found; ++iter) {
            if (iter == 0 && use_guess) {
                t = guess;
            } else {
                if (iter & 1)
                    t = 0.5 * lo + 0.5 * hi;
                else
                    t = lo + (hi - lo) * ((double)(k - clo) / (double)(chi - clo));
                if (!(t > lo && t < hi)) {
                    t = 0.5 * lo + 0.5 * hi;
                    if (!(t > lo && t < hi)) break;  // bracket collapsed: exact ties
                }
            }
            int c = 0;
            for (int j = lane; j < n; j += 32) c += ((double)row[j * stride] <= t) ? 1 : 0;
            c = warp_sum(c);
            if (c == k) {
                found = true;
            } else if (c < k) {
                lo = t;
                clo = c;
            } else {
                hi = t;
                chi = c;
            }
        }
    }
    if (!found) t = lo;
    double vk = -INF, vk1 = INF;
    for (int j = lane; j < n; j += 32) {
        double v = (double)row[j * stride];
        if (v <= t)
            vk = fmax(vk, v);
        else
            vk1 = fmin(vk1, v);
    }
    vk = warp_max(vk);
    vk1 = warp_min(vk1);
    double sep = vk + 0.5 * (vk1 - vk);
    if (!(sep < vk1)) sep = vk;
    return sep;
}

template <int BINS, int THREADS, int MODE>
__global__ void __launch_bounds__(THREADS) k_sweep(SweepArgs A) {
    constexpr int CPL = kSweepCPL;
    constexpr int NW = THREADS / 32;
    constexpr int XS = BINS + 1;  // staged row: BINS values + the squared norm
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    ACOSS_DYNSMEM(smem_raw);
    __shared__ int s_item;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int m = A.m, RD = m + 1, R = A.R;
    const int Wmax = A.Wmax;

    // shared-memory carve-up: [bit rows][x staging][ring][strip]
    uint32_t* bitrow = reinterpret_cast<uint32_t*>(smem_raw);
    const int bitrow_words = Wmax / 32 + 2;
    size_t off = ((size_t)2 * bitrow_words * sizeof(uint32_t) + 15) & ~(size_t)15;
    double* xs = reinterpret_cast<double*>(smem_raw + off);
    off += (size_t)kStageRows * XS * sizeof(double);
    double* ring = A.use_gscratch ? (A.gscratch + (int64_t)blockIdx.x * A.gscratch_stride)
                                  : reinterpret_cast<double*>(smem_raw + off);
    double* strip = ring + (size_t)RD * Wmax;

    for (;;) {
        if (tid == 0) s_item = (int)atomicAdd(A.counter, 1u);
        __syncthreads();
        const int it = s_item;
        __syncthreads();
        if (it >= A.n_items) break;

        const uint32_t item = A.items[it];
        int pidx, orient;
        if (MODE == 0) {
            pidx = (int)(item >> 1);
            orient = (int)(item & 1u);
        } else {
            pidx = (int)item;
            orient = A.pairs[pidx].swap;
        }
        const PairWork pw = A.pairs[pidx];
        const int rsong = orient ? pw.song_j : pw.song_i;
        const int csong = orient ? pw.song_i : pw.song_j;
        const int Lr = orient ? pw.Lj : pw.Li;
        const int Lc = orient ? pw.Li : pw.Lj;
        const int rshift = orient ? 0 : pw.oti;  // OTI rolls the first song of the pair
        const int cshift = orient ? pw.oti : 0;
        const int Mo = Lr - m + 1, No = Lc - m + 1;  // output rows / columns
        if (Mo <= 0 || No <= 0) continue;
        const int kk = orient ? pw.k_col : pw.k_row;  // neighbours per output row
        const int G = (Lc + THREADS * CPL - 1) / (THREADS * CPL);
        const int W = Wmax;
        const double* rframes = A.frames + A.offs[rsong] * BINS;
        const double* cframes = A.frames + A.offs[csong] * BINS;
        const double* rnorms = A.norms + A.offs[rsong];
        const double* cnorms = A.norms + A.offs[csong];
        const int64_t rsep_off = pw.sep_off + (orient ? pw.M : 0);
        const int64_t csep_off = pw.sep_off + (orient ? 0 : pw.M);
        const bool nonmutual = (A.flags & kFlagNonMutual) != 0;

        double y2[CPL][BINS], ny[CPL], cs[CPL], Pprev[CPL];
#pragma unroll
        for (int r = 0; r < CPL; ++r) {
            Pprev[r] = 0.0;
            cs[r] = -INF;
            ny[r] = 0.0;
        }
        double guess = INF - INF;  // NaN: no guess yet

        for (int a = 0; a < Lr; ++a) {
            // ---- stage the next kStageRows frames of the row song (rolled by rshift)
            if ((a % kStageRows) == 0) {
                for (int idx = tid; idx < kStageRows * XS; idx += THREADS) {
                    const int rr = idx / XS, k = idx - rr * XS;
                    const int a2 = a + rr;
                    double v = 0.0;
                    if (a2 < Lr) {
                        if (k < BINS) {
                            int src = k - rshift;
                            if (src < 0) src += BINS;
                            v = rframes[(int64_t)a2 * BINS + src];
                        } else {
                            v = rnorms[a2];
                        }
                    }
                    xs[idx] = v;
                }
                __syncthreads();
            }
            const double* xr = xs + (a % kStageRows) * XS;
            double x[BINS];
#pragma unroll
            for (int k = 0; k < BINS; ++k) x[k] = xr[k];
            const double nx = xr[BINS];

            const int i = a - (m - 1);  // output row finished by this row
            double rs = -INF;
            if (MODE == 1 && i >= 0) rs = (nonmutual && orient) ? INF : A.seps[rsep_off + i];

            const int slot0 = a % RD;
            const double* ring1 = ring + (size_t)((a + RD - 1) % RD) * W;  // row a-1
            const double* ringm = ring + (size_t)((a + 1) % RD) * W;       // row a-m
            double* ring0 = ring + (size_t)slot0 * W;

            for (int g = 0; g < G; ++g) {
                const int b0 = (g * THREADS + tid) * CPL;
                if (G > 1 || a == 0) {
#pragma unroll
                    for (int r = 0; r < CPL; ++r) {
                        const int b = b0 + r;
                        if (b < Lc) {
#pragma unroll
                            for (int k = 0; k < BINS; ++k) {
                                int src = k - cshift;
                                if (src < 0) src += BINS;
                                y2[r][k] = -2.0 * cframes[(int64_t)b * BINS + src];
                            }
                            ny[r] = cnorms[b];
                        } else {
#pragma unroll
                            for (int k = 0; k < BINS; ++k) y2[r][k] = 0.0;
                            ny[r] = 0.0;
                        }
                        if (MODE == 1) {
                            const int j = b - (m - 1);
                            cs[r] = -INF;
                            if (j >= 0 && j < No) cs[r] = (nonmutual && !orient) ? INF : A.seps[csep_off + j];
                        }
                    }
                }
                double P[CPL], T[CPL];
#pragma unroll
                for (int r = 0; r < CPL; ++r) {
                    double acc = nx + ny[r];
#pragma unroll
                    for (int k = 0; k < BINS; ++k) acc = fma(x[k], y2[r][k], acc);
                    const double C = fmax(acc, 0.0);
                    const int b = b0 + r;
                    double pd;
                    if (r == 0 || G > 1)
                        pd = (a > 0 && b > 0) ? ring1[b - 1] : 0.0;
                    else
                        pd = Pprev[r - 1];
                    P[r] = pd + C;
                    const double po = (a >= m && b >= m) ? ringm[b - m] : 0.0;
                    T[r] = P[r] - po;
                }
#pragma unroll
                for (int r = 0; r < CPL; ++r) {
                    ring0[b0 + r] = P[r];
                    Pprev[r] = P[r];
                }
                if (MODE == 0) {
                    if (i >= 0) {
                        double* srow = strip + (size_t)(i % R) * W;
#pragma unroll
                        for (int r = 0; r < CPL; ++r) srow[b0 + r] = T[r];
                    }
                } else {
                    unsigned nib = 0;
#pragma unroll
                    for (int r = 0; r < CPL; ++r) nib |= ((T[r] <= rs) && (T[r] <= cs[r])) ? (1u << r) : 0u;
                    unsigned v = nib << (4 * (lane & 7));
                    v |= __shfl_xor_sync(kFullMask, v, 1);
                    v |= __shfl_xor_sync(kFullMask, v, 2);
                    v |= __shfl_xor_sync(kFullMask, v, 4);
                    if ((lane & 7) == 0) bitrow[(a & 1) * bitrow_words + (g * THREADS + tid) / 8] = v;
                }
            }
            __syncthreads();

            if (MODE == 0) {
                if (i >= 0 && (((i + 1) % R) == 0 || a == Lr - 1)) {
                    const int nrows = (i % R) + 1;
                    for (int rr = warp; rr < nrows; rr += NW) {
                        const double* srow = strip + (size_t)rr * W + (m - 1);
                        const double sep = warp_select_sep(srow, No, 1, kk, guess, lane);
                        guess = sep;
                        if (lane == 0) A.seps[rsep_off + (i - (nrows - 1) + rr)] = sep;
                    }
                    __syncthreads();
                }
            } else {
                if (i >= 0) {
                    const int wp = pw.crp_wp;
                    const unsigned sh = (unsigned)(m - 1) & 31u;
                    const int wo = (m - 1) >> 5;
                    const uint32_t* br = bitrow + (a & 1) * bitrow_words;
                    uint32_t* out = A.crp + pw.crp_off + (int64_t)i * wp;
                    const int nwords = G * THREADS * CPL / 32;  // words written this row
                    for (int w = tid; w < wp; w += THREADS) {
                        const uint32_t lo = (w + wo < nwords) ? br[w + wo] : 0u;
                        const uint32_t hi = (w + wo + 1 < nwords) ? br[w + wo + 1] : 0u;
                        out[w] = __funnelshift_r(lo, hi, sh);
                    }
                }
            }
        }
        __syncthreads();
    }
}

}  // namespace acoss
