// kernels_feat.cuh -- the step just before the scoring path:
// Serra09.load_features (benchmarking/Serra09.py:96-114) on the device.
//   * global chroma  g = sum_t chroma[t, :] / max(.)            (Serra09.py:24-28, 101-102)
//   * block aggregation over `fac` frames, librosa.util.sync(data, arange(0, n, fac), aggregate):
//     slices [fac k, min(fac (k + 1), n)), median for chroma (Serra09.py:104), mean for mfcc (:110)
// All arithmetic is done in the input's precision and in numpy's order, so the results are the
// bits numpy produces: np.sum(axis=0) adds row after row; np.median averages the two middle
// order statistics of an even-sized slice as (a + b) / 2; np.mean sums a contiguous slice
// pairwise (8 partial sums, then the tail) and divides by the count.
#pragma once
#include "common.cuh"

namespace acoss {

constexpr int kFeatMaxFac = 64;  // frames per aggregation block (the reference uses 40)

template <typename T> __device__ __forceinline__ T feat_add(T a, T b);
template <> __device__ __forceinline__ float feat_add<float>(float a, float b) { return __fadd_rn(a, b); }
template <> __device__ __forceinline__ double feat_add<double>(double a, double b) { return __dadd_rn(a, b); }
template <typename T> __device__ __forceinline__ T feat_div(T a, T b);
template <> __device__ __forceinline__ float feat_div<float>(float a, float b) { return __fdiv_rn(a, b); }
template <> __device__ __forceinline__ double feat_div<double>(double a, double b) { return __ddiv_rn(a, b); }

struct FeatArgs {
    const void* in;           // chunk arena on the device
    const int64_t* in_off;    // [songs] element offset of the song inside the chunk arena
    const int32_t* nframes;   // [songs]
    void* out;                // [sum L][bins] frame-major
    const int64_t* out_off;   // [songs] element offset of the song's output
    void* gchroma;            // [songs][bins] or null
    int n_songs, bins, fac, layout, aggregate, zero_nonfinite;
    int blocks_per_song;      // k_feat_aggregate: CTA b handles song b / blocks_per_song
};

// CTA -> (song, block of outputs); threads over (output frame, bin)
template <typename T>
__global__ void __launch_bounds__(128) k_feat_aggregate(FeatArgs A) {
    const int s = (int)(blockIdx.x / (unsigned)A.blocks_per_song);
    const int bx = (int)(blockIdx.x % (unsigned)A.blocks_per_song);
    const int n = A.nframes[s], bins = A.bins, fac = A.fac;
    const int L = (n + fac - 1) / fac;
    const int o = bx * (int)blockDim.x + (int)threadIdx.x;
    if (o >= L * bins) return;
    const int k = o / bins, b = o - k * bins;
    const int w = imin(fac, n - k * fac);
    const T* in = reinterpret_cast<const T*>(A.in) + A.in_off[s];
    T v[kFeatMaxFac];
    for (int i = 0; i < w; ++i) {
        const int t = k * fac + i;
        T x = A.layout == 0 ? in[(int64_t)t * bins + b] : in[(int64_t)b * n + t];
        if (A.zero_nonfinite && !(fabs((double)x) <= 1.7976931348623157e308)) x = (T)0;  // NaN, +-Inf -> 0 (Serra09.py:107-109)
        v[i] = x;
    }
    T res;
    if (A.aggregate == 0) {
        // median: order statistics (w - 1) / 2 and w / 2 by rank counting (stable for equal values)
        const int lo = (w - 1) >> 1, hi = w >> 1;
        T vlo = (T)0, vhi = (T)0;
        bool has_nan = false;
        for (int i = 0; i < w; ++i) {
            const T x = v[i];
            if (x != x) has_nan = true;
            int rank = 0;
            for (int j = 0; j < w; ++j) rank += (v[j] < x || (v[j] == x && j < i)) ? 1 : 0;
            if (rank == lo) vlo = x;
            if (rank == hi) vhi = x;
        }
        res = (lo == hi) ? vlo : feat_div<T>(feat_add<T>(vlo, vhi), (T)2);
        if (has_nan) res = (T)NAN;  // np.median propagates NaN
    } else {
        // mean: numpy's pairwise sum of a contiguous run of w < 128 values, then / w
        T sum;
        if (w < 8) {
            sum = (T)0;
            for (int i = 0; i < w; ++i) sum = feat_add<T>(sum, v[i]);
        } else {
            T r[8];
            for (int q = 0; q < 8; ++q) r[q] = v[q];
            int i = 8;
            for (; i < w - (w % 8); i += 8)
                for (int q = 0; q < 8; ++q) r[q] = feat_add<T>(r[q], v[i + q]);
            sum = feat_add<T>(feat_add<T>(feat_add<T>(r[0], r[1]), feat_add<T>(r[2], r[3])),
                              feat_add<T>(feat_add<T>(r[4], r[5]), feat_add<T>(r[6], r[7])));
            for (; i < w; ++i) sum = feat_add<T>(sum, v[i]);
        }
        res = feat_div<T>(sum, (T)w);
    }
    reinterpret_cast<T*>(A.out)[A.out_off[s] + (int64_t)k * bins + b] = res;
}

// One warp per song, lane = bin (bins <= 32 per pass): sequential sum over the frames, then / max.
template <typename T>
__global__ void __launch_bounds__(128) k_feat_gchroma(FeatArgs A) {
    const int s = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const int lane = threadIdx.x & 31;
    if (s >= A.n_songs) return;
    const int n = A.nframes[s], bins = A.bins;
    const T* in = reinterpret_cast<const T*>(A.in) + A.in_off[s];
    T* g = reinterpret_cast<T*>(A.gchroma) + (int64_t)s * bins;
    // bins may exceed 32 (36-bin chroma): two passes keep every lane's sums in registers
    T acc0 = (T)0, acc1 = (T)0;
    const int b0 = lane, b1 = lane + 32;
    for (int t = 0; t < n; ++t) {
        if (b0 < bins) acc0 = feat_add<T>(acc0, in[(int64_t)t * bins + b0]);
        if (b1 < bins) acc1 = feat_add<T>(acc1, in[(int64_t)t * bins + b1]);
    }
    // np.max: NaN propagates; otherwise the largest sum
    T m = (b0 < bins) ? acc0 : (T)-INFINITY;
    if (b1 < bins && (acc1 > m || acc1 != acc1)) m = acc1;
    bool nan = (m != m);
    for (int d = 16; d >= 1; d >>= 1) {
        const T o = __shfl_xor_sync(kFullMask, m, d);
        if (o > m) m = o;
        nan = nan || (o != o);
    }
    nan = __any_sync(kFullMask, nan);
    if (nan) m = (T)NAN;
    if (b0 < bins) g[b0] = feat_div<T>(acc0, m);
    if (b1 < bins) g[b1] = feat_div<T>(acc1, m);
}

}  // namespace acoss
