// kernels_dense.cuh -- binarisation of caller-supplied dense dissimilarity matrices
// (csm_to_binary / csm_to_binary_mutual, CRPUtils.py:169-219) for the EarlySNF tail
// (EarlySNF.py:86-90, EarlySNF_Old.py:200-218), and small format helpers.
#pragma once
#include "common.cuh"
#include "kernels_sweep.cuh"

namespace acoss {

struct DenseDesc {
    int64_t csm_off;  // elements
    int64_t sep_off;  // doubles: [rows][cols]
    int64_t crp_off;  // words
    int32_t rows, cols, k_row, k_col, wp, pad;
};

// One warp per matrix line: lines 0..rows-1 are rows (k_row smallest kept), lines
// rows..rows+cols-1 are columns (k_col smallest kept, strided reads).
template <typename T>
__global__ void __launch_bounds__(128) k_dense_seps(const T* __restrict__ csm, const DenseDesc* __restrict__ descs,
                                                    const int64_t* __restrict__ line_start, int n_mats,
                                                    int64_t total_lines, double* __restrict__ seps) {
    const int64_t gw = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (gw >= total_lines) return;
    // binary search for the matrix owning this line
    int lo = 0, hi = n_mats - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (line_start[mid] <= gw)
            lo = mid;
        else
            hi = mid - 1;
    }
    const DenseDesc d = descs[lo];
    const int line = (int)(gw - line_start[lo]);
    const double NANV = __longlong_as_double(0x7ff8000000000000LL);
    double sep;
    if (line < d.rows)
        sep = warp_select_sep(csm + d.csm_off + (int64_t)line * d.cols, d.cols, (int64_t)1, d.k_row, NANV, lane);
    else
        sep = warp_select_sep(csm + d.csm_off + (line - d.rows), d.rows, (int64_t)d.cols, d.k_col, NANV, lane);
    if (lane == 0) seps[d.sep_off + line] = sep;
}

// One thread per packed output word.
template <typename T>
__global__ void __launch_bounds__(256) k_dense_bits(const T* __restrict__ csm, const DenseDesc* __restrict__ descs,
                                                    const int64_t* __restrict__ word_start, int n_mats,
                                                    int64_t total_words, const double* __restrict__ seps,
                                                    uint32_t* __restrict__ crp, int mutual) {
    const int64_t gi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (gi >= total_words) return;
    int lo = 0, hi = n_mats - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (word_start[mid] <= gi)
            lo = mid;
        else
            hi = mid - 1;
    }
    const DenseDesc d = descs[lo];
    const int64_t wi = gi - word_start[lo];
    const int i = (int)(wi / d.wp), w = (int)(wi - (int64_t)i * d.wp);
    const double rs = seps[d.sep_off + i];
    uint32_t out = 0;
    for (int t = 0; t < 32; ++t) {
        const int j = w * 32 + t;
        if (j < d.cols) {
            const double v = (double)csm[d.csm_off + (int64_t)i * d.cols + j];
            bool on = v <= rs;
            if (mutual) on = on && (v <= seps[d.sep_off + d.rows + j]);
            out |= on ? (1u << t) : 0u;
        }
    }
    crp[d.crp_off + wi] = out;
}

__global__ void __launch_bounds__(256) k_f32_to_f64(const float* __restrict__ in, double* __restrict__ out, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (double)in[i];
}

}  // namespace acoss
