// kernels_eval.cuh -- the step just after the scoring path:
// the rank computation of CoverAlgorithm.getEvalStatistics (benchmarking/CoverAlgorithm.py:330-418).
//
// The reference permutes the N x N score matrix so that cliques are contiguous (:345-357), puts
// -inf on the diagonal (:361), sorts every row in descending order (np.argsort(-D, 1), :362) and
// then walks each sorted row looking for the members of the row's own clique (:367-382).  The
// only thing that loop extracts is, per row, the 1-based positions of the clique's other songs
// in the sorted row ("iranks").  The position of member c in the sorted row i is
//      1 + #{ j != i : D[i, j] > D[i, c]  or  (D[i, j] == D[i, c] and pos(j) < pos(c)) }
// (pos = index in the clique-contiguous order; equal scores keep that order, i.e. a stable sort --
// the reference's introsort leaves the order of equal scores undefined).  So no sort is needed:
// one streaming pass over row i with one counter per clique member.  The matrix is read once, in
// place (no permuted copies): N^2 * 4 bytes, HBM-bound.
#pragma once
#include "common.cuh"

namespace acoss {

constexpr int kEvalThreads = 256;
constexpr int kEvalMembers = 16;  // most clique members counted per pass over the row (k_eval_ranks<NM>: NM <= 16)

struct EvalArgs {
    const float* D;            // [n][n] row-major scores, as CoverAlgorithm.Ds[type] (symmetrised, diagonal ignored)
    const int32_t* order;      // [n] clique-contiguous order (CoverAlgorithm.py:355): position p holds song order[p]
    const int32_t* pos;        // [n] inverse of order
    const int32_t* cstart;     // [rows] position of the first member of the clique of position p
    const int32_t* csize;      // [rows] size K of that clique (>= 2)
    const int64_t* out_off;    // [rows] offset of the row's K - 1 ranks in raw / out
    int32_t* raw;              // ranks in member order (scratch)
    int32_t* out;              // ranks ascending = the reference's iranks[0:-1]
    int64_t n;
    int rows;
};

// One CTA per position p < rows (cliques are sorted by size, so the rows of cliques with K >= 2
// are the first `rows` positions).  NM = members counted per pass: the launch picks the smallest
// instantiation that covers the largest clique (NM = 1 for sets of pairs such as covers80, 12 for
// the benchmark subset's cliques of 13), because the inner loop costs 3 instructions per member
// and element and decides whether the kernel is HBM- or issue-bound.
template <int NM>
__global__ void __launch_bounds__(kEvalThreads) k_eval_ranks(EvalArgs A) {
    __shared__ float s_val[NM];
    __shared__ int s_pos[NM];
    __shared__ int s_cnt[NM];
    const int p = (int)blockIdx.x;
    if (p >= A.rows) return;
    const int tid = (int)threadIdx.x, lane = tid & 31;
    const int i = A.order[p];
    const int start = A.cstart[p], K = A.csize[p];
    const int64_t n = A.n;
    const float* __restrict__ row = A.D + (int64_t)i * n;
    int32_t* raw = A.raw + A.out_off[p];
    int32_t* out = A.out + A.out_off[p];
    const int nm_all = K - 1;  // members other than the row's own song
    for (int m0 = 0; m0 < nm_all; m0 += NM) {
        const int nm = imin(NM, nm_all - m0);
        __syncthreads();
        if (tid < NM) {
            float v = INFINITY;  // unused slots: nothing is greater than +inf
            int q = -1;
            if (tid < nm) {
                const int t = m0 + tid;
                q = start + t + ((start + t >= p) ? 1 : 0);  // skip the row's own position
                v = row[A.order[q]];
            }
            s_val[tid] = v;
            s_pos[tid] = q;
            s_cnt[tid] = 0;
        }
        __syncthreads();
        float mv[NM];
        int cnt[NM];
        float mmin = INFINITY;
#pragma unroll
        for (int t = 0; t < NM; ++t) {
            mv[t] = s_val[t];
            cnt[t] = 0;
            mmin = fminf(mmin, mv[t]);
        }
        // four elements in flight per thread; a group whose largest score is below every member's adds nothing
        for (int64_t j0 = (int64_t)tid; j0 < n; j0 += 4 * kEvalThreads) {
            float v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int64_t j = j0 + (int64_t)u * kEvalThreads;
                v[u] = (j < n && j != i) ? __ldg(row + j) : -INFINITY;  // the diagonal is -inf (CoverAlgorithm.py:361)
            }
            const float vmax = fmaxf(fmaxf(v[0], v[1]), fmaxf(v[2], v[3]));
            if (!(vmax >= mmin)) continue;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                bool eq = false;
#pragma unroll
                for (int t = 0; t < NM; ++t) {
                    cnt[t] += (v[u] > mv[t]) ? 1 : 0;
                    eq = eq || (v[u] == mv[t]);
                }
                if (eq) {
                    // equal scores: the clique-contiguous order decides (stable sort)
                    const int64_t j = j0 + (int64_t)u * kEvalThreads;
                    if (j < n && j != i) {
                        const int pj = A.pos[j];
#pragma unroll
                        for (int t = 0; t < NM; ++t)
                            cnt[t] += (v[u] == mv[t] && pj < s_pos[t]) ? 1 : 0;
                    }
                }
            }
        }
#pragma unroll
        for (int t = 0; t < NM; ++t) {
            const int s = warp_sum(cnt[t]);
            if (lane == 0 && s != 0) atomicAdd(&s_cnt[t], s);
        }
        __syncthreads();
        if (tid < nm) raw[m0 + tid] = 1 + s_cnt[tid];
    }
    __syncthreads();
    // ascending order: ranks are distinct, so the slot of a rank is the number of smaller ones
    for (int t = tid; t < nm_all; t += kEvalThreads) {
        const int r = raw[t];
        int slot = 0;
        for (int u = 0; u < nm_all; ++u) slot += (raw[u] < r) ? 1 : 0;
        out[slot] = r;
    }
}

// Larger cliques: binary search instead of one compare per member.  The (<= 15) members of a pass are sorted by
// (score ascending, clique position descending) -- the order in which a row element "beats" them -- so the set of
// members an element beats is a prefix of that list; its length b is found by a 4-step binary search on the scores
// (plus a short walk over members of exactly equal score), counted in a per-thread histogram column in shared
// memory (conflict-free), and member s ends up with  sum_{b > s} hist[b]  elements in front of it.  About 21
// instructions per element for any clique size, against 3 per member for the direct kernel.
constexpr int kEvalBsMembers = 15;
constexpr int kEvalBsFlight = 8;  // row elements in flight per thread

__global__ void __launch_bounds__(kEvalThreads) k_eval_ranks_bs(EvalArgs A) {
    __shared__ float s_rawv[16], s_val[16];
    __shared__ int s_rawp[16], s_pos[16], s_tot[16];
    __shared__ int s_hist[16 * kEvalThreads];
    const int p = (int)blockIdx.x;
    if (p >= A.rows) return;
    const int tid = (int)threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int i = A.order[p];
    const int start = A.cstart[p], K = A.csize[p];
    const int64_t n = A.n;
    const float* __restrict__ row = A.D + (int64_t)i * n;
    int32_t* raw = A.raw + A.out_off[p];
    int32_t* out = A.out + A.out_off[p];
    const int nm_all = K - 1;
    for (int m0 = 0; m0 < nm_all; m0 += kEvalBsMembers) {
        const int nm = imin(kEvalBsMembers, nm_all - m0);
        __syncthreads();
        if (tid < 16) {
            float v = INFINITY;  // pads sort last and are never beaten
            int q = -1;
            if (tid < nm) {
                const int t = m0 + tid;
                q = start + t + ((start + t >= p) ? 1 : 0);
                v = row[A.order[q]];
            }
            s_rawv[tid] = v;
            s_rawp[tid] = q;
            s_tot[tid] = 0;
        }
#pragma unroll
        for (int b = 0; b < 16; ++b) s_hist[b * kEvalThreads + tid] = 0;
        __syncthreads();
        if (tid < 16) {
            const float v = s_rawv[tid];
            const int q = s_rawp[tid];
            int r = 0;
            for (int u = 0; u < 16; ++u) {
                const float vu = s_rawv[u];
                const int qu = s_rawp[u];
                r += (vu < v || (vu == v && (qu > q || (qu == q && u < tid)))) ? 1 : 0;
            }
            s_val[r] = v;
            s_pos[r] = q;
        }
        __syncthreads();
        const float mmin = s_val[0];
        for (int64_t j0 = (int64_t)tid; j0 < n; j0 += kEvalBsFlight * kEvalThreads) {
            float v[kEvalBsFlight];
#pragma unroll
            for (int u = 0; u < kEvalBsFlight; ++u) {
                const int64_t j = j0 + (int64_t)u * kEvalThreads;
                v[u] = (j < n && j != i) ? __ldg(row + j) : -INFINITY;
            }
#pragma unroll
            for (int u = 0; u < kEvalBsFlight; ++u) {
                const float x = v[u];
                if (!(x >= mmin) || x == -INFINITY) continue;  // below every member / the diagonal / past the row
                int b = (s_val[7] < x) ? 8 : 0;
                b += (s_val[b + 3] < x) ? 4 : 0;
                b += (s_val[b + 1] < x) ? 2 : 0;
                b += (s_val[b] < x) ? 1 : 0;
                if (s_val[b] == x) {
                    // equal scores: the element also beats the members that come later in the clique order
                    const int pj = A.pos[j0 + (int64_t)u * kEvalThreads];
                    while (b < 16 && s_val[b] == x && pj < s_pos[b]) ++b;
                }
                if (b) s_hist[b * kEvalThreads + tid] += 1;
            }
        }
        __syncthreads();
        // bucket totals: warp w sums buckets 2w and 2w + 1
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int b = 2 * warp + h;
            int c = 0;
#pragma unroll
            for (int e = 0; e < kEvalThreads / 32; ++e) c += s_hist[b * kEvalThreads + lane + 32 * e];
            c = warp_sum(c);
            if (lane == 0) s_tot[b] = c;
        }
        __syncthreads();
        if (tid < nm) {
            int c = 0;
            for (int b = tid + 1; b < 16; ++b) c += s_tot[b];
            raw[m0 + tid] = 1 + c;
        }
    }
    __syncthreads();
    for (int t = tid; t < nm_all; t += kEvalThreads) {
        const int r = raw[t];
        int slot = 0;
        for (int u = 0; u < nm_all; ++u) slot += (raw[u] < r) ? 1 : 0;
        out[slot] = r;
    }
}

}  // namespace acoss
