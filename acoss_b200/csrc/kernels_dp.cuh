// kernels_dp.cuh -- Qmax / Dmax / constrained Smith-Waterman over a bit-packed binary
// cross-recurrence plot (the CUDA equivalent of benchmarking/SequenceAlignment.c).
//
// One warp scores one pair.  All three recurrences read only earlier ROWS
// (SequenceAlignment.c:87-91,125-127,159-162), so a row is computed for all columns at once:
// lane l owns CPL consecutive columns and keeps the last two (Qmax, SW) / three (Dmax) rows
// of its columns in registers; the three values it needs from its left neighbour arrive by
// __shfl_up_sync.  Plots wider than 32*CPL columns are processed as consecutive column blocks,
// the last columns of a block being handed to the next through a small per-pair scratch.
//
// Qmax and Dmax values are multiples of 0.5 below 2^23 in the reference's float table, so
// they are computed exactly in int32 half-units (match +2, gap -1, bonus +2 per skipped
// one-cell); SW keeps the reference's float32 expression order (its -0.7 is not dyadic).
#pragma once
#include "common.cuh"

namespace acoss {

struct DpArgs {
    const PairWork* pairs;
    const uint32_t* items;  // pair indices handled by this launch (one per warp)
    int n_items;
    const uint32_t* crp;
    int4* bnd;              // boundary scratch, 2 records of 2 x int4 per row per wide pair
    float* qmax_out;        // indexed by pair
    float* dmax_out;
    float* sw_out;
    uint32_t flags;
};

// 32 bits of a packed row starting at bit `start` (may be negative or run past the row: zeros).
__device__ __forceinline__ uint32_t fetch_bits(const uint32_t* row, int wp, int start) {
    const int w = start >> 5;
    const unsigned sh = (unsigned)start & 31u;
    const uint32_t lo = (w >= 0 && w < wp) ? __ldg(row + w) : 0u;
    const uint32_t hi = (w + 1 >= 0 && w + 1 < wp) ? __ldg(row + w + 1) : 0u;
    return __funnelshift_r(lo, hi, sh);
}

__device__ __forceinline__ float sw_step(float h, float ms, unsigned here, unsigned before) {
    // Delta(before, here) of SequenceAlignment.c:43-54 added after the match score
    const float gap = here ? 0.0f : (before ? -0.5f : -0.7f);
    return __fadd_rn(__fadd_rn(h, ms), gap);
}

template <int CPL, bool WITH_SW>
__global__ void __launch_bounds__(128) k_dp(DpArgs A) {
    const int gw = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const int lane = threadIdx.x & 31;
    if (gw >= A.n_items) return;  // whole warps leave together
    const int pidx = (int)A.items[gw];
    const PairWork pw = A.pairs[pidx];
    const int rows = pw.crp_rows, cols = pw.crp_cols, wp = pw.crp_wp;
    const uint32_t* bits = A.crp + pw.crp_off;
    const bool stale = (A.flags & kFlagDmaxClean) == 0;

    int qbest = 0, dbest = 0;
    float hbest = 0.0f;
    const int nblk = (cols + 32 * CPL - 1) / (32 * CPL);
    for (int blk = 0; blk < nblk && rows >= 3 && cols >= 3; ++blk) {
        const int j0 = blk * 32 * CPL + lane * CPL;
        const int nvalid = imax(0, imin(CPL, cols - j0));
        // bits of the field that map to real columns (field bit t <-> column j0-2+t)
        const int vb = cols - (j0 - 2);
        const unsigned fmask = ((1u << (CPL + 2)) - 1u) & (vb >= 32 ? 0xffffffffu : (vb <= 0 ? 0u : ((1u << vb) - 1u)));
        int4* bnd_out = (pw.bnd_off >= 0) ? A.bnd + pw.bnd_off + (int64_t)(blk & 1) * rows * 2 : nullptr;
        const int4* bnd_in = (pw.bnd_off >= 0) ? A.bnd + pw.bnd_off + (int64_t)((blk & 1) ^ 1) * rows * 2 : nullptr;
        int q1[CPL], q2[CPL], d1[CPL], d2[CPL], d3[CPL];
        float h1[CPL], h2[CPL];
#pragma unroll
        for (int c = 0; c < CPL; ++c) {
            q1[c] = q2[c] = d1[c] = d2[c] = d3[c] = 0;
            h1[c] = h2[c] = 0.0f;
        }
        unsigned sp1 = 0, sp2 = 0;  // bit fields of rows i-1, i-2 (bit c+2 <-> column j0+c)
        int cq1 = 0, cd1 = 0, cd2 = 0;
        float ch1 = 0.0f;
        for (int i = 0; i < rows; ++i) {
            const unsigned s0 = fetch_bits(bits + (int64_t)i * wp, wp, j0 - 2) & fmask;
            // values of the columns just left of this lane, rows i-1 (fresh) and i-2, i-3 (carried)
            int qL1 = __shfl_up_sync(kFullMask, q1[CPL - 1], 1);
            int qL2 = __shfl_up_sync(kFullMask, q1[CPL - 2], 1);
            int dL1 = __shfl_up_sync(kFullMask, d1[CPL - 1], 1);
            int dL2 = __shfl_up_sync(kFullMask, d1[CPL - 2], 1);
            int dL3 = __shfl_up_sync(kFullMask, d1[CPL - 3], 1);
            float hL1 = 0.0f, hL2 = 0.0f;
            if (WITH_SW) {
                hL1 = __shfl_up_sync(kFullMask, h1[CPL - 1], 1);
                hL2 = __shfl_up_sync(kFullMask, h1[CPL - 2], 1);
            }
            if (lane == 0) {
                qL1 = qL2 = dL1 = dL2 = dL3 = 0;
                hL1 = hL2 = 0.0f;
                if (blk > 0 && i > 0) {
                    const int4 ra = bnd_in[(int64_t)(i - 1) * 2];
                    qL1 = ra.x;
                    qL2 = ra.y;
                    dL1 = ra.z;
                    dL2 = ra.w;
                    const int4 rb = bnd_in[(int64_t)(i - 1) * 2 + 1];
                    dL3 = rb.x;
                    hL1 = __int_as_float(rb.y);
                    hL2 = __int_as_float(rb.z);
                }
            }
            const int q2L1 = cq1, d2L1 = cd1, d3L1 = cd2;
            const float h2L1 = ch1;
            cq1 = qL1;
            cd2 = cd1;
            cd1 = dL1;
            ch1 = hL1;

            int q0[CPL], d0[CPL];
            float h0[CPL];
#pragma unroll
            for (int c = 0; c < CPL; ++c) {
                const unsigned s = (s0 >> (c + 2)) & 1u;
                // ---- Qmax (SequenceAlignment.c:121-141)
                if (i >= 2) {
                    const int p1 = (c >= 1) ? q1[c >= 1 ? c - 1 : 0] : qL1;
                    const int p2 = (c >= 1) ? q2[c >= 1 ? c - 1 : 0] : q2L1;
                    const int p3 = (c >= 2) ? q1[c >= 2 ? c - 2 : 0] : (c == 1 ? qL1 : qL2);
                    const int pq = imax3(p1, p2, p3);
                    q0[c] = s ? pq + 2 : imax(pq - 1, 0);
                } else {
                    q0[c] = 0;
                }
                // ---- Dmax (SequenceAlignment.c:155-177)
                if (i >= 3) {
                    const int s1j = (int)((sp1 >> (c + 2)) & 1u);   // S[i-1][j]
                    const int s2j = (int)((sp2 >> (c + 2)) & 1u);   // S[i-2][j]
                    const int s0a = (int)((s0 >> (c + 1)) & 1u);    // S[i][j-1]
                    const int s0b = (int)((s0 >> c) & 1u);          // S[i][j-2]
                    const int a1 = (c >= 1) ? d1[c >= 1 ? c - 1 : 0] : dL1;
                    const int a2 = ((c >= 1) ? d2[c >= 1 ? c - 1 : 0] : d2L1) + 2 * s1j;
                    const int a3 = ((c >= 2) ? d1[c >= 2 ? c - 2 : 0] : (c == 1 ? dL1 : dL2)) + 2 * s0a;
                    const int a4 = ((c >= 1) ? d3[c >= 1 ? c - 1 : 0] : d3L1) + 2 * (s2j + s1j);
                    const int a5 = ((c >= 3) ? d1[c >= 3 ? c - 3 : 0] : (c == 2 ? dL1 : (c == 1 ? dL2 : dL3))) +
                                   2 * (s0b + s0a);
                    const int pd = imax(imax3(a1, a2, a3), imax(a4, a5));
                    d0[c] = s ? pd + 2 : imax(pd - 1, 0);
                } else if (i == 2) {
                    d0[c] = stale ? 2 * (int)s : 0;  // row 2 of the table qmax left behind
                } else {
                    d0[c] = 0;
                }
                // ---- constrained Smith-Waterman (SequenceAlignment.c:83-97), 0-based on S
                if (WITH_SW) {
                    if (i >= 2) {
                        const float ms = s ? 1.0f : -1.0f;
                        const unsigned b11 = (sp1 >> (c + 1)) & 1u;  // S[i-1][j-1]
                        const unsigned b21 = (sp2 >> (c + 1)) & 1u;  // S[i-2][j-1]
                        const unsigned b12 = (sp1 >> c) & 1u;        // S[i-1][j-2]
                        const float g1 = (c >= 1) ? h1[c >= 1 ? c - 1 : 0] : hL1;
                        const float g2 = (c >= 1) ? h2[c >= 1 ? c - 1 : 0] : h2L1;
                        const float g3 = (c >= 2) ? h1[c >= 2 ? c - 2 : 0] : (c == 1 ? hL1 : hL2);
                        const float e1 = sw_step(g1, ms, s, b11);
                        const float e2 = sw_step(g2, ms, s, b21);
                        const float e3 = sw_step(g3, ms, s, b12);
                        h0[c] = fmaxf(fmaxf(e1, e2), fmaxf(e3, 0.0f));
                    } else {
                        h0[c] = 0.0f;
                    }
                } else {
                    h0[c] = 0.0f;
                }
            }
            // column boundary of the table: columns 0,1 are never written; column 2 of the
            // Dmax table holds what qmax left there (S[i][2]) in the Serra09 call pattern
            if (j0 == 0) {
                q0[0] = q0[1] = 0;
                d0[0] = d0[1] = 0;
                d0[2] = (stale && i >= 2) ? 2 * (int)((s0 >> 4) & 1u) : 0;
                h0[0] = h0[1] = 0.0f;
            }
            // running maxima over real cells only
            if (i >= 2) {
#pragma unroll
                for (int c = 0; c < CPL; ++c)
                    if (c < nvalid) {
                        qbest = imax(qbest, q0[c]);
                        if (WITH_SW) hbest = fmaxf(hbest, h0[c]);
                    }
            }
            if (i >= 3) {
#pragma unroll
                for (int c = 0; c < CPL; ++c)
                    if (c < nvalid && (j0 + c) >= 3) dbest = imax(dbest, d0[c]);
            }
            if (bnd_out != nullptr && lane == 31 && blk + 1 < nblk) {
                bnd_out[(int64_t)i * 2] = make_int4(q0[CPL - 1], q0[CPL - 2], d0[CPL - 1], d0[CPL - 2]);
                bnd_out[(int64_t)i * 2 + 1] =
                    make_int4(d0[CPL - 3], __float_as_int(h0[CPL - 1]), __float_as_int(h0[CPL - 2]), 0);
            }
#pragma unroll
            for (int c = 0; c < CPL; ++c) {
                q2[c] = q1[c];
                q1[c] = q0[c];
                d3[c] = d2[c];
                d2[c] = d1[c];
                d1[c] = d0[c];
                h2[c] = h1[c];
                h1[c] = h0[c];
            }
            sp2 = sp1;
            sp1 = s0;
        }
        __syncwarp();
    }
    // warp reduce
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        qbest = imax(qbest, __shfl_xor_sync(kFullMask, qbest, o));
        dbest = imax(dbest, __shfl_xor_sync(kFullMask, dbest, o));
        hbest = fmaxf(hbest, __shfl_xor_sync(kFullMask, hbest, o));
    }
    if (lane == 0) {
        if (A.qmax_out) A.qmax_out[pidx] = (rows >= 3 && cols >= 3) ? 0.5f * (float)qbest : 0.0f;
        if (A.dmax_out) A.dmax_out[pidx] = (rows >= 4 && cols >= 4) ? 0.5f * (float)dbest : 0.0f;
        if (WITH_SW && A.sw_out) A.sw_out[pidx] = (rows >= 3 && cols >= 3) ? hbest : 0.0f;
    }
}

}  // namespace acoss
