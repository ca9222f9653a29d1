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

// One DP row for all three recurrences.  S1, S2, S3 are the register slots that hold rows i-1,
// i-2, i-3; the new row overwrites slot S3 in place: cells are visited right to left, so a cell
// reads its left neighbours of row i-3 before they are replaced (no register copies per row).
// STEADY: the row index is known to be >= 3 (every row after the first group of three): no per-row predicates.
template <int CPL, bool WITH_SW, int S1, int S2, int S3, bool STEADY>
__device__ __forceinline__ void dp_row(int i, unsigned s0, unsigned sp1, unsigned sp2, int (&q)[3][CPL], int (&d)[3][CPL],
                                       float (&h)[3][CPL], int (&e)[CPL], int (&f)[CPL], int qL1, int qL2, int q2L1,
                                       int dL1, int dL2, int dL3, int d2L1, int d3L1, float hL1, float hL2, float h2L1,
                                       bool stale, bool first_cols, int nvalid, int j0, int& qbest, int& dbest,
                                       float& hbest) {
    // sv[t] = 2 * S[i][j0 - 2 + t]  (half-units of one matched cell)
    int sv[CPL + 2];
#pragma unroll
    for (int t = 0; t < CPL + 2; ++t) sv[t] = (int)((s0 >> t) & 1u) * 2;
#pragma unroll
    for (int c = CPL - 1; c >= 0; --c) {
        const int s2x = sv[c + 2];
        const int delta = s2x + (s2x >> 1) - 1;  // +2 on a match, -1 on a gap
        // ---- Qmax (SequenceAlignment.c:121-141)
        if (STEADY || i >= 2) {
            const int p1 = (c >= 1) ? q[S1][c >= 1 ? c - 1 : 0] : qL1;
            const int p2 = (c >= 1) ? q[S2][c >= 1 ? c - 1 : 0] : q2L1;
            const int p3 = (c >= 2) ? q[S1][c >= 2 ? c - 2 : 0] : (c == 1 ? qL1 : qL2);
            q[S3][c] = imax(imax3(p1, p2, p3) + delta, 0);
        } else {
            q[S3][c] = 0;
        }
        // ---- Dmax (SequenceAlignment.c:155-177); e = 2 S[i-1][j], f = 2 (S[i-2][j] + S[i-1][j])
        if (STEADY || i >= 3) {
            const int a1 = (c >= 1) ? d[S1][c >= 1 ? c - 1 : 0] : dL1;
            const int a2 = ((c >= 1) ? d[S2][c >= 1 ? c - 1 : 0] : d2L1) + e[c];
            const int a3 = ((c >= 2) ? d[S1][c >= 2 ? c - 2 : 0] : (c == 1 ? dL1 : dL2)) + sv[c + 1];
            const int a4 = ((c >= 1) ? d[S3][c >= 1 ? c - 1 : 0] : d3L1) + f[c];
            const int a5 = ((c >= 3) ? d[S1][c >= 3 ? c - 3 : 0] : (c == 2 ? dL1 : (c == 1 ? dL2 : dL3))) + sv[c] +
                           sv[c + 1];
            d[S3][c] = imax(imax3(imax3(a1, a2, a3), a4, a5) + delta, 0);
        } else if (i == 2) {
            d[S3][c] = stale ? s2x : 0;  // row 2 of the table qmax left behind
        } else {
            d[S3][c] = 0;
        }
        f[c] = e[c] + s2x;
        e[c] = s2x;
        // ---- constrained Smith-Waterman (SequenceAlignment.c:83-97), 0-based on S
        if (WITH_SW) {
            if (STEADY || i >= 2) {
                const unsigned s = (s0 >> (c + 2)) & 1u;
                const float ms = s ? 1.0f : -1.0f;
                const unsigned b11 = (sp1 >> (c + 1)) & 1u;  // S[i-1][j-1]
                const unsigned b21 = (sp2 >> (c + 1)) & 1u;  // S[i-2][j-1]
                const unsigned b12 = (sp1 >> c) & 1u;        // S[i-1][j-2]
                const float g1 = (c >= 1) ? h[S1][c >= 1 ? c - 1 : 0] : hL1;
                const float g2 = (c >= 1) ? h[S2][c >= 1 ? c - 1 : 0] : h2L1;
                const float g3 = (c >= 2) ? h[S1][c >= 2 ? c - 2 : 0] : (c == 1 ? hL1 : hL2);
                h[S3][c] = fmaxf(fmaxf(sw_step(g1, ms, s, b11), sw_step(g2, ms, s, b21)),
                                 fmaxf(sw_step(g3, ms, s, b12), 0.0f));
            } else {
                h[S3][c] = 0.0f;
            }
        }
    }
    // column boundary of the table: columns 0,1 are never written; column 2 of the Dmax table
    // holds what qmax left there (S[i][2]) in the Serra09 call pattern
    if (first_cols) {
        q[S3][0] = q[S3][1] = 0;
        d[S3][0] = d[S3][1] = 0;
        d[S3][2] = (stale && (STEADY || i >= 2)) ? sv[4] : 0;
        if (WITH_SW) h[S3][0] = h[S3][1] = 0.0f;
    }
    // Steady rows: the maxima may run over the columns beyond the plot as well.  Those cells have S = 0 (fmask),
    // so a Qmax cell there is max(0, max(predecessors) - 1) and a Dmax cell max(0, max5(...) - 1) where every
    // bonus term comes with a valid matched cell of the same row that scores at least one half-unit more: by
    // induction each such value is strictly below some real cell's, and the maximum is unchanged.
    if (STEADY && !WITH_SW) {
#pragma unroll
        for (int c = 0; c < CPL; ++c) qbest = imax(qbest, q[S3][c]);
        if (!first_cols) {
#pragma unroll
            for (int c = 0; c < CPL; ++c) dbest = imax(dbest, d[S3][c]);
        } else {
#pragma unroll
            for (int c = 3; c < CPL; ++c) dbest = imax(dbest, d[S3][c]);  // columns 0..2 are the table's border
        }
        return;
    }
    // running maxima over real cells only
    if (STEADY || i >= 2) {
        if (nvalid == CPL) {
#pragma unroll
            for (int c = 0; c < CPL; ++c) {
                qbest = imax(qbest, q[S3][c]);
                if (WITH_SW) hbest = fmaxf(hbest, h[S3][c]);
            }
        } else {
#pragma unroll
            for (int c = 0; c < CPL; ++c)
                if (c < nvalid) {
                    qbest = imax(qbest, q[S3][c]);
                    if (WITH_SW) hbest = fmaxf(hbest, h[S3][c]);
                }
        }
    }
    if (STEADY || i >= 3) {
        if (nvalid == CPL && !first_cols) {
#pragma unroll
            for (int c = 0; c < CPL; ++c) dbest = imax(dbest, d[S3][c]);
        } else {
#pragma unroll
            for (int c = 0; c < CPL; ++c)
                if (c < nvalid && (j0 + c) >= 3) dbest = imax(dbest, d[S3][c]);
        }
    }
}

template <int CPL, bool WITH_SW>
__global__ void __launch_bounds__(128, (CPL >= 16 ? 4 : 5)) k_dp(DpArgs A) {
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
        unsigned fmask =
            ((1u << (CPL + 2)) - 1u) & (vb >= 32 ? 0xffffffffu : (vb <= 0 ? 0u : ((1u << vb) - 1u)));
#ifndef ACOSS_EMU
        // loop invariant, but ptxas rematerialises its 7 instructions in every row: pin it in a register
        asm volatile("mov.b32 %0, %0;" : "+r"(fmask));
#endif
        int4* bnd_out = (pw.bnd_off >= 0) ? A.bnd + pw.bnd_off + (int64_t)(blk & 1) * rows * 2 : nullptr;
        const int4* bnd_in = (pw.bnd_off >= 0) ? A.bnd + pw.bnd_off + (int64_t)((blk & 1) ^ 1) * rows * 2 : nullptr;
        int q[3][CPL], d[3][CPL], e[CPL], f[CPL];
        float h[3][CPL];
#pragma unroll
        for (int c = 0; c < CPL; ++c) {
            q[0][c] = q[1][c] = q[2][c] = 0;
            d[0][c] = d[1][c] = d[2][c] = 0;
            h[0][c] = h[1][c] = h[2][c] = 0.0f;
            e[c] = f[c] = 0;
        }
        unsigned sp1 = 0, sp2 = 0;  // bit fields of rows i-1, i-2 (bit c+2 <-> column j0+c)
        int cq1 = 0, cd1 = 0, cd2 = 0;
        float ch1 = 0.0f;
        const bool first_cols = (j0 == 0);
        const bool write_bnd = bnd_out != nullptr && lane == 31 && blk + 1 < nblk;

#define ACOSS_DP_ROW(S1, S2, S3, U, ST)                                                                          \
    {                                                                                                         \
        const unsigned s0 = __funnelshift_r(plo[U], phi[U], fsh) & fmask;                                     \
        int qL1 = __shfl_up_sync(kFullMask, q[S1][CPL - 1], 1);                                               \
        int qL2 = __shfl_up_sync(kFullMask, q[S1][CPL - 2], 1);                                               \
        int dL1 = __shfl_up_sync(kFullMask, d[S1][CPL - 1], 1);                                               \
        int dL2 = __shfl_up_sync(kFullMask, d[S1][CPL - 2], 1);                                               \
        int dL3 = __shfl_up_sync(kFullMask, d[S1][CPL - 3], 1);                                               \
        float hL1 = 0.0f, hL2 = 0.0f;                                                                         \
        if (WITH_SW) {                                                                                        \
            hL1 = __shfl_up_sync(kFullMask, h[S1][CPL - 1], 1);                                               \
            hL2 = __shfl_up_sync(kFullMask, h[S1][CPL - 2], 1);                                               \
        }                                                                                                     \
        if (lane == 0) {                                                                                      \
            qL1 = qL2 = dL1 = dL2 = dL3 = 0;                                                                  \
            hL1 = hL2 = 0.0f;                                                                                 \
            if (blk > 0 && (ST || i > 0)) {                                                                   \
                const int4 ra = pra[U];                                                                       \
                qL1 = ra.x & 0xffff; qL2 = (int)((unsigned)ra.x >> 16); dL1 = ra.y; dL2 = ra.z; dL3 = ra.w;   \
                if (WITH_SW) {                                                                                \
                    const int4 rb = prb[U];                                                                   \
                    hL1 = __int_as_float(rb.x); hL2 = __int_as_float(rb.y);                                   \
                }                                                                                             \
            }                                                                                                 \
        }                                                                                                     \
        const int q2L1 = cq1, d2L1 = cd1, d3L1 = cd2;                                                         \
        const float h2L1 = ch1;                                                                               \
        cq1 = qL1; cd2 = cd1; cd1 = dL1; ch1 = hL1;                                                           \
        dp_row<CPL, WITH_SW, S1, S2, S3, ST>(i, s0, sp1, sp2, q, d, h, e, f, qL1, qL2, q2L1, dL1, dL2, dL3, d2L1, \
                                         d3L1, hL1, hL2, h2L1, stale, first_cols, nvalid, j0, qbest, dbest,   \
                                         hbest);                                                              \
        if (write_bnd) {                                                                                      \
            /* one record per row: Qmax values are < 2^16 half-units (plots of up to 32767 rows or columns) */ \
            bnd_out[(int64_t)i * 2] = make_int4(q[S3][CPL - 1] | (q[S3][CPL - 2] << 16), d[S3][CPL - 1],      \
                                                d[S3][CPL - 2], d[S3][CPL - 3]);                              \
            if (WITH_SW)                                                                                      \
                bnd_out[(int64_t)i * 2 + 1] = make_int4(__float_as_int(h[S3][CPL - 1]),                       \
                                                        __float_as_int(h[S3][CPL - 2]), 0, 0);                \
        }                                                                                                     \
        sp2 = sp1;                                                                                            \
        sp1 = s0;                                                                                             \
    }

        // rows in groups of three so that the slot rotation is static: the row written in
        // position u of a group lands in slot u (previous rows: slots (u+2)%3 and (u+1)%3).
        // The packed bits of the NEXT group are requested before this group is computed, so the
        // L2 / HBM latency of the CRP read is off the critical path.
        const int w0 = (j0 - 2) >> 5;
        const unsigned fsh = (unsigned)(j0 - 2) & 31u;
        const bool okA = (w0 >= 0 && w0 < wp), okB = (w0 + 1 >= 0 && w0 + 1 < wp);
        unsigned plo[3], phi[3], nlo[3], nhi[3];
        int4 pra[3], prb[3], nra[3], nrb[3];  // hand-over records of rows i-1 (lane 0 of blocks > 0)
        const bool rd_bnd = (lane == 0 && blk > 0);
#pragma unroll
        for (int u = 0; u < 3; ++u) {
            const uint32_t* rp = bits + (int64_t)u * wp + w0;
            plo[u] = (okA && u < rows) ? __ldg(rp) : 0u;
            phi[u] = (okB && u < rows) ? __ldg(rp + 1) : 0u;
            pra[u] = prb[u] = make_int4(0, 0, 0, 0);
            if (rd_bnd && u >= 1 && u - 1 < rows) {
                pra[u] = bnd_in[(int64_t)(u - 1) * 2];
                if (WITH_SW) prb[u] = bnd_in[(int64_t)(u - 1) * 2 + 1];
            }
        }
        for (int i0 = 0; i0 < rows; i0 += 3) {
#pragma unroll
            for (int u = 0; u < 3; ++u) {
                const int in = i0 + 3 + u;
                const uint32_t* rp = bits + (int64_t)in * wp + w0;
                nlo[u] = (okA && in < rows) ? __ldg(rp) : 0u;
                nhi[u] = (okB && in < rows) ? __ldg(rp + 1) : 0u;
                nra[u] = nrb[u] = make_int4(0, 0, 0, 0);
                if (rd_bnd && in < rows) {
                    nra[u] = bnd_in[(int64_t)(in - 1) * 2];
                    if (WITH_SW) nrb[u] = bnd_in[(int64_t)(in - 1) * 2 + 1];
                }
            }
            int i = i0;
            if (i0 >= 3) {
                ACOSS_DP_ROW(2, 1, 0, 0, true)
                i = i0 + 1;
                if (i < rows) ACOSS_DP_ROW(0, 2, 1, 1, true)
                i = i0 + 2;
                if (i < rows) ACOSS_DP_ROW(1, 0, 2, 2, true)
            } else {
                ACOSS_DP_ROW(2, 1, 0, 0, false)
                i = i0 + 1;
                if (i < rows) ACOSS_DP_ROW(0, 2, 1, 1, false)
                i = i0 + 2;
                if (i < rows) ACOSS_DP_ROW(1, 0, 2, 2, false)
            }
#pragma unroll
            for (int u = 0; u < 3; ++u) {
                plo[u] = nlo[u];
                phi[u] = nhi[u];
                pra[u] = nra[u];
                prb[u] = nrb[u];
            }
        }
#undef ACOSS_DP_ROW
        __syncwarp();
    }
    qbest = __reduce_max_sync(kFullMask, qbest);
    dbest = __reduce_max_sync(kFullMask, dbest);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) hbest = fmaxf(hbest, __shfl_xor_sync(kFullMask, hbest, o));
    if (lane == 0) {
        if (A.qmax_out) A.qmax_out[pidx] = (rows >= 3 && cols >= 3) ? 0.5f * (float)qbest : 0.0f;
        if (A.dmax_out) A.dmax_out[pidx] = (rows >= 4 && cols >= 4) ? 0.5f * (float)dbest : 0.0f;
        if (WITH_SW && A.sw_out) A.sw_out[pidx] = (rows >= 3 && cols >= 3) ? hbest : 0.0f;
    }
}

}  // namespace acoss
