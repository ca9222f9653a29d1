// kernels_dp16.cuh -- Qmax and Dmax (SequenceAlignment.c:113-180) on packed 16-bit lanes.
//
// Same traversal as k_dp (kernels_dp.cuh: one warp per pair, lane = CPL consecutive columns, rows in groups of three
// with a static slot rotation, cells right to left so that a row overwrites row i-3 in place), but a cell's Qmax and
// Dmax values travel together in ONE 32-bit register, X = (Dmax << 16) | Qmax in half-units, and the two recurrences
// run on the DPX instructions (VIADDMNMX.S16x2[.RELU]), which issue at the rate of their 32-bit forms
// (profiles/exp/ubench/dpx_rate_b200.txt): five instructions per cell for both tables,
//
//     m = max(P1,           P2 + (0, e))                 P1 = X[i-1][j-1], P2 = X[i-2][j-1]
//     m = max(m,            P3 + (0, 2 S[i][j-1]))       P3 = X[i-1][j-2]
//     m = max(m,            P4 + (NEG, f))               P4 = X[i-3][j-1]     (Dmax only)
//     m = max(m,            P5 + (NEG, 2 S[i][j-2] + 2 S[i][j-1]))   P5 = X[i-1][j-3]   (Dmax only)
//     X = max(m + (delta, delta), 0)
//
// where e = 2 S[i-1][j], f = 2 (S[i-2][j] + S[i-1][j]), delta = +2 on a match and -1 on a gap, and NEG = -16384 in
// the Qmax lane keeps the two Dmax-only predecessors out of the Qmax maximum.  Values: Qmax <= 2 min(rows, cols),
// Dmax <= 1.5 (rows + cols) half-units (every step of a path adds at most 6 while advancing i + j by 4), so plots
// with rows + cols <= kDp16MaxSide keep every lane inside (-32768, 16384): no wrap-around, results identical to the
// int32 kernel.  Larger plots and calls that want Smith-Waterman stay on k_dp.
#pragma once
#include "kernels_dp.cuh"

namespace acoss {

constexpr int kDp16MaxSide = 10000;            // rows + cols of a plot the packed kernel accepts
constexpr int kDp16MaxCols = 1536;             // wider plots (dozens of column blocks, a hand-over record per row and block) are
                                               // faster on k_dp: 75 vs 90 ms per step of the long-song stress set
constexpr unsigned kDp16Neg = 0x0000C000u;     // (-16384, 0): Qmax lane of the Dmax-only addends

#ifdef ACOSS_EMU
__device__ __forceinline__ unsigned dp16_addmax(unsigned a, unsigned b, unsigned c) {
    const int lo = imax((int)(short)(a & 0xffffu) + (int)(short)(b & 0xffffu), (int)(short)(c & 0xffffu));
    const int hi = imax((int)(short)(a >> 16) + (int)(short)(b >> 16), (int)(short)(c >> 16));
    return ((unsigned)hi << 16) | ((unsigned)lo & 0xffffu);
}
__device__ __forceinline__ unsigned dp16_addrelu(unsigned a, unsigned b) {
    const int lo = imax((int)(short)(a & 0xffffu) + (int)(short)(b & 0xffffu), 0);
    const int hi = imax((int)(short)(a >> 16) + (int)(short)(b >> 16), 0);
    return ((unsigned)hi << 16) | ((unsigned)lo & 0xffffu);
}
__device__ __forceinline__ unsigned dp16_max(unsigned a, unsigned b) {
    const int lo = imax((int)(short)(a & 0xffffu), (int)(short)(b & 0xffffu));
    const int hi = imax((int)(short)(a >> 16), (int)(short)(b >> 16));
    return ((unsigned)hi << 16) | ((unsigned)lo & 0xffffu);
}
#else
__device__ __forceinline__ unsigned dp16_addmax(unsigned a, unsigned b, unsigned c) { return __viaddmax_s16x2(a, b, c); }
__device__ __forceinline__ unsigned dp16_addrelu(unsigned a, unsigned b) { return __viaddmax_s16x2_relu(a, b, 0u); }
__device__ __forceinline__ unsigned dp16_max(unsigned a, unsigned b) { return __vmaxs2(a, b); }
#endif

// One row.  X[S1], X[S2], X[S3] hold rows i-1, i-2, i-3; the new row replaces X[S3].  E[c] = (0, 2 S[i-1][j]),
// F[c] = (NEG, 2 (S[i-2][j] + S[i-1][j])).  L1..L3: X[i-1] of the three columns left of this lane's first; C1, C2:
// X[i-2], X[i-3] of the column left of it.  STEADY: i >= 3.
template <int CPL, int S1, int S2, int S3, bool STEADY>
__device__ __forceinline__ void dp16_row(int i, unsigned s0, unsigned (&X)[3][CPL], unsigned (&E)[CPL], unsigned (&F)[CPL],
                                         unsigned L1, unsigned L2, unsigned L3, unsigned C1, unsigned C2, bool stale,
                                         bool first_cols, unsigned& best) {
    // H[t] = (0, 2 S[i][j0 - 2 + t])
    unsigned H[CPL + 2];
#pragma unroll
    for (int t = 0; t < CPL + 2; ++t) H[t] = (s0 & (1u << t)) ? 0x20000u : 0u;
#pragma unroll
    for (int c = CPL - 1; c >= 0; --c) {
        const unsigned Hc = H[c + 2];
        const unsigned delta = (s0 & (1u << (c + 2))) ? 0x00020002u : 0xffffffffu;
        const unsigned P1 = (c >= 1) ? X[S1][c >= 1 ? c - 1 : 0] : L1;
        const unsigned P2 = (c >= 1) ? X[S2][c >= 1 ? c - 1 : 0] : C1;
        const unsigned P3 = (c >= 2) ? X[S1][c >= 2 ? c - 2 : 0] : (c == 1 ? L1 : L2);
        const unsigned P4 = (c >= 1) ? X[S3][c >= 1 ? c - 1 : 0] : C2;
        const unsigned P5 = (c >= 3) ? X[S1][c >= 3 ? c - 3 : 0] : (c == 2 ? L1 : (c == 1 ? L2 : L3));
        unsigned m = dp16_addmax(P2, E[c], P1);
        m = dp16_addmax(P3, H[c + 1], m);
        m = dp16_addmax(P4, F[c], m);
        m = dp16_addmax(P5, H[c] + H[c + 1] + kDp16Neg, m);
        unsigned x = dp16_addrelu(m, delta);
        if (!STEADY) {
            // rows 0, 1: both tables zero; row 2: Qmax as computed, the Dmax table still holds what qmax left there
            if (i < 2) x = 0u;
            else if (i == 2) x = (x & 0xffffu) | (stale ? Hc : 0u);
        }
        X[S3][c] = x;
        F[c] = E[c] + Hc + kDp16Neg;
        E[c] = Hc;
    }
    // column boundary of the table: columns 0, 1 are never written; column 2 of the Dmax table holds what qmax left
    // there (2 S[i][2]) in the Serra09 call pattern
    if (first_cols) {
        X[S3][0] = X[S3][1] = 0u;
        X[S3][2] = (X[S3][2] & 0xffffu) | ((stale && (STEADY || i >= 2)) ? H[4] : 0u);
    }
    // Running maxima.  Steady rows: over all columns of the lane, the zero-padded ones included (kernels_dp.cuh,
    // dp_row: a padded cell is strictly below some real cell); the table's border columns 0..2 carry no Dmax value.
    // Rows 0..2: Qmax only.
    if (STEADY && !first_cols) {
#pragma unroll
        for (int c = 0; c < CPL; ++c) best = dp16_max(best, X[S3][c]);
    } else {
#pragma unroll
        for (int c = 0; c < CPL; ++c)
            best = dp16_max(best, (STEADY && c >= 3) ? X[S3][c] : (X[S3][c] & 0xffffu));
    }
}

template <int CPL>
__global__ void __launch_bounds__(128, (CPL >= 16 ? 5 : 6)) k_dp16(DpArgs A) {
    static_assert(CPL >= 4, "the border fix-up touches columns 0..2 of a lane");
    const int gw = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const int lane = threadIdx.x & 31;
    if (gw >= A.n_items) return;  // whole warps leave together
    const int pidx = (int)A.items[gw];
    const PairWork pw = A.pairs[pidx];
    const int rows = pw.crp_rows, cols = pw.crp_cols, wp = pw.crp_wp;
    const uint32_t* bits = A.crp + pw.crp_off;
    const bool stale = (A.flags & kFlagDmaxClean) == 0;

    unsigned best = 0u;
    const int nblk = (cols + 32 * CPL - 1) / (32 * CPL);
    for (int blk = 0; blk < nblk && rows >= 3 && cols >= 3; ++blk) {
        const int j0 = blk * 32 * CPL + lane * CPL;
        // bits of the field that map to real columns (field bit t <-> column j0-2+t)
        const int vb = cols - (j0 - 2);
        unsigned fmask =
            ((1u << (CPL + 2)) - 1u) & (vb >= 32 ? 0xffffffffu : (vb <= 0 ? 0u : ((1u << vb) - 1u)));
#ifndef ACOSS_EMU
        asm volatile("mov.b32 %0, %0;" : "+r"(fmask));  // loop invariant: keep it in a register (see k_dp)
#endif
        int4* bnd_out = (pw.bnd_off >= 0) ? A.bnd + pw.bnd_off + (int64_t)(blk & 1) * rows * 2 : nullptr;
        const int4* bnd_in = (pw.bnd_off >= 0) ? A.bnd + pw.bnd_off + (int64_t)((blk & 1) ^ 1) * rows * 2 : nullptr;
        unsigned X[3][CPL], E[CPL], F[CPL];
#pragma unroll
        for (int c = 0; c < CPL; ++c) {
            X[0][c] = X[1][c] = X[2][c] = 0u;
            E[c] = 0u;
            F[c] = kDp16Neg;
        }
        unsigned c1 = 0u, c2 = 0u;  // X[i-2], X[i-3] of the column left of the lane
        const bool first_cols = (j0 == 0);
        const bool write_bnd = bnd_out != nullptr && lane == 31 && blk + 1 < nblk;

#define ACOSS_DP16_ROW(S1, S2, S3, U, ST)                                                                     \
    {                                                                                                         \
        const unsigned s0 = __funnelshift_r(plo[U], phi[U], fsh) & fmask;                                     \
        unsigned L1 = __shfl_up_sync(kFullMask, X[S1][CPL - 1], 1);                                           \
        unsigned L2 = __shfl_up_sync(kFullMask, X[S1][CPL - 2], 1);                                           \
        unsigned L3 = __shfl_up_sync(kFullMask, X[S1][CPL - 3], 1);                                           \
        if (lane == 0) {                                                                                      \
            L1 = L2 = L3 = 0u;                                                                                \
            if (blk > 0 && (ST || i > 0)) {                                                                   \
                const int4 ra = pra[U];                                                                       \
                L1 = (unsigned)ra.x; L2 = (unsigned)ra.y; L3 = (unsigned)ra.z;                                \
            }                                                                                                 \
        }                                                                                                     \
        const unsigned C1 = c1, C2 = c2;                                                                      \
        c2 = c1; c1 = L1;                                                                                     \
        dp16_row<CPL, S1, S2, S3, ST>(i, s0, X, E, F, L1, L2, L3, C1, C2, stale, first_cols, best);           \
        if (write_bnd)                                                                                        \
            bnd_out[(int64_t)i * 2] = make_int4((int)X[S3][CPL - 1], (int)X[S3][CPL - 2], (int)X[S3][CPL - 3], 0); \
    }

        const int w0 = (j0 - 2) >> 5;
        unsigned fsh = (unsigned)(j0 - 2) & 31u;
#ifndef ACOSS_EMU
        asm volatile("mov.b32 %0, %0;" : "+r"(fsh));
#endif
        const bool okA = (w0 >= 0 && w0 < wp), okB = (w0 + 1 >= 0 && w0 + 1 < wp);
        unsigned plo[3], phi[3], nlo[3], nhi[3];
        int4 pra[3], nra[3];  // hand-over records of rows i-1 (lane 0 of blocks > 0)
        const bool rd_bnd = (lane == 0 && blk > 0);
#pragma unroll
        for (int u = 0; u < 3; ++u) {
            const uint32_t* rp = bits + (int64_t)u * wp + w0;
            plo[u] = (okA && u < rows) ? __ldg(rp) : 0u;
            phi[u] = (okB && u < rows) ? __ldg(rp + 1) : 0u;
            pra[u] = make_int4(0, 0, 0, 0);
            if (rd_bnd && u >= 1 && u - 1 < rows) pra[u] = bnd_in[(int64_t)(u - 1) * 2];
        }
        for (int i0 = 0; i0 < rows; i0 += 3) {
            // the packed bits of the next group are requested before this group is computed
#pragma unroll
            for (int u = 0; u < 3; ++u) {
                const int in = i0 + 3 + u;
                const uint32_t* rp = bits + (int64_t)in * wp + w0;
                nlo[u] = (okA && in < rows) ? __ldg(rp) : 0u;
                nhi[u] = (okB && in < rows) ? __ldg(rp + 1) : 0u;
                nra[u] = make_int4(0, 0, 0, 0);
                if (rd_bnd && in < rows) nra[u] = bnd_in[(int64_t)(in - 1) * 2];
            }
            int i = i0;
            if (i0 >= 3) {
                ACOSS_DP16_ROW(2, 1, 0, 0, true)
                i = i0 + 1;
                if (i < rows) ACOSS_DP16_ROW(0, 2, 1, 1, true)
                i = i0 + 2;
                if (i < rows) ACOSS_DP16_ROW(1, 0, 2, 2, true)
            } else {
                ACOSS_DP16_ROW(2, 1, 0, 0, false)
                i = i0 + 1;
                if (i < rows) ACOSS_DP16_ROW(0, 2, 1, 1, false)
                i = i0 + 2;
                if (i < rows) ACOSS_DP16_ROW(1, 0, 2, 2, false)
            }
#pragma unroll
            for (int u = 0; u < 3; ++u) {
                plo[u] = nlo[u];
                phi[u] = nhi[u];
                pra[u] = nra[u];
            }
        }
#undef ACOSS_DP16_ROW
        __syncwarp();
    }
    const int qbest = __reduce_max_sync(kFullMask, (int)(best & 0xffffu));
    const int dbest = __reduce_max_sync(kFullMask, (int)(best >> 16));
    if (lane == 0) {
        if (A.qmax_out) A.qmax_out[pidx] = (rows >= 3 && cols >= 3) ? 0.5f * (float)qbest : 0.0f;
        if (A.dmax_out) A.dmax_out[pidx] = (rows >= 4 && cols >= 4) ? 0.5f * (float)dbest : 0.0f;
    }
}

}  // namespace acoss
