// kernels_sweep3.cuh -- warp-specialised cross-similarity sweep (the production sweep).
//
// ncu on the previous generation (kernels_sweep2.cuh; profiles/r01_b_*) showed the per-row
// neighbour selection taking ~80 % of the kernel's stall samples: every warp first computed R
// rows, then every warp selected one row -- a latency-bound phase (dependent count passes) in
// which the FP64 pipe idles, followed by a barrier.  Here the two jobs run CONCURRENTLY in one
// CTA:
//   * NWc compute warps (thread t owns columns 4t..4t+3, as before: 12-DFMA chain per cell from
//     the diagonal predecessor, (m+1)-deep prefix ring in shared memory) write each finished row
//     of windowed distances into one of two strip buffers (natural column order, two STS.128);
//   * NS = ceil(NWc/2) selection warps take a full strip, load each of their rows into
//     REGISTERS (they have no column frames to keep), find the row separator by counting passes
//     that are now pure register DSETP + one REDUX, and either emit the mid-point separator
//     (MODE 0) or ballot the mutual-kNN bits straight into packed CRP words (MODE 2).
// Hand-over is by named barriers (bar.arrive / bar.sync): FULL[b] compute -> selection,
// EMPTY[b] selection -> compute, plus a compute-only row barrier.  Mathematics, orientations,
// separators and CRP layout are exactly those of kernels_sweep2.cuh.
#pragma once
#include "common.cuh"
#include "kernels_sweep2.cuh"

namespace acoss {

constexpr int kBarRow = 1;    // compute warps, once per row
constexpr int kBarFull = 2;   // +b
constexpr int kBarEmpty = 4;  // +b
constexpr int kBarSel = 6;    // selection warps only
constexpr int kWsMaxThreads = 384;  // 384 x 168 registers = 63 K: one CTA always fits

__host__ __device__ inline int ws_sel_warps(int TC) {
    const int nwc = TC / 32;
    int ns = (nwc + 1) / 2;
    const int room = (kWsMaxThreads - TC) / 32;
    if (ns > room) ns = room;
    return ns < 1 ? 1 : ns;
}

__host__ __device__ inline int ws_cs_len(int TC) { return (4 * TC > 1024 ? 4 * TC : 1024) + 32; }

__host__ __device__ inline size_t ws_smem_bytes(int TC, int m) {
    const int padt = fast_padt(m);
    const int R = 2 * ws_sel_warps(TC);
    size_t d = (size_t)(m + 1) * 4 * (TC + padt)  // ring
               + (size_t)2 * R * 4 * TC           // two strip buffers
               + (size_t)kStageRows * kFastXS     // row-song staging
               + (size_t)2 * R                    // windowed row norms of the strip rows
               + (size_t)ws_cs_len(TC);           // column separators (MODE 2)
    return (d * sizeof(double) + 15) & ~(size_t)15;
}

// ---- a strip row as seen by a selection warp -------------------------------------------------
// RegRow<VPL>: the row lives in VPL registers per lane (column j = lane + 32 u), +inf padded.
template <int VPL>
struct RegRow {
    double v[VPL];
    __device__ __forceinline__ void load(const double* row, int n, int lane) {
        const double INF = __longlong_as_double(0x7ff0000000000000LL);
#pragma unroll
        for (int u = 0; u < VPL; ++u) {
            const int j = lane + 32 * u;
            v[u] = (j < n) ? row[j] : INF;
        }
    }
    __device__ __forceinline__ int count(double t) const {
        int c = 0;
#pragma unroll
        for (int u = 0; u < VPL; ++u) c += (v[u] <= t) ? 1 : 0;
        return warp_sum(c);
    }
    __device__ __forceinline__ void minmax(double& vmin, double& vmax) const {
        const double INF = __longlong_as_double(0x7ff0000000000000LL);
        vmin = INF;
        vmax = -INF;
#pragma unroll
        for (int u = 0; u < VPL; ++u) {
            vmin = fmin(vmin, v[u]);
            if (v[u] < INF) vmax = fmax(vmax, v[u]);
        }
        vmin = warp_min(vmin);
        vmax = warp_max(vmax);
    }
    __device__ __forceinline__ void neighbours(double t, double& vk, double& vk1) const {
        const double INF = __longlong_as_double(0x7ff0000000000000LL);
        vk = -INF;
        vk1 = INF;
#pragma unroll
        for (int u = 0; u < VPL; ++u) {
            if (v[u] <= t)
                vk = fmax(vk, v[u]);
            else
                vk1 = fmin(vk1, v[u]);
        }
        vk = warp_max(vk);
        vk1 = warp_min(vk1);
    }
    __device__ __forceinline__ double at(int u, const double*, int, int) const { return v[u]; }
};

// MemRow: rows too long for registers are re-read from shared memory on every pass.
struct MemRow {
    const double* row;
    int n, lane;
    __device__ __forceinline__ void load(const double* r, int n_, int lane_) {
        row = r;
        n = n_;
        lane = lane_;
    }
    __device__ __forceinline__ int count(double t) const {
        int c = 0;
        for (int j = lane; j < n; j += 32) c += (row[j] <= t) ? 1 : 0;
        return warp_sum(c);
    }
    __device__ __forceinline__ void minmax(double& vmin, double& vmax) const {
        const double INF = __longlong_as_double(0x7ff0000000000000LL);
        vmin = INF;
        vmax = -INF;
        for (int j = lane; j < n; j += 32) {
            vmin = fmin(vmin, row[j]);
            vmax = fmax(vmax, row[j]);
        }
        vmin = warp_min(vmin);
        vmax = warp_max(vmax);
    }
    __device__ __forceinline__ void neighbours(double t, double& vk, double& vk1) const {
        const double INF = __longlong_as_double(0x7ff0000000000000LL);
        vk = -INF;
        vk1 = INF;
        for (int j = lane; j < n; j += 32) {
            const double v = row[j];
            if (v <= t)
                vk = fmax(vk, v);
            else
                vk1 = fmin(vk1, v);
        }
        vk = warp_max(vk);
        vk1 = warp_min(vk1);
    }
};

// Threshold t with exactly k row values <= t (same search as strip_select in kernels_sweep2.cuh:
// previous separator + gallop, then alternating interpolation / bisection; min/max bracket as the
// fallback).  Returns false when exact ties straddle the k-th position.
template <class Row>
__device__ inline bool row_select(const Row& R, int k, double g, SelState& st, double& t_out) {
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    double lo = -INF, hi = INF;
    int clo = 0, chi = 0;
    bool have_lo = false, have_hi = false, found = false;
    double t = g;
    if (st.have && g > -INF && g < INF && st.spacing > 0.0) {
        int c = R.count(t);
        if (c == k) {
            found = true;
        } else {
            double step = st.spacing;
            for (int it = 0; it < 64 && !found; ++it) {
                if (c < k) {
                    lo = t; clo = c; have_lo = true;
                    if (have_hi) break;
                    t = fma(step, (double)(1.5f * (float)(k - c)), lo);
                } else {
                    hi = t; chi = c; have_hi = true;
                    if (have_lo) break;
                    t = fma(-step, (double)(1.5f * (float)(c - k)), hi);
                }
                step *= 2.0;
                c = R.count(t);
                if (c == k) found = true;
            }
            if (!found && have_lo && have_hi && !(lo < hi)) have_lo = have_hi = false;
        }
    }
    if (!found && !(have_lo && have_hi)) {
        double vmin, vmax;
        R.minmax(vmin, vmax);
        lo = vmin;
        clo = R.count(lo);
        hi = vmax;
        chi = R.count(hi);
        if (clo == k) {
            found = true;
            t = lo;
        } else if (clo > k) {
            t_out = -INF;
            return false;
        }
    }
    if (!found) {
        for (int iter = 0; iter < 300 && !found; ++iter) {
            if (iter & 1)
                t = 0.5 * lo + 0.5 * hi;
            else
                t = fma(hi - lo, (double)__fdividef((float)(k - clo), (float)(chi - clo)), lo);  // fp32 ratio: no DDIV
            if (!(t > lo && t < hi)) {
                t = 0.5 * lo + 0.5 * hi;
                if (!(t > lo && t < hi)) break;
            }
            const int c = R.count(t);
            if (c == k) {
                found = true;
            } else if (c < k) {
                lo = t; clo = c;
            } else {
                hi = t; chi = c;
            }
        }
        if (chi > clo && hi > lo && hi < INF && lo > -INF) st.spacing = (hi - lo) * (double)__fdividef(1.0f, (float)(chi - clo));
    }
    if (!found) t = lo;
    t_out = t;
    return found;
}

// One strip row, processed by one selection warp: separator search, then either the mid-point
// separator (MODE 0) or the packed mutual-kNN bits of the row (MODE 2).
template <int MODE, class Row>
__device__ __forceinline__ void ws_finish_row(const Row& R, int No, int kk, double wxr, SelState& st, bool use_row,
                                              double& t_out) {
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    double t = INF;
    if (kk <= 0) {
        t = -INF;
    } else if (kk < No && (MODE == 0 || use_row)) {
        row_select(R, kk, st.guess - wxr, st, t);
        st.guess = t + wxr;
        st.have = 1;
        if (MODE == 0) {
            double vk, vk1;
            R.neighbours(t, vk, vk1);
            double sep = vk + 0.5 * (vk1 - vk);  // slack for the other orientation's rounding
            if (!(sep < vk1)) sep = vk;
            t = sep;
        }
    }
    t_out = t;
}

template <int VPL>
__device__ __forceinline__ void ws_row_bits(const RegRow<VPL>& R, double t, double wxr, const double* cs_sm,
                                            uint32_t* out, int wp, int lane) {
    unsigned mine = 0;
#pragma unroll
    for (int u = 0; u < VPL; ++u) {
        const double v = R.v[u];
        const bool on = (v <= t) && ((v + wxr) <= cs_sm[lane + 32 * u]);
        const unsigned w = __ballot_sync(kFullMask, on);
        if (lane == u) mine = w;
    }
    if (lane < wp) out[lane] = mine;
}

__device__ __forceinline__ void ws_row_bits(const MemRow& R, double t, double wxr, const double* cs_sm, uint32_t* out,
                                            int wp, int lane) {
    for (int u = 0; u < wp; ++u) {
        const int j = lane + 32 * u;
        bool on = false;
        if (j < R.n) {
            const double v = R.row[j];
            on = (v <= t) && ((v + wxr) <= cs_sm[j]);
        }
        const unsigned w = __ballot_sync(kFullMask, on);
        if (lane == 0) out[u] = w;
    }
}

template <int BINS, int MODE, int MW>
__global__ void ACOSS_MAXNREG(168) k_sweep_ws(SweepArgs A) {
    constexpr int CPL = 4;
    constexpr int XS = kFastXS;
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    ACOSS_DYNSMEM(smem_raw);
    __shared__ int s_item;

    const int TT = (int)blockDim.x;
    const int TC = A.Wmax / 4;            // compute threads (columns / 4)
    const int TS = TT - TC, NS = TS >> 5;  // selection threads / warps
    const int R = 2 * NS;                  // rows per strip: two per selection warp
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool is_compute = tid < TC;
    const int m = (MW > 0) ? MW : A.m;
    const int RD = m + 1;
    const int PADT = fast_padt(m), TWP = TC + PADT, RS = 4 * TWP, SSN = 4 * TC;

    double* const ring = reinterpret_cast<double*>(smem_raw);
    double* const strip = ring + RD * RS;             // [2][R][SSN], natural column order
    double* const xs = strip + 2 * R * SSN;
    double* const wxs = xs + kStageRows * XS;         // [2][R]
    double* const cs_sm = wxs + 2 * R;                // [ws_cs_len(TC)]
    const int cslen = ws_cs_len(TC);

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
        const int Mo = Lr - m + 1, No = Lc - m + 1;
        if (Mo <= 0 || No <= 0) continue;
        const int nstrips = (Mo + R - 1) / R;

        if (is_compute) {
            // ================================================================== compute warps
            const int rshift = orient ? 0 : pw.oti;
            const int cshift = orient ? pw.oti : 0;
            const double* rframes = A.frames + A.offs[rsong] * BINS;
            const double* cframes = A.frames + A.offs[csong] * BINS;
            const double* rnorms = A.norms + A.offs[rsong];
            const double* cnorms = A.norms + A.offs[csong];
            double* const ring_t = ring + tid + PADT;
            int pooff[CPL];
#pragma unroll
            for (int r = 0; r < CPL; ++r) {
                const int d = r - m;
                const int q = (d >= 0) ? (d >> 2) : -((-d + 3) >> 2);
                pooff[r] = q + (d - 4 * q) * TWP;
            }
            for (int p = tid; p < RD * RS; p += TC) ring[p] = 0.0;
            double y2[CPL][BINS], wy[CPL], Pprev[CPL];
#pragma unroll
            for (int r = 0; r < CPL; ++r) {
                const int b = tid * CPL + r;
                Pprev[r] = 0.0;
                wy[r] = 0.0;
                if (b < Lc) {
#pragma unroll
                    for (int k = 0; k < BINS; ++k) {
                        int src = k - cshift;
                        if (src < 0) src += BINS;
                        y2[r][k] = -2.0 * cframes[(int64_t)b * BINS + src];
                    }
                } else {
#pragma unroll
                    for (int k = 0; k < BINS; ++k) y2[r][k] = 0.0;
                }
                const int j = b - (m - 1);
                if (j >= 0 && j < No) {
                    double w = 0.0;
                    for (int q = 0; q < m; ++q) w += cnorms[b - q];
                    wy[r] = w;
                }
            }
            int slot0 = 0, slot1 = RD - 1, slotm = 1 % RD, si = 0, s = 0;
            for (int a = 0; a < Lr; ++a) {
                if ((a & (kStageRows - 1)) == 0) {
                    for (int idx = tid; idx < kStageRows * XS; idx += TC) {
                        const int rr = idx / XS, k = idx - rr * XS;
                        const int a2 = a + rr;
                        double v = 0.0;
                        if (a2 < Lr) {
                            if (k < BINS) {
                                int src = k - rshift;
                                if (src < 0) src += BINS;
                                v = rframes[(int64_t)a2 * BINS + src];
                            } else if (k == XS - 1 && a2 >= m - 1) {
                                for (int q = 0; q < m; ++q) v += rnorms[a2 - q];
                            }
                        }
                        xs[idx] = v;
                    }
                    named_bar_sync(kBarRow, TC);
                }
                const double* xr = xs + (a & (kStageRows - 1)) * XS;
                double x[BINS];
#pragma unroll
                for (int k = 0; k + 1 < BINS; k += 2) {
                    const double2 v = *reinterpret_cast<const double2*>(xr + k);
                    x[k] = v.x;
                    x[k + 1] = v.y;
                }
                if (BINS & 1) x[BINS - 1] = xr[BINS - 1];
                const int i = a - (m - 1);
                // the strip buffer about to be written must have been released by the selection warps
                if (i >= 0 && si == 0 && s >= 2) named_bar_sync(kBarEmpty + (s & 1), TT);

                const double* r1 = ring_t + slot1 * RS;
                const double* rm = ring_t + slotm * RS;
                double* r0 = ring_t + slot0 * RS;
                double P[CPL], z[CPL];
#pragma unroll
                for (int r = 0; r < CPL; ++r) {
                    double acc = (r == 0) ? r1[3 * TWP - 1] : Pprev[r - 1];
#pragma unroll
                    for (int k = 0; k < BINS; ++k) acc = fma(x[k], y2[r][k], acc);
                    P[r] = acc;
                }
#pragma unroll
                for (int r = 0; r < CPL; ++r) {
                    z[r] = (P[r] - rm[pooff[r]]) + wy[r];
                    r0[r * TWP] = P[r];
                    Pprev[r] = P[r];
                }
                if (i >= 0) {
                    double2* srow = reinterpret_cast<double2*>(strip + ((s & 1) * R + si) * SSN + 4 * tid);
                    double2 z01, z23;
                    z01.x = z[0]; z01.y = z[1];
                    z23.x = z[2]; z23.y = z[3];
                    srow[0] = z01;
                    srow[1] = z23;
                    if (tid == 0) wxs[(s & 1) * R + si] = xr[XS - 1];
                }
                named_bar_sync(kBarRow, TC);
                slot1 = slot0;
                slot0 = (slot0 + 1 == RD) ? 0 : slot0 + 1;
                slotm = (slotm + 1 == RD) ? 0 : slotm + 1;
                if (i >= 0) {
                    if (si == R - 1 || a == Lr - 1) {
                        __threadfence_block();
                        named_bar_arrive(kBarFull + (s & 1), TT);
                        ++s;
                        si = 0;
                    } else {
                        ++si;
                    }
                }
            }
        } else {
            // ================================================================ selection warps
            const int sw = warp - (TC >> 5);
            const int kk = orient ? pw.k_col : pw.k_row;
            const int64_t rsep_off = pw.sep_off + (orient ? pw.M : 0);
            const int64_t csep_off = pw.sep_off + (orient ? 0 : pw.M);
            const bool nonmutual = (A.flags & kFlagNonMutual) != 0;
            const bool use_row = !(nonmutual && orient == 1);
            const bool use_col = !(nonmutual && orient == 0);
            const bool skip_sel = (A.flags & (1u << 30)) != 0;  // profiling knob
            const int wp = pw.crp_wp;
            if (MODE == 2) {
                for (int j = tid - TC; j < cslen; j += TS)
                    cs_sm[j] = (j < No) ? (use_col ? A.seps[csep_off + j] : INF) : -INF;
                named_bar_sync(kBarSel, TS);
            }
            SelState st;
            st.guess = 0.0;
            st.spacing = 0.0;
            st.have = 0;
            for (int s = 0; s < nstrips; ++s) {
                const int b = s & 1;
                named_bar_sync(kBarFull + b, TT);
                const int nrows = imin(R, Mo - s * R);
                for (int q = 0; q < 2; ++q) {
                    const int rr = sw * 2 + q;
                    if (rr >= nrows) break;
                    const double* rowp = strip + (b * R + rr) * SSN + (m - 1);
                    const double wxr = wxs[b * R + rr];
                    const int irow = s * R + rr;
                    uint32_t* out = A.crp + pw.crp_off + (int64_t)irow * wp;
                    double t;
                    if (No <= 512) {
                        RegRow<16> row;
                        row.load(rowp, No, lane);
                        if (skip_sel) t = 1.0; else ws_finish_row<MODE>(row, No, kk, wxr, st, use_row, t);
                        if (MODE == 2) ws_row_bits(row, t, wxr, cs_sm, out, wp, lane);
                    } else if (No <= 1024) {
                        RegRow<32> row;
                        row.load(rowp, No, lane);
                        if (skip_sel) t = 1.0; else ws_finish_row<MODE>(row, No, kk, wxr, st, use_row, t);
                        if (MODE == 2) ws_row_bits(row, t, wxr, cs_sm, out, wp, lane);
                    } else {
                        MemRow row;
                        row.load(rowp, No, lane);
                        if (skip_sel) t = 1.0; else ws_finish_row<MODE>(row, No, kk, wxr, st, use_row, t);
                        if (MODE == 2) ws_row_bits(row, t, wxr, cs_sm, out, wp, lane);
                    }
                    if (MODE == 0 && lane == 0) A.seps[rsep_off + irow] = t + wxr;
                }
                if (s + 2 < nstrips) {
                    __threadfence_block();
                    named_bar_arrive(kBarEmpty + b, TT);
                }
            }
        }
        __syncthreads();
    }
}

}  // namespace acoss
