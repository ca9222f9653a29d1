// kernels_sweep2.cuh -- the fast cross-similarity sweep for column songs of up to 1024 frames
// (everything else goes through the general kernel in kernels_sweep.cuh).
//
// Same mathematics as kernels_sweep.cuh (get_csm + sliding_csm + csm_to_binary[_mutual],
// CRPUtils.py:24-45,67-84,169-219), reorganised after the first ncu profile
// (profiles/r01_sweep_v1.md: 355 instructions per thread-row of which 48 DFMA, generic-address
// ring, runtime modulo, 4-way shared-memory bank conflicts):
//   * the squared norms are separable:  sum_k C[a-k][b-k] = D[a][b] + wx[a] + wy[b] with
//     D = windowed sum of -2 x.y only, wx / wy = windowed sums of the row / column norms.  The
//     12-DFMA chain starts directly from the diagonal predecessor, so a cell costs 12 DFMA +
//     2 DADD.  (The reference clamps each C at 0 before the window, CRPUtils.py:83; C < 0 only
//     by rounding (1e-16), far below the tie margin, so the clamp is dropped.)
//   * one CTA of T = 32*ceil(Lc/128) threads per (pair, orientation); thread t owns columns
//     4t..4t+3.  The (m+1)-deep prefix ring and the R-row strip live in shared memory in a
//     [4][T] structure-of-arrays layout: every LDS/STS.64 is lane-consecutive (conflict-free).
//     The ring is zero-padded on the left and zeroed per item, so no access is conditional.
//   * MODE 0 emits mid-point row separators (float64, absolute distance^2 units);
//     MODE 2 selects the row separators of a strip and immediately turns the strip into packed
//     CRP bits using the column separators of the other orientation: the pair is swept twice,
//     not three times, and the distances are never written outside shared memory.
#pragma once
#include "common.cuh"
#include "kernels_sweep.cuh"

#ifdef ACOSS_EMU
#define ACOSS_MAXNREG(n)
#else
#define ACOSS_MAXNREG(n) __maxnreg__(n)
#endif

namespace acoss {

constexpr int kFastXS = 14;       // staged row: 12/13 bins (+ pad), [XS-1] = windowed norm sum
constexpr int kFastMaxCols = 1408;  // 352 threads x 4 columns: 352 x 168 registers still fit one SM
constexpr int kFastMaxStripRows = 8;

__host__ __device__ inline int fast_padt(int m) { return (m + 3) / 4 + 1; }

__host__ __device__ inline size_t fast_smem_bytes(int T, int m, int R) {
    const int padt = fast_padt(m);
    size_t d = (size_t)(m + 1) * 4 * (T + padt)  // ring
               + (size_t)R * 4 * T               // strip
               + (size_t)kStageRows * kFastXS    // row-song staging
               + 2 * (size_t)R;                  // per-strip-row separators and row norms
    size_t bytes = d * sizeof(double) + (size_t)R * (T / 8 + 2) * sizeof(uint32_t);
    return (bytes + 15) & ~(size_t)15;
}

struct SelState {
    double guess;    // separator of the previous row handled by this warp, absolute units
    double spacing;  // typical distance between neighbouring values around the separator
    int have;
};

// count of strip-row values <= t (invalid positions hold +inf); n4 is a multiple of 128
__device__ __forceinline__ int strip_count(const double* row, int n4, double t, int lane) {
    int c = 0;
    const double* q = row + lane;
    for (int p = 0; p < n4; p += 128, q += 128) {
        const double v0 = q[0], v1 = q[32], v2 = q[64], v3 = q[96];
        c += (v0 <= t) ? 1 : 0;
        c += (v1 <= t) ? 1 : 0;
        c += (v2 <= t) ? 1 : 0;
        c += (v3 <= t) ? 1 : 0;
    }
    return warp_sum(c);
}

// Finds t with exactly k values <= t.  Starts from the previous row's separator and gallops
// with the previous spacing; falls back to a min/max bracket.  Returns false when exact ties
// straddle the k-th position (t then selects fewer than k).
__device__ inline bool strip_select(const double* row, int n4, int k, double g, SelState& st, int lane, double& t_out) {
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    double lo = -INF, hi = INF;
    int clo = 0, chi = 0;
    bool have_lo = false, have_hi = false, found = false;
    double t = g;
    if (st.have && g == g && st.spacing > 0.0) {
        int c = strip_count(row, n4, t, lane);
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
                c = strip_count(row, n4, t, lane);
                if (c == k) found = true;
            }
            if (!found && have_lo && have_hi && !(lo < hi)) have_lo = have_hi = false;
        }
    }
    if (!found && !(have_lo && have_hi)) {
        double vmin = INF, vmax = -INF;
        for (int p = lane; p < n4; p += 32) {
            const double v = row[p];
            vmin = fmin(vmin, v);
            if (v < INF) vmax = fmax(vmax, v);
        }
        vmin = warp_min(vmin);
        vmax = warp_max(vmax);
        lo = vmin;
        clo = strip_count(row, n4, lo, lane);
        hi = vmax;
        chi = strip_count(row, n4, hi, lane);
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
            const int c = strip_count(row, n4, t, lane);
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


// MW > 0: the embedding window m is a compile-time constant (every ring offset becomes an
// immediate or one add); MW == 0: m is read from the arguments (any window, slower addressing).
template <int BINS, int MODE, int MW>
__global__ void ACOSS_MAXNREG(168) k_sweep_fast(SweepArgs A) {
    constexpr int CPL = 4;
    constexpr int XS = kFastXS;
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    ACOSS_DYNSMEM(smem_raw);
    __shared__ int s_item;

    const int T = (int)blockDim.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, NW = T >> 5;
    const int m = (MW > 0) ? MW : A.m;
    const int RD = m + 1, R = A.R;
    const int PADT = fast_padt(m), TWP = T + PADT, RS = 4 * TWP, SS = 4 * T;

    double* const ring = reinterpret_cast<double*>(smem_raw);
    double* const strip = ring + RD * RS;
    double* const xs = strip + R * SS;
    double* const rsz = xs + kStageRows * XS;  // [R] row separators (relative units) of the strip
    double* const wxs = rsz + R;               // [R] windowed row norms of the strip rows
    uint32_t* const bitrows = reinterpret_cast<uint32_t*>(wxs + R);
    const int BW = T / 8 + 2;

    // Column 4*tid + r of a ring row sits at ring_t[r * TWP]; the cell m columns to the left at
    // ring_t[pooff(r)]; the left neighbour of column 4*tid at ring_t[3 * TWP - 1].
    double* const ring_t = ring + tid + PADT;
    int pooff[CPL];
#pragma unroll
    for (int r = 0; r < CPL; ++r) {
        const int d = r - m;
        const int q = (d >= 0) ? (d >> 2) : -((-d + 3) >> 2);
        pooff[r] = q + (d - 4 * q) * TWP;
    }

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
        const int rshift = orient ? 0 : pw.oti;
        const int cshift = orient ? pw.oti : 0;
        const int Mo = Lr - m + 1, No = Lc - m + 1;
        if (Mo <= 0 || No <= 0) continue;
        const int kk = orient ? pw.k_col : pw.k_row;
        const double* rframes = A.frames + A.offs[rsong] * BINS;
        const double* cframes = A.frames + A.offs[csong] * BINS;
        const double* rnorms = A.norms + A.offs[rsong];
        const double* cnorms = A.norms + A.offs[csong];
        const int64_t rsep_off = pw.sep_off + (orient ? pw.M : 0);
        const int64_t csep_off = pw.sep_off + (orient ? 0 : pw.M);
        const bool nonmutual = (A.flags & kFlagNonMutual) != 0;
        const bool use_row = !(nonmutual && orient == 1);
        const bool use_col = !(nonmutual && orient == 0);

        // ---- per-item setup: zero the ring, +inf the strip, load this thread's columns
        for (int p = tid; p < RD * RS; p += T) ring[p] = 0.0;
        for (int p = tid; p < R * SS; p += T) strip[p] = INF;
        double y2[CPL][BINS], wy[CPL], cs[CPL], Pprev[CPL];
        unsigned vmask = 0;
#pragma unroll
        for (int r = 0; r < CPL; ++r) {
            const int b = tid * CPL + r;
            Pprev[r] = 0.0;
            wy[r] = 0.0;
            cs[r] = -INF;
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
                vmask |= 1u << r;
                double w = 0.0;
                for (int q = 0; q < m; ++q) w += cnorms[b - q];
                wy[r] = w;
                if (MODE == 2) cs[r] = use_col ? A.seps[csep_off + j] : INF;
            }
        }
        SelState st;
        st.guess = 0.0;
        st.spacing = 0.0;
        st.have = 0;

        int slot0 = 0, slot1 = RD - 1, slotm = 1 % RD, si = 0;
        for (int a = 0; a < Lr; ++a) {
            if ((a & (kStageRows - 1)) == 0) {
                for (int idx = tid; idx < kStageRows * XS; idx += T) {
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
                __syncthreads();
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

            const double* r1 = ring_t + slot1 * RS;
            const double* rm = ring_t + slotm * RS;
            double* r0 = ring_t + slot0 * RS;
            double P[CPL];
#pragma unroll
            for (int r = 0; r < CPL; ++r) {
                double acc = (r == 0) ? r1[3 * TWP - 1] : Pprev[r - 1];
#pragma unroll
                for (int k = 0; k < BINS; ++k) acc = fma(x[k], y2[r][k], acc);
                P[r] = acc;
            }
            double* srow = strip + si * SS + tid;
#pragma unroll
            for (int r = 0; r < CPL; ++r) {
                const double z = (P[r] - rm[pooff[r]]) + wy[r];
                r0[r * TWP] = P[r];
                Pprev[r] = P[r];
                if (i >= 0 && (vmask & (1u << r))) srow[r * T] = z;
            }
            if (i >= 0 && tid == 0) wxs[si] = xr[XS - 1];
            __syncthreads();
            slot1 = slot0;
            slot0 = (slot0 + 1 == RD) ? 0 : slot0 + 1;
            slotm = (slotm + 1 == RD) ? 0 : slotm + 1;

            if (i >= 0) {
                if (si == R - 1 || a == Lr - 1) {
                    const int nrows = si + 1;
                    const int i0 = i - si;
                    // ---- selection: one warp per strip row
                    for (int rr = warp; rr < nrows; rr += NW) {
                        const double* sel = strip + rr * SS;
                        const double wxr = wxs[rr];
                        double t = INF;
                        if (kk <= 0) {
                            t = -INF;
                        } else if (A.flags & (1u << 30)) {
                            t = 1.0;  // profiling knob: skip the selection (results are meaningless)
                        } else if (kk < No && (MODE == 0 || use_row)) {
                            const double g = st.guess - wxr;
                            st.have = st.have && (g > -INF) && (g < INF);
                            strip_select(sel, SS, kk, g, st, lane, t);
                            st.guess = t + wxr;
                            st.have = 1;
                            if (MODE == 0) {
                                // mid-point between the k-th and (k+1)-th smallest: slack for the
                                // other orientation's differently rounded distances
                                double vk = -INF, vk1 = INF;
                                for (int p = lane; p < SS; p += 32) {
                                    const double v = sel[p];
                                    if (v <= t)
                                        vk = fmax(vk, v);
                                    else
                                        vk1 = fmin(vk1, v);
                                }
                                vk = warp_max(vk);
                                vk1 = warp_min(vk1);
                                double sep = vk + 0.5 * (vk1 - vk);
                                if (!(sep < vk1)) sep = vk;
                                t = sep;
                            }
                        }
                        if (MODE == 0) {
                            if (lane == 0) A.seps[rsep_off + i0 + rr] = t + wxr;
                        } else {
                            if (lane == 0) rsz[rr] = t;
                        }
                    }
                    __syncthreads();
                    if (MODE == 2) {
                        // ---- strip -> packed bits (column space), then shifted copy-out
                        for (int rr = 0; rr < nrows; ++rr) {
                            const double rz = rsz[rr], wxr = wxs[rr];
                            const double* own = strip + rr * SS + tid;
                            unsigned nib = 0;
#pragma unroll
                            for (int r = 0; r < CPL; ++r) {
                                const double v = own[r * T];
                                nib |= ((v <= rz) && ((v + wxr) <= cs[r])) ? (1u << r) : 0u;
                            }
                            unsigned w = nib << (4 * (lane & 7));
                            w |= __shfl_xor_sync(kFullMask, w, 1);
                            w |= __shfl_xor_sync(kFullMask, w, 2);
                            w |= __shfl_xor_sync(kFullMask, w, 4);
                            if ((lane & 7) == 0) bitrows[rr * BW + (tid >> 3)] = w;
                        }
                        __syncthreads();
                        const int wp = pw.crp_wp;
                        const unsigned sh = (unsigned)(m - 1) & 31u;
                        const int wo = (m - 1) >> 5;
                        const int nwords = T / 8;
                        for (int rr = 0; rr < nrows; ++rr) {
                            const uint32_t* br = bitrows + rr * BW;
                            uint32_t* out = A.crp + pw.crp_off + (int64_t)(i0 + rr) * wp;
                            for (int w = tid; w < wp; w += T) {
                                const uint32_t lo = (w + wo < nwords) ? br[w + wo] : 0u;
                                const uint32_t hi = (w + wo + 1 < nwords) ? br[w + wo + 1] : 0u;
                                out[w] = __funnelshift_r(lo, hi, sh);
                            }
                        }
                    }
                    si = 0;
                } else {
                    ++si;
                }
            }
        }
        __syncthreads();
    }
}

}  // namespace acoss
