// kernels_sweep6.cuh -- the cross-similarity sweep on the FP64 tensor cores (DMMA), for 12 bins,
// m = 9 and column songs of up to 631 frames.
//
// Keys, separators, selection and the float64 fallback are those of kernels_sweep4.cuh; the
// distances are produced differently.  The SIMT sweeps keep the column song in 96 registers per
// thread (4 columns x 12 bins), which caps the SM at 12 warps and leaves the kernel latency bound
// (profiles/r01_e, r01_f: FP64 pipe 15-33 % busy, issue slots 47 %).  Here, per PERIOD of 16 rows:
//   STAGE 1  16 rows x all columns of  |y|^2 - 2 x.y  by mma.sync.m8n8k4.f64 (DMMA): a warp owns 32
//            columns, whose fragments are 12 doubles per lane (pre-scaled by -2, OTI-rotated) plus
//            the columns' squared norms as accumulator seeds; the row fragments come from the
//            TMA-staged row block.  The 8 x 8 tiles go to a shared-memory slab.
//   STAGE 2  every thread owns one DIAGONAL of the pair's matrix (thread t: column - row = t mod
//            NT) and walks it through the slab: running window sum W += v - v_9_rows_ago with the
//            last 16 values in registers, so there is no prefix sum, no ring in shared memory and
//            no dependency between threads.  W starts at 1.5 * 2^e, i.e. it IS the fixed-point key
//            (s4_fix: one funnel shift) at every step; the column's windowed norm is already inside.
//            Threads whose diagonal stays inside the matrix for the whole period run a branch-free
//            path (load, 2 DADD, shift, store per cell).
//   SELECT   the 16 rows of the previous period are resolved, one row per warp (s4_resolve_row,
//            keys in registers), before the barrier that releases the slab.
// Two CTA-wide barriers per 16 rows; <= 128 registers per thread, 16 warps per SM at 512 threads.
#pragma once
#include "kernels_sweep4.cuh"

namespace acoss {

// profiling aid (build with -DACOSS_S6_PROFILE, then debug flag bit 26): cycles spent by every warp's lane 0 in
// [0] stage 2, [1] selection, [2] stage 1, [3] waiting at the period barrier, [4] periods, [5] items.
// Compile-time because the 18 registers of the counters cost the 128-register kernels spills.
#if defined(ACOSS_S6_PROFILE) && !defined(ACOSS_EMU)
#define ACOSS_S6_PROF 1
__device__ unsigned long long g_s6_prof[8];
#else
#define ACOSS_S6_PROF 0
#endif

constexpr int kS6Rows = 16;         // rows per period
constexpr int kS6MaxThreads = 640;  // column songs of up to 631 frames (102 registers per thread)
constexpr int kS6Pad = 9;            // NT >= Lc + kS6Pad: zero columns that flush the window (m = 9)

// CTA sizes: 128, 256, 512 (slab double-buffered, one barrier per period) and 640 (single slab, two barriers)
__host__ __device__ constexpr bool s6_double_buffered(int NT) { return NT <= 512; }
__host__ __device__ constexpr int s6_threads_for(int Lc) {
    return Lc + 9 <= 128 ? 128 : (Lc + 9 <= 256 ? 256 : (Lc + 9 <= 512 ? 512 : 640));
}
__host__ __device__ constexpr int s6_bucket_of(int NT) { return NT == 128 ? 0 : (NT == 256 ? 1 : (NT == 512 ? 2 : 3)); }
constexpr int kS6Buckets = 4;

struct S6Layout {
    int xs, dots, fsepx, strip, csk, fcol, fblo, cand, misc, mbar, total, lcp;
};
__host__ __device__ inline S6Layout s6_layout(int NT, int bins) {
    S6Layout L;
    L.lcp = NT + 8;  // row stride of the slab (doubles): = 8 mod 16, so the 8 x 8 tiles store conflict-free
    int o = 0;
    L.xs = o;    o += 2 * kS4Stage * bins * 8;
    L.dots = o;  o += (s6_double_buffered(NT) ? 2 : 1) * kS6Rows * L.lcp * 8;
    L.fsepx = o; o += kS4MaxFlag * 8;
    L.strip = o; o += 2 * kS6Rows * NT * 4;
    L.csk = o;   o += NT * 4;
    L.fcol = o;  o += kS4MaxFlag * 4;
    L.fblo = o;  o += kS4MaxFlag * 4;
    L.cand = o;  o += kS6Rows * 2 * kS4MaxFlag * 4;
    L.misc = o;  o += (16 + kS6Rows) * 4;
    o = (o + 7) & ~7;
    L.mbar = o;  o += 2 * 8;
    L.total = (o + 15) & ~15;
    return L;
}
__host__ __device__ constexpr int s6_min_blocks(int NT) { return NT <= 128 ? 3 : (NT <= 256 ? 2 : 1); }

// D(8x8) += A(8x4) B(4x8) in float64.  Lane l holds A[l/4][l%4], B[l%4][l/4] and
// D[l/4][2(l%4)], D[l/4][2(l%4)+1].
__device__ __forceinline__ void s6_dmma(double& c0, double& c1, double a, double b) {
#ifndef ACOSS_EMU
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
#else
    const int lane = threadIdx.x & 31, row = lane >> 2;
    for (int j = 0; j < 2; ++j) {
        const int col = 2 * (lane & 3) + j;
        double acc = j ? c1 : c0;
        for (int k = 0; k < 4; ++k) {
            const double aa = __shfl_sync(kFullMask, a, row * 4 + k);
            const double bb = __shfl_sync(kFullMask, b, col * 4 + k);
            acc = fma(aa, bb, acc);
        }
        if (j) c1 = acc; else c0 = acc;
    }
#endif
}

template <int MODE, int MW, int NT>
__global__ void __launch_bounds__(NT, s6_min_blocks(NT)) k_sweep6(Sweep4Args A) {
    constexpr int BINS = 12;
    constexpr int NW = NT / 32, NQ = NT / 128, LCP = NT + 8;
    constexpr bool DB = s6_double_buffered(NT);
    static_assert(MW <= kS6Pad && NT % 128 == 0, "the zero pad must flush the window");
    ACOSS_DYNSMEM(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int m = MW;
    const int G = A.gap, BANDW = 2 * A.gap + 6;
    const S6Layout LY = s6_layout(NT, BINS);

    double* const xs = reinterpret_cast<double*>(smem_raw + LY.xs);
    double* const dots = reinterpret_cast<double*>(smem_raw + LY.dots);  // [DB ? 2 : 1][16][LCP]
    double* const fsepx = reinterpret_cast<double*>(smem_raw + LY.fsepx);
    int* const stripi = reinterpret_cast<int*>(smem_raw + LY.strip);     // [2][16][NT] keys
    int* const csk_s = reinterpret_cast<int*>(smem_raw + LY.csk);
    int* const fcol = reinterpret_cast<int*>(smem_raw + LY.fcol);
    int* const fblo = reinterpret_cast<int*>(smem_raw + LY.fblo);
    int* const cand = reinterpret_cast<int*>(smem_raw + LY.cand);
    int* const misc = reinterpret_cast<int*>(smem_raw + LY.misc);
    int* const s_item = misc;
    int* const s_nflag = misc + 1;
    uint64_t* const mbar = reinterpret_cast<uint64_t*>(smem_raw + LY.mbar);

    if (tid == 0) {
        s4_mbar_init(&mbar[0], 1);
        s4_mbar_init(&mbar[1], 1);
        s4_fence_mbar_init();
    }
    __syncthreads();
    unsigned par0 = 0, par1 = 0;

    for (;;) {
        if (tid == 0) {
            *s_item = (int)atomicAdd(A.counter, 1u);
            *s_nflag = 0;
            misc[8] = misc[9] = misc[10] = 0;
        }
        __syncthreads();
        const int it = *s_item;
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
        const PairWork* const pwp = A.pairs + pidx;
        const int rsong = orient ? pwp->song_j : pwp->song_i;
        const int csong = orient ? pwp->song_i : pwp->song_j;
        const int Lr = orient ? pwp->Lj : pwp->Li;
        const int Lc = orient ? pwp->Li : pwp->Lj;
        const int Mo = Lr - m + 1, No = Lc - m + 1;
        if (Mo <= 0 || No <= 0) continue;
        const int kk = orient ? pwp->k_col : pwp->k_row;
        const double* rframes = A.frames + A.offs[rsong] * BINS;
        const double* rwn = A.wnorm + A.offs[rsong];
        const int64_t rsep_off = pwp->sep_off + (orient ? pwp->M : 0);
        const int64_t csep_off = pwp->sep_off + (orient ? 0 : pwp->M);
        const bool nonmutual = (A.flags & kFlagNonMutual) != 0;
        const bool use_row = !(nonmutual && orient == 1);
        const bool use_col = !(nonmutual && orient == 0);
        const double BIGE = ldexp(1.0, pwp->expo);
        const double OFF = 1.5 * BIGE;
        S4Pair pp;
        pp.fi = A.frames + A.offs[pwp->song_i] * BINS;
        pp.fj = A.frames + A.offs[pwp->song_j] * BINS;
        pp.bins = BINS;
        pp.oti = pwp->oti;
        pp.m = m;
        pp.orient = orient;

        // ---- per-item setup
        if (tid == 0) {
            const unsigned n0 = (unsigned)imin(kS4Stage, Lr) * BINS * 8u;
            s4_bulk_load(xs, rframes, n0, &mbar[0]);
        }
        // column fragments of this warp's 32 columns: block bi = columns 32 warp + 8 bi .. + 7; lane
        // holds column (l / 4) of the block, bins 4 s + (l % 4), rolled by OTI and scaled by -2; the
        // accumulator seeds are the squared norms of columns 2 (l % 4), 2 (l % 4) + 1 of the block
        double yf[4][3], cn[4][2];
        {
            const double* cframes = A.frames + A.offs[csong] * BINS;
            const double* cnorm = A.norms + A.offs[csong];
            const int crot = orient ? (BINS - pwp->oti) % BINS : pwp->oti;
#pragma unroll
            for (int bi = 0; bi < 4; ++bi) {
                const int col = 32 * warp + 8 * bi + (lane >> 2);
#pragma unroll
                for (int s = 0; s < 3; ++s) {
                    int src = 4 * s + (lane & 3) + crot;
                    if (src >= BINS) src -= BINS;
                    yf[bi][s] = (col < Lc) ? -2.0 * cframes[(int64_t)col * BINS + src] : 0.0;
                }
                const int c2 = 32 * warp + 8 * bi + 2 * (lane & 3);
                cn[bi][0] = (c2 < Lc) ? cnorm[c2] : 0.0;
                cn[bi][1] = (c2 + 1 < Lc) ? cnorm[c2 + 1] : 0.0;
            }
        }
        {
            const int j = tid - (m - 1);
            if (MODE == 2) {
                int v = -1;
                if (j >= 0 && j < No) {
                    v = use_col ? A.sepk[csep_off + j] : (kKeyInf - 1);
                    if (v < 0) {
                        const int blo = -v - 1;
                        const int slot = atomicAdd(s_nflag, 1);
                        if (slot < kS4MaxFlag) {
                            fcol[slot] = tid;
                            fblo[slot] = blo;
                            fsepx[slot] = A.sepx[csep_off + j];
                            v = blo - 1;
                        } else {
                            misc[10] = 1;  // list overflow: the encoded value stays (generic path)
                        }
                    }
                }
                csk_s[tid] = v;
            }
            // warps without columns never write their part of the slab: it is the zero pad
            if (32 * warp >= Lc)
                for (int r = 0; r < (DB ? 2 : 1) * kS6Rows; ++r) dots[r * LCP + tid] = 0.0;
            // every key slot starts invalid
            uint4* s4p = reinterpret_cast<uint4*>(stripi);
            const uint4 inf4 = make_uint4((unsigned)kKeyInf, (unsigned)kKeyInf, (unsigned)kKeyInf, (unsigned)kKeyInf);
            for (int p = tid; p < 2 * kS6Rows * NT / 4; p += NT) s4p[p] = inf4;
        }
        // The matrix is padded to NT >= Lc + 9 columns whose values are zero, and thread t walks the
        // cyclic diagonal column = (row + t) mod NT: by the time it re-enters at column 0 the window
        // has seen 9 zeros, so W is back at its offset and nothing needs to be reset.
        int cb = tid;  // column at the first row of the period
        double W = OFF;
        double ring[16];
#pragma unroll
        for (int q = 0; q < 16; ++q) ring[q] = 0.0;
        S4Sel st;
        st.spacing = 1.0f;
        st.scale = 0.0f;
        int spar = 0;
        __syncthreads();
        const int nflag = (MODE == 2) ? imin(*s_nflag, kS4MaxFlag) : 0;
        const int nper = (Lr + kS6Rows - 1) / kS6Rows;
        const int est_rank = (kk >= 1 && kk < No) ? s4_est_rank(kk, No) : -1;

        // ---- stage 1 of period q: this warp's 32 columns x 16 rows into slab q & 1
        auto stage1 = [&](int q) {
            const int a0 = q * kS6Rows;
            const int buf = (a0 >> 5) & 1;
            if ((q & 1) == 0) {
                // a new block of 32 rows: prefetch the next one into the other buffer (its readers,
                // stage 1 of the previous block, are all behind a barrier), wait for this one
                if (tid == 0 && a0 + kS4Stage < Lr) {
                    const unsigned nn = (unsigned)imin(kS4Stage, Lr - (a0 + kS4Stage)) * BINS * 8u;
                    s4_bulk_load(xs + (buf ^ 1) * kS4Stage * BINS, rframes + (int64_t)(a0 + kS4Stage) * BINS, nn,
                                 &mbar[buf ^ 1]);
                }
                if (buf == 0) {
                    s4_mbar_wait(&mbar[0], par0);
                    par0 ^= 1u;
                } else {
                    s4_mbar_wait(&mbar[1], par1);
                    par1 ^= 1u;
                }
            }
            if (32 * warp < Lc) {
                double* const slab = dots + (DB ? (q & 1) * kS6Rows * LCP : 0);
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    if (a0 + 8 * h < Lr) {
                        const double* xr =
                            xs + (buf * kS4Stage + (a0 & (kS4Stage - 1)) + 8 * h + (lane >> 2)) * BINS + (lane & 3);
                        const double xa0 = xr[0], xa1 = xr[4], xa2 = xr[8];
                        double* dst = slab + (8 * h + (lane >> 2)) * LCP + 32 * warp + 2 * (lane & 3);
#pragma unroll
                        for (int bi = 0; bi < 4; ++bi) {
                            double c0 = cn[bi][0], c1 = cn[bi][1];
                            s6_dmma(c0, c1, xa0, yf[bi][0]);
                            s6_dmma(c0, c1, xa1, yf[bi][1]);
                            s6_dmma(c0, c1, xa2, yf[bi][2]);
                            double2 cc;
                            cc.x = c0;
                            cc.y = c1;
                            *reinterpret_cast<double2*>(dst + 8 * bi) = cc;
                        }
                    }
                }
            }
        };
        if (DB) {
            stage1(0);
            __syncthreads();
        }

#if ACOSS_S6_PROF
        const bool prof = (A.flags & (1u << 26)) != 0 && lane == 0;
        long long tp0 = 0, tp1 = 0, tp2 = 0, tp3 = 0, tp4 = 0, pa0 = 0, pa1 = 0, pa2 = 0, pa3 = 0;
        if (prof && tid == 0) atomicAdd(&g_s6_prof[5], 1ull);
#define S6_TICK(v) if (prof) v = clock64()
#else
#define S6_TICK(v)
#endif
        for (int p = 0; p <= nper; ++p) {
            const int a0 = p * kS6Rows;
            S6_TICK(tp0);
            if (!DB) {
                if (p < nper) stage1(p);
                __syncthreads();  // the slab of period p is complete
            }
            const double* const dslab = dots + (DB ? (p & 1) * kS6Rows * LCP : 0);

            // ---- stage 2: walk this thread's diagonal through the slab
            if (p < nper) {
                int* const sl = stripi + (p & 1) * kS6Rows * NT;
                if (cb + kS6Rows <= NT) {
                    const double* dp = dslab + cb;
                    int* sp = sl + cb;
                    if (a0 >= MW - 1 && cb >= MW - 1 && cb + kS6Rows <= Lc) {
                        // inside the matrix for all 16 rows, every key valid
#pragma unroll
                        for (int hh = 0; hh < 2; ++hh) {
                            double v[8];
#pragma unroll
                            for (int r = 0; r < 8; ++r) v[r] = dp[(8 * hh + r) * (LCP + 1)];
#pragma unroll
                            for (int r = 0; r < 8; ++r) {
                                W += v[r] - ring[(8 * hh + r - MW) & 15];
                                ring[8 * hh + r] = v[r];
                                sp[(8 * hh + r) * (NT + 1)] = s4_fix(W);
                            }
                        }
                    } else {
                        // no wrap, but some keys are invalid (first rows / first columns / the zero pad)
#pragma unroll
                        for (int hh = 0; hh < 2; ++hh) {
                            double v[8];
#pragma unroll
                            for (int r = 0; r < 8; ++r) v[r] = dp[(8 * hh + r) * (LCP + 1)];
#pragma unroll
                            for (int r = 0; r < 8; ++r) {
                                const int rr = 8 * hh + r;
                                W += v[r] - ring[(rr - MW) & 15];
                                ring[rr] = v[r];
                                const bool ok = (unsigned)(cb + rr - (MW - 1)) < (unsigned)No && a0 + rr >= MW - 1;
                                sp[rr * (NT + 1)] = ok ? s4_fix(W) : kKeyInf;
                            }
                        }
                    }
                } else {
                    // the diagonal wraps from the zero pad back to column 0 inside this period
#pragma unroll
                    for (int r = 0; r < kS6Rows; ++r) {
                        int col = cb + r;
                        if (col >= NT) col -= NT;
                        const double v = dslab[r * LCP + col];
                        W += v - ring[(r - MW) & 15];
                        ring[r] = v;
                        const bool ok = (unsigned)(col - (MW - 1)) < (unsigned)No && a0 + r >= MW - 1;
                        sl[r * NT + col] = ok ? s4_fix(W) : kKeyInf;
                    }
                }
                cb += kS6Rows;
                if (cb >= NT) cb -= NT;
            }

            S6_TICK(tp1);
            // ---- selection of period p - 1: one row per warp
            if (p >= 1) {
                const int ap = (p - 1) * kS6Rows;
                const int rfirst = imax(0, (m - 1) - ap), rlast = imin(kS6Rows - 1, Lr - 1 - ap);
                if (rlast >= rfirst) {
                    for (int r = warp; r <= rlast; r += NW) {
                        if (r < rfirst) continue;
                        S4RowArgs ra;
                        ra.row4 = reinterpret_cast<const uint4*>(stripi + (((p - 1) & 1) * kS6Rows + r) * NT);
                        ra.nq = NQ;
                        ra.srow = r - rfirst;
                        ra.nrows = rlast - rfirst + 1;
                        ra.hist_back = 4;
                        ra.use_hist = (A.flags & (1u << 28)) != 0;
                        ra.est_rank = (A.flags & (1u << 27)) ? est_rank : -1;  // measured: the history guess is as fast
                        ra.irow = ap + r - (m - 1);
                        ra.lane = lane;
                        ra.kk = kk;
                        ra.No = No;
                        ra.m = m;
                        ra.G = G;
                        ra.BANDW = BANDW;
                        ra.use_row = use_row;
                        ra.rwn = rwn;
                        ra.BIGE = BIGE;
                        ra.rsep_off = rsep_off;
                        ra.csep_off = csep_off;
                        ra.misc = misc;
                        ra.spar = spar;
                        ra.cand_s = cand + r * 2 * kS4MaxFlag;
                        ra.s_wcnt = misc + 16 + r;
                        ra.fcol = fcol;
                        ra.fblo = fblo;
                        ra.fsepx = fsepx;
                        ra.csk_s = csk_s;
                        ra.nflag = nflag;
                        ra.nb_off = pwp->nb_off;
                        ra.nb_stride = pwp->nb_stride;
                        s4_resolve_row<BINS, MODE, NQ>(A, pp, ra, st);
                    }
                    spar ^= 1;
                }
            }
            S6_TICK(tp2);
            if (DB && p + 1 < nper) stage1(p + 1);  // overlaps the other warps' selection
            S6_TICK(tp3);
            __syncthreads();  // the keys of period p and (DB) the slab of period p + 1 are complete
#if ACOSS_S6_PROF
            if (prof) {
                tp4 = clock64();
                pa0 += tp1 - tp0;
                pa1 += tp2 - tp1;
                pa2 += tp3 - tp2;
                pa3 += tp4 - tp3;
            }
#endif
        }
#if ACOSS_S6_PROF
        if (prof) {
            atomicAdd(&g_s6_prof[0], (unsigned long long)pa0);
            atomicAdd(&g_s6_prof[1], (unsigned long long)pa1);
            atomicAdd(&g_s6_prof[2], (unsigned long long)pa2);
            atomicAdd(&g_s6_prof[3], (unsigned long long)pa3);
            if (tid == 0) atomicAdd(&g_s6_prof[4], (unsigned long long)(nper + 1));
        }
#endif
    }
}

}  // namespace acoss
