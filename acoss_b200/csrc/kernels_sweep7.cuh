// kernels_sweep7.cuh -- the cross-similarity sweep, seventh generation: the default for 12 / 13
// bins, m = 9 and column songs of up to 1015 frames.
//
// Distances, keys, separators and the float64 repair of ambiguous rows are those of
// kernels_sweep4.cuh / _sweep6.cuh (get_csm + sliding_csm + csm_to_binary[_mutual],
// CRPUtils.py:24-45,67-84,169-219).  What changed after the round-1 captures (profiles/r01_g: one
// row's selection = ~800 dependent warp instructions, 6 200 of the 12 200 cycles of a 16-row
// period, barrier stalls 25-42 %):
//
//   * SELECTION IN CHAINS OF CONSECUTIVE ROWS.  A strip of R = 32..64 rows of keys is kept in shared
//     memory and warp w resolves rows w R/NW .. (w+1) R/NW - 1 one after the other, the row's keys in
//     registers.  Row r+1 of the matrix is row r shifted by one column plus one new frame, so the
//     separator of row r is an excellent first threshold for row r+1 (measured on the bench songs: the
//     count at that threshold is within 6 ranks of k for 68 % of the rows, median error 3 ranks, against
//     a median error of 25-38 ranks for a separator 16-64 rows back, which is what gens 4-6 used): one
//     counting pass and one candidate walk (s4_select / s4_walk of kernels_sweep4.cuh) resolve most
//     rows.  The first row of a chain starts from the order-statistic estimate of its own keys
//     (s4_estimate).  Every warp has the same number of rows, so the strip needs one barrier.
//   * 8-ROW PERIODS.  The DMMA slab is 8 rows (one m8n8k4 tile row) and double buffered where
//     shared memory allows, which leaves room for the strip.
//   * PACKED CRP FROM THE SWEEP.  MODE 2 turns a strip into CRP words with one __ballot_sync per
//     (row, 32 columns) -- thread = CRP column -- and writes them straight into the layout the
//     DP kernel reads.  The nibble-byte arena and k_nb_to_packed of gens 4-6 are gone: a CRP
//     leaves the SM once, as 1 bit per cell.
//   * 13 bins (the mfcc half of Serra09.similarity, Serra09.py:177-184) run on the same kernel:
//     K is padded to 16 for the DMMA fragments.
//   * Column capacities 128, 256, ..., 1024 in steps of 128 (the diagonal walk and the selection cost NC per row,
//     whatever the song's length): one diagonal per thread up to 256 columns, two beyond.
#pragma once
#include <type_traits>

#include "kernels_sweep4.cuh"

namespace acoss {

constexpr int kS7Pad = 9;      // NC >= Lc + kS7Pad zero columns flush the window (m <= 9)
constexpr int kS7Per = 8;      // rows per period (one DMMA tile row)

#if defined(ACOSS_S7_STATS) && !defined(ACOSS_EMU)
#define ACOSS_S7_STAT 1
// [0] rows resolved, [1] unsafe rows, [2] strips, [3] items; [4 + k] cycles of lane 0 of every warp in phase k:
// 0 item setup, 1 stage 1, 2 stage 2, 3 period barrier, 4 strip init + barrier, 5 selection, 6 barrier behind the
// selection, 7 export / ballots, 8 flag scan + float64 fix-ups, 9 CRP copy-out, 10 last barrier of the strip, 11 item fetch
__device__ unsigned long long g_s7_stats[16];
#else
#define ACOSS_S7_STAT 0
#endif

template <int NT_, int CPT_>
struct S7Cfg {
    static constexpr int NT = NT_, CPT = CPT_;
    static constexpr int NC = NT * CPT;                    // padded columns
    static constexpr int NW = NT / 32;
    static constexpr int NQ = NC / 128;                    // uint4 of a strip row per lane (selection: keys in registers)
    // Rows per strip, slab buffering and CTAs per SM by column capacity (shared memory: R * NC * 4 bytes of keys):
    //   128, 256            one diagonal per thread, 64-row strips, two slabs, 4 / 2 CTAs per SM
    //   384, 512            two diagonals per thread (192 / 256 threads), 48 / 32-row strips, one slab, 2 CTAs per SM
    //                       (independent barrier domains)
    //   640                 320 threads x 2, 40-row strips, two slabs, 1 CTA per SM
    //   768, 896, 1024      384 / 448 / 512 threads x 2, 48 / 32 / 32-row strips, one slab, 1 CTA per SM
    //   512 as 512 x 1      measured alternative (debug flag): 64-row strips, two slabs, 1 CTA per SM
    static constexpr bool WIDE1 = (NC == 512 && NT == 512);
    static constexpr int R = (NC <= 256 || WIDE1) ? 64 : NC == 384 ? 48 : NC == 512 ? 32 : NC == 640 ? 40 : NC == 768 ? 48 : 32;
    static constexpr bool DBS = NC <= 256 || WIDE1 || NC == 640;   // double-buffered slab
    static constexpr int MINB = NC <= 128 ? 4 : ((NC <= 512 && NT <= 256) ? 2 : 1);
    static constexpr int LCP = NC + 8;                     // slab row stride (doubles), = 8 mod 16
    static constexpr int WPR = NC / 32;                    // CRP words per strip row
    static constexpr int RPW = (R + NW - 1) / NW;          // rows per warp: the chain resolved with lag-1 guesses
    static_assert(NT % 32 == 0 && NC % 128 == 0 && R % kS7Per == 0 && R <= NT, "shape");
};

constexpr int kS7Buckets = 8;                 // column capacities 128, 256, ..., 1024
constexpr int kS7MaxCols = 1024 - kS7Pad;

struct S7Layout {
    int xs, slab, csq, strip, csk, bits, rinfo, rsepx, cand, misc, mbar, total;
};
template <int NT, int CPT>
__host__ __device__ inline S7Layout s7_layout(int bins, int mode) {
    using C = S7Cfg<NT, CPT>;
    S7Layout L;
    int o = 0;
    L.xs = o;    o += 2 * kS4Stage * bins * 8;
    L.slab = o;  o += (C::DBS ? 2 : 1) * kS7Per * C::LCP * 8;
    L.csq = o;   o += C::NC * 8;
    L.rsepx = o; o += C::R * 8;
    L.strip = o; o += C::R * C::NC * 4;
    L.csk = o;   o += (mode == 2 ? C::NC * 4 : 0);
    L.bits = o;  o += (mode == 2 ? C::R * C::WPR * 4 : 0);
    L.rinfo = o; o += 4 * C::R * 4;                      // thr, kwx, blo, flag
    L.cand = o;  o += C::NW * (kS4MaxFlag + 1) * 4;      // per warp: candidate columns + their count
    L.misc = o;  o += 32 * 4;                            // [0] item, [5] flagged-column mark, [6] flagged row in the strip, [16..31] phase cycles
    o = (o + 7) & ~7;
    L.mbar = o;  o += 2 * 8;
    L.total = (o + 15) & ~15;
    return L;
}

// D(8x8) += A(8x4) B(4x8) in float64.  Lane l holds A[l/4][l%4], B[l%4][l/4] and
// D[l/4][2(l%4)], D[l/4][2(l%4)+1].
__device__ __forceinline__ void s7_dmma(double& c0, double& c1, double a, double b) {
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

struct Sweep7Args {
    const double* frames;   // [total_frames][BINS]
    const double* wnorm;    // [total_frames] windowed squared norms
    const double* norms;    // [total_frames] squared norms of the frames
    const int64_t* offs;
    const PairWork* pairs;
    const uint32_t* items;  // MODE 0: pair*2 + orientation; MODE 2: pair
    int n_items;
    unsigned int* counter;
    int32_t* sepk;          // integer separators
    double* sepx;           // exact float64 separators of flagged entries
    uint32_t* crp;          // MODE 2 output: packed CRPs (PairWork::crp_off, crp_wp)
    int gap;                // G: keys closer than 2G to a boundary make the row "unsafe"
    uint32_t flags;
};

template <int MODE, int BINS, int MW, int NT, int CPT>
__global__ void __launch_bounds__(NT, (S7Cfg<NT, CPT>::MINB)) k_sweep7(Sweep7Args A) {
    using C = S7Cfg<NT, CPT>;
    constexpr int NC = C::NC, NW = C::NW, NQ = C::NQ, R = C::R, RPW = C::RPW, LCP = C::LCP, WPR = C::WPR;
    constexpr bool DBS = C::DBS;
    constexpr int KS = (BINS + 3) / 4;  // DMMA k-steps
    constexpr int PPS = R / kS7Per;     // periods per strip
    static_assert(MW <= kS7Pad && MW <= 15, "the zero pad must flush the window");
    ACOSS_DYNSMEM(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int m = MW;
    const int G = A.gap, BANDW = 2 * A.gap + 6;
    const S7Layout LY = s7_layout<NT, CPT>(BINS, MODE);

    double* const xs = reinterpret_cast<double*>(smem_raw + LY.xs);
    double* const slab = reinterpret_cast<double*>(smem_raw + LY.slab);  // [DBS ? 2 : 1][8][LCP]
    double* const csq = reinterpret_cast<double*>(smem_raw + LY.csq);    // squared norms of the columns
    double* const rsepx = reinterpret_cast<double*>(smem_raw + LY.rsepx);
    int* const strip = reinterpret_cast<int*>(smem_raw + LY.strip);      // [R][NC] keys
    int* const csk_s = reinterpret_cast<int*>(smem_raw + LY.csk);        // [NC] by CRP column (MODE 2)
    uint32_t* const bits = reinterpret_cast<uint32_t*>(smem_raw + LY.bits);  // [R][wp] CRP words (MODE 2)
    int* const rthr = reinterpret_cast<int*>(smem_raw + LY.rinfo);       // row-relative certainly-in bound
    int* const rkwx = rthr + R;                                          // the row's windowed norm in key units
    int* const rblo = rthr + 2 * R;
    int* const rflag = rthr + 3 * R;                                     // 1: flagged (float64 separator in rsepx)
    int* const cand = reinterpret_cast<int*>(smem_raw + LY.cand) + warp * (kS4MaxFlag + 1);
    int* const misc = reinterpret_cast<int*>(smem_raw + LY.misc);
    int* const s_item = misc;        // [0] item, [5] any flagged column
    uint64_t* const mbar = reinterpret_cast<uint64_t*>(smem_raw + LY.mbar);

    if (tid == 0) {
        s4_mbar_init(&mbar[0], 1);
        s4_mbar_init(&mbar[1], 1);
        s4_fence_mbar_init();
    }
    __syncthreads();
    unsigned par0 = 0, par1 = 0;
#if ACOSS_S7_STAT
    int* const s_prof = misc + 16;
    int tl_ = (int)clock();
    if (tid < 16) s_prof[tid] = 0;
#define S7_LAP(k)                                   \
    if (lane == 0) {                                \
        const int t_ = (int)clock();                \
        atomicAdd(&s_prof[k], t_ - tl_);            \
        tl_ = t_;                                   \
    }
#else
#define S7_LAP(k)
#endif
    S4Sel st;  // key spacing around the separator, carried from row to row
    st.spacing = 1.0f;
    st.scale = 0.0f;

    for (;;) {
        if (tid == 0) {
            *s_item = (int)atomicAdd(A.counter, 1u);
            misc[5] = 0;
        }
        __syncthreads();
        const int it = *s_item;
        __syncthreads();
        if (it >= A.n_items) break;
        S7_LAP(11);

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
        const int wp = pwp->crp_wp;
        S4Pair pp;
        pp.fi = A.frames + A.offs[pwp->song_i] * BINS;
        pp.fj = A.frames + A.offs[pwp->song_j] * BINS;
        pp.bins = BINS;
        pp.oti = pwp->oti;
        pp.m = m;
        pp.orient = orient;
        const bool do_select = kk >= 1 && kk < No && (MODE == 0 || use_row) && !(A.flags & (1u << 30));

        // ---- per-item setup
        // row frames: 32 rows per stage, by one bulk copy when the source is 16-byte aligned (always for 12 bins)
        const bool bulk_ok = ((BINS * 8) % 16 == 0) || ((A.offs[rsong] & 1) == 0);
        auto stage_rows = [&](int blk, int buf) {  // uniform; returns after issuing
            const int r0 = blk * kS4Stage;
            const int nr = imin(kS4Stage, Lr - r0);
            const unsigned nb = (unsigned)(nr * BINS * 8);
            if (bulk_ok && (nb % 16u) == 0u) {
                if (tid == 0) s4_bulk_load(xs + buf * kS4Stage * BINS, rframes + (int64_t)r0 * BINS, nb, &mbar[buf]);
            } else {
                // cooperative copy; the barrier still flips so that the waits stay uniform
                for (int e = tid; e < nr * BINS; e += NT) xs[buf * kS4Stage * BINS + e] = rframes[(int64_t)r0 * BINS + e];
                __threadfence_block();
                if (tid == 0) s4_bulk_load(xs + buf * kS4Stage * BINS, rframes, 0u, &mbar[buf]);
            }
        };
        stage_rows(0, 0);
        // column fragments: block bi of set h = columns 32 (warp + NW h) + 8 bi .. + 7; lane holds column
        // (l / 4) of the block, bins 4 s + (l % 4), rolled by OTI and scaled by -2
        double yf[CPT][4][KS];
        {
            const double* cframes = A.frames + A.offs[csong] * BINS;
            const double* cnorm = A.norms + A.offs[csong];
            const int crot = orient ? (BINS - pwp->oti) % BINS : pwp->oti;
#pragma unroll
            for (int h = 0; h < CPT; ++h) {
#pragma unroll
                for (int bi = 0; bi < 4; ++bi) {
                    const int col = 32 * (warp + NW * h) + 8 * bi + (lane >> 2);
#pragma unroll
                    for (int s = 0; s < KS; ++s) {
                        const int b = 4 * s + (lane & 3);
                        int src = b + crot;
                        if (src >= BINS) src -= BINS;
                        yf[h][bi][s] = (col < Lc && b < BINS) ? -2.0 * cframes[(int64_t)col * BINS + src] : 0.0;
                    }
                }
                const int c = tid + NT * h;
                // accumulator seeds of the dot products: the column's squared norm ON TOP OF the fixed-point offset
                // 1.5 * 2^e, so that every slab entry lies in the binade [2^e, 2^(e+1)) and its bit pattern is a
                // fixed-point number (stage 2 then adds and subtracts the patterns as 64-bit integers: exact, and
                // off the FP64 pipe the DMMAs need)
                csq[c] = ((c < Lc) ? cnorm[c] : 0.0) + OFF;
                // column sets that hold no column of the song are never written by stage 1: they are the zero pad
                if (32 * (warp + NW * h) >= Lc)
                    for (int r = 0; r < (DBS ? 2 : 1) * kS7Per; ++r) slab[r * LCP + c] = OFF;
                if (MODE == 2) {
                    // separators of the CRP columns (window index j = c)
                    int v = -1;
                    if (c < No) {
                        v = use_col ? A.sepk[csep_off + c] : (kKeyInf - 1);
                        if (v < 0) {
                            misc[5] = 1;       // a flagged column: its band is decided in float64 after the ballots
                            v = (-v - 1) - 1;  // certainly-in bound below the band
                        }
                    }
                    csk_s[c] = v;
                }
            }
        }
        // The matrix is padded to NC >= Lc + 9 columns whose values are zero, and a thread walks the cyclic
        // diagonal column = (row + t) mod NC: by the time it re-enters at column 0 the window has seen 9
        // zeros, so W is back at its offset and nothing needs to be reset.
        int cb[CPT];
        const long long OFFB = __double_as_longlong(OFF);
        long long W[CPT], ring[CPT][16];  // bit patterns: W = offset + window sum, ring = the last 16 slab values
#pragma unroll
        for (int h = 0; h < CPT; ++h) {
            cb[h] = tid + NT * h;
            W[h] = OFFB;
#pragma unroll
            for (int q = 0; q < 16; ++q) ring[h][q] = OFFB;
        }
        const int nper = (Lr + kS7Per - 1) / kS7Per;
        const int est_rank = (kk >= 1 && kk < No) ? s4_est_rank(kk, No) : 0;
        __syncthreads();
        const bool any_cflag = (MODE == 2) && misc[5] != 0;
        double cn[CPT == 1 ? 4 : 1][2];  // one diagonal per thread: the seeds of this warp's 32 columns stay in registers
        if (CPT == 1) {
#pragma unroll
            for (int bi = 0; bi < (CPT == 1 ? 4 : 1); ++bi) {
                const double2 c2 = *reinterpret_cast<const double2*>(csq + 32 * warp + 8 * bi + 2 * (lane & 3));
                cn[bi][0] = c2.x;
                cn[bi][1] = c2.y;
            }
        }
        S7_LAP(0);

        // ---- stage 1 of period q: this warp's column sets x 8 rows into the slab
        auto stage1 = [&](int q) {
            const int a0 = q * kS7Per;
            const int buf = (a0 >> 5) & 1;
            if ((a0 & (kS4Stage - 1)) == 0) {
                // a new block of 32 rows: prefetch the next one into the other buffer (its readers, stage 1 of
                // the previous block, are all behind a barrier), wait for this one
                if (a0 + kS4Stage < Lr) stage_rows((a0 >> 5) + 1, buf ^ 1);
                if (buf == 0) {
                    s4_mbar_wait(&mbar[0], par0);
                    par0 ^= 1u;
                } else {
                    s4_mbar_wait(&mbar[1], par1);
                    par1 ^= 1u;
                }
            }
            double* const sl = slab + (DBS ? (q & 1) * kS7Per * LCP : 0);
            const double* xr = xs + (buf * kS4Stage + (a0 & (kS4Stage - 1)) + (lane >> 2)) * BINS + (lane & 3);
            double xa[KS];
#pragma unroll
            for (int s = 0; s < KS; ++s) xa[s] = (4 * s + (lane & 3) < BINS) ? xr[4 * s] : 0.0;
#pragma unroll
            for (int h = 0; h < CPT; ++h) {
                const int cbase = 32 * (warp + NW * h);
                if (cbase < Lc) {
                    double* dst = sl + (lane >> 2) * LCP + cbase + 2 * (lane & 3);
                    const double2* cq = reinterpret_cast<const double2*>(csq + cbase + 2 * (lane & 3));
#pragma unroll
                    for (int bi = 0; bi < 4; ++bi) {
                        double c0, c1;
                        if (CPT == 1) {
                            c0 = cn[CPT == 1 ? bi : 0][0];
                            c1 = cn[CPT == 1 ? bi : 0][1];
                        } else {
                            const double2 c2 = cq[4 * bi];
                            c0 = c2.x;
                            c1 = c2.y;
                        }
#pragma unroll
                        for (int s = 0; s < KS; ++s) s7_dmma(c0, c1, xa[s], yf[h][bi][s]);
                        double2 cc;
                        cc.x = c0;
                        cc.y = c1;
                        *reinterpret_cast<double2*>(dst + 8 * bi) = cc;
                    }
                }
            }
        };

        // ---- stage 2 of period p: walk this thread's diagonals through the slab; H = p & 1 picks the ring half
        auto stage2 = [&](auto Hc, int p) {
            constexpr int H = decltype(Hc)::value;
            const int a0 = p * kS7Per;
            const long long* const dslab = reinterpret_cast<const long long*>(slab + (DBS ? (p & 1) * kS7Per * LCP : 0));
            int* const sl = strip + (a0 % R) * NC;
            auto fixb = [](long long w) {  // bits 52..21 of the pattern (s4_fix on the double it encodes)
                return (int)__funnelshift_r((unsigned)(unsigned long long)w, (unsigned)((unsigned long long)w >> 32), 21);
            };
#pragma unroll
            for (int h = 0; h < CPT; ++h) {
                const int c = cb[h];
                if (c + kS7Per <= NC) {
                    const long long* dp = dslab + c;
                    int* sp = sl + c;
                    long long v[kS7Per];
#pragma unroll
                    for (int r = 0; r < kS7Per; ++r) v[r] = dp[r * (LCP + 1)];
                    if (a0 >= MW - 1 && c >= MW - 1 && c + kS7Per <= Lc) {
                        // inside the matrix for all 8 rows, every key valid
#pragma unroll
                        for (int r = 0; r < kS7Per; ++r) {
                            W[h] += v[r] - ring[h][(8 * H + r - MW) & 15];
                            ring[h][8 * H + r] = v[r];
                            sp[r * (NC + 1)] = fixb(W[h]);
                        }
                    } else {
                        // no wrap, but some keys are invalid (first rows / first columns / the zero pad)
#pragma unroll
                        for (int r = 0; r < kS7Per; ++r) {
                            W[h] += v[r] - ring[h][(8 * H + r - MW) & 15];
                            ring[h][8 * H + r] = v[r];
                            const bool ok = (unsigned)(c + r - (MW - 1)) < (unsigned)No && a0 + r >= MW - 1;
                            sp[r * (NC + 1)] = ok ? fixb(W[h]) : kKeyInf;
                        }
                    }
                } else {
                    // the diagonal wraps from the zero pad back to column 0 inside this period
#pragma unroll
                    for (int r = 0; r < kS7Per; ++r) {
                        int col = c + r;
                        if (col >= NC) col -= NC;
                        const long long v = dslab[r * LCP + col];
                        W[h] += v - ring[h][(8 * H + r - MW) & 15];
                        ring[h][8 * H + r] = v;
                        const bool ok = (unsigned)(col - (MW - 1)) < (unsigned)No && a0 + r >= MW - 1;
                        sl[r * NC + col] = ok ? fixb(W[h]) : kKeyInf;
                    }
                }
                cb[h] = (c + kS7Per >= NC) ? c + kS7Per - NC : c + kS7Per;
            }
        };

        if (DBS) {
            stage1(0);
            S7_LAP(1);
            __syncthreads();
            S7_LAP(3);
        }
        bool pre_done = false;  // !DBS: stage 1 of the coming period already ran (behind the selection)
        bool pre2 = false;      // DBS: stage 1 of the period after the coming one ran behind the selection too
        for (int p = 0; p < nper; ++p) {
            if (!DBS) {
                if (!pre_done) stage1(p);
                pre_done = false;
                S7_LAP(1);
                __syncthreads();  // the slab of period p is complete
                S7_LAP(3);
            }
            const bool end_strip = ((p + 1) % PPS == 0) || (p + 1 == nper);
            // Half of the warps of every scheduler (warps 4..7, 12..15, ...) run the next period's dot products BEFORE
            // their diagonal walk, the other half after it: the FP64 tensor pipe and the integer / LSU pipes are then
            // busy at the same time instead of one after the other.
            const bool s1_first = DBS && !end_strip && ((warp >> 2) & 1) && !pre2;
            if (s1_first) {
                stage1(p + 1);
                S7_LAP(1);
            }
            if (p & 1) stage2(std::integral_constant<int, 1>(), p);
            else stage2(std::integral_constant<int, 0>(), p);
            S7_LAP(2);
            if (!end_strip) {
                if (DBS && !s1_first && !pre2) stage1(p + 1);
                pre2 = false;
                S7_LAP(1);
                __syncthreads();  // keys of period p, (DBS) slab of period p + 1
                S7_LAP(3);
                continue;
            }

            // ================= the strip is complete: frame rows a_first .. a_last
            const int sp_idx = p / PPS;
            const int a_base = sp_idx * R;
            const int s_first = imax(0, (m - 1) - a_base);                // first strip row that is a window
            const int s_last = imin(R - 1, Lr - 1 - a_base);
            const int nsel = s_last - s_first + 1;
            if (tid == 0) misc[6] = 0;
            if (tid < R) {
                const int a = a_base + tid;
                rkwx[tid] = (a < Lr) ? s4_fix(rwn[a] + BIGE) : 0;
                rflag[tid] = 0;
                rthr[tid] = (kk <= 0) ? -1 : ((A.flags & (1u << 30)) ? (1 << 30) : (kKeyInf - 1));
            }
            __syncthreads();  // the keys of the strip are complete
            S7_LAP(4);

            // ---- this warp's chain of RPW consecutive rows: separator, then (MODE 2) the row's CRP words
            if (nsel > 0 && (do_select || MODE == 2)) {
                bool have_prev = false, have_prev2 = false;
                int prev_abs = 0, prev2_abs = 0;
#pragma unroll 1
                for (int q = 0; q < RPW; ++q) {
                    const int srow = warp * RPW + q;
                    if (srow < s_first || srow > s_last) continue;
                    const int* krow = strip + srow * NC;
                    const int kwx = rkwx[srow];
                    int thr = rthr[srow];
                    if (do_select) {
                        const uint4* row4 = reinterpret_cast<const uint4*>(krow);
                        S4Keys<NQ> keys;
                        keys.load_strided(krow, lane);
                        int vk, vk1, t0;
                        bool have_g = true;
                        if (have_prev) {
                            // the previous row's separator, extrapolated by half the last step (absolute key units)
                            t0 = prev_abs + (have_prev2 ? (prev_abs - prev2_abs) / 2 : 0) - kwx;
                        } else if ((A.flags & (1u << 22)) && sp_idx > 0) {
                            // A/B (debug flag): first row of the chain from the last row of the previous strip
                            // (w RPW + 1 rows back) instead of the order-statistic estimate of its own keys
                            t0 = misc[10 + ((sp_idx - 1) & 1)] - kwx;
                        } else {
                            t0 = s4_estimate(keys, est_rank, lane);   // first row of the chain: from its own keys
                            have_g = t0 < kKeyInf - 1;
                        }
                        s4_select(keys, kk, No, t0, have_g, st, lane, vk, vk1);
                        have_prev2 = have_prev;
                        prev2_abs = prev_abs;
                        const int gapk = vk1 - vk;
                        if (gapk >= 2 * G) {
                            thr = vk + (gapk >> 1);
                            prev_abs = thr + kwx;
                        } else {
                            // the two keys are closer than 2G: decide in float64
                            const int irow = a_base + srow - (m - 1);
                            double sx = 0.0;
                            int ncand = 0;
                            unsigned in_mask = 0;
                            const bool ok = s4_resolve_unsafe(row4, NQ, kk, vk, vk1, pp, irow, m, lane, cand, cand + kS4MaxFlag,
                                                              sx, ncand, in_mask);
                            if (ok) {
                                thr = vk - 3 - 1;
                                if (lane == 0) {
                                    misc[6] = 1;
                                    rflag[srow] = 1;
                                    rblo[srow] = vk - 3;
                                    rsepx[srow] = sx;
                                }
                            } else {
                                // tie cluster (the reference's own choice is undefined here): equal keys across the k-th position are
                                // all left out, like the float64 kernels do (warp_select_sep); fewer than k neighbours
                                thr = (vk1 == vk) ? vk - 1 : vk;
                            }
                            prev_abs = vk + kwx;
#if ACOSS_S7_STAT
                            if (lane == 0) atomicAdd(&g_s7_stats[1], 1ull);
#endif
                        }
                        if (lane == 0) rthr[srow] = thr;
                        if (srow == s_last && lane == 0) misc[10 + (sp_idx & 1)] = prev_abs;
                        have_prev = true;
#if ACOSS_S7_STAT
                        if (lane == 0) atomicAdd(&g_s7_stats[0], 1ull);
#endif
                    }
                    if (MODE == 2) {
                        // CRP words of the row: lane = column within the word (window j <-> frame column j + m - 1)
                        unsigned mine = 0;
#pragma unroll
                        for (int w = 0; w < WPR; ++w) {
                            if (w < wp) {
                                const int j = 32 * w + lane;
                                const int key = (j + (MW - 1) < NC) ? krow[j + (MW - 1)] : kKeyInf;
                                const unsigned bw = __ballot_sync(kFullMask, key <= imin(thr, csk_s[j] - kwx));
                                if (lane == w) mine = bw;
                            }
                        }
                        if (lane < wp) bits[srow * wp + lane] = mine;
                    }
                }
            }
            S7_LAP(5);
            // early finishers start the next period's dot products while the others still select
            if (p + 1 < nper && !pre2) {
                stage1(p + 1);
                pre_done = true;
                // (not when period p + 2 starts a new block of staged rows: its prefetch would overwrite the rows that
                // slower warps still need for period p + 1)
                if (DBS && p + 2 < nper && ((p + 2) % PPS != 0) && (((p + 2) * kS7Per) & (kS4Stage - 1)) != 0) {
                    stage1(p + 2);  // both slabs are free behind the strip's last diagonal walk
                    pre2 = true;
                }
            }
            S7_LAP(1);
            __syncthreads();  // every row of the strip has its separator (MODE 2: and its CRP words)
            S7_LAP(6);

            if (MODE == 0) {
                if (tid >= s_first && tid <= s_last) {
                    const int irow = a_base + tid - (m - 1);
                    const int thr = rthr[tid], kwx = rkwx[tid];
                    if (rflag[tid]) {
                        A.sepk[rsep_off + irow] = -(rblo[tid] + kwx) - 1;
                        A.sepx[rsep_off + irow] = rsepx[tid];
                    } else {
                        // kk <= 0 -> never in; kk >= No -> everything valid
                        A.sepk[rsep_off + irow] = (thr < 0) ? 0 : ((thr >= kKeyInf - 1) ? kKeyInf - 1 : thr + kwx);
                    }
                }
                S7_LAP(7);
            } else if (nsel > 0) {
                S7_LAP(7);
                const bool fix = any_cflag || misc[6] != 0;  // a flagged column of the item or a flagged row of the strip
                if (fix) {
                    // cells of flagged rows / flagged columns: recompute the decision with the float64 bands
                    for (int pass = 0; pass < 2; ++pass) {
#pragma unroll
                        for (int h = 0; h < CPT; ++h) {
                            const int j = tid + NT * h;
                            if (j >= No) continue;
                            const int craw = use_col ? A.sepk[csep_off + j] : (kKeyInf - 1);
                            const bool cflag = craw < 0;
                            if (pass == 1 && !cflag) continue;
                            const int cblo = -craw - 1;
                            const double csepx = cflag ? A.sepx[csep_off + j] : 0.0;
                            for (int s = s_first; s <= s_last; ++s) {
                                if (pass == 0 && (!rflag[s] || cflag)) continue;
                                const int key = strip[s * NC + j + (MW - 1)];
                                if (key == kKeyInf) continue;
                                const int irow = a_base + s - (m - 1);
                                const int kabs = key + rkwx[s];
                                double zx = 0.0;
                                bool have_z = false;
                                bool colin, rowin;
                                if (cflag) {
                                    if (kabs < cblo) colin = true;
                                    else if (kabs > cblo + BANDW) colin = false;
                                    else {
                                        zx = orient ? s4_cell_exact(pp.fi, pp.fj, BINS, pp.oti, m, j, irow)
                                                    : s4_cell_exact(pp.fi, pp.fj, BINS, pp.oti, m, irow, j);
                                        have_z = true;
                                        colin = zx <= csepx;
                                    }
                                } else {
                                    colin = kabs <= craw;
                                }
                                if (rflag[s]) {
                                    if (key < rblo[s]) rowin = true;
                                    else if (key > rblo[s] + BANDW) rowin = false;
                                    else {
                                        if (!have_z)
                                            zx = orient ? s4_cell_exact(pp.fi, pp.fj, BINS, pp.oti, m, j, irow)
                                                        : s4_cell_exact(pp.fi, pp.fj, BINS, pp.oti, m, irow, j);
                                        rowin = zx <= rsepx[s];
                                    }
                                } else {
                                    rowin = key <= rthr[s];
                                }
                                uint32_t* wd = bits + s * wp + (j >> 5);
                                const uint32_t bit = 1u << (j & 31);
                                if (colin && rowin) atomicOr(wd, bit);
                                else atomicAnd(wd, ~bit);
                            }
                        }
                        __syncthreads();
                    }
                }
                S7_LAP(8);
                // ---- the strip's rows are one contiguous run of words in the packed CRP
                {
                    uint32_t* dst = A.crp + pwp->crp_off + (int64_t)(a_base + s_first - (m - 1)) * wp;
                    const uint32_t* src = bits + s_first * wp;
                    const int nwords = nsel * wp;
                    for (int e = tid; e < nwords; e += NT) dst[e] = src[e];
                }
                S7_LAP(9);
            }
            __syncthreads();  // the strip, the row records and the CRP words are free again
            S7_LAP(10);
#if ACOSS_S7_STAT
            if (tid == 0) atomicAdd(&g_s7_stats[2], 1ull);
#endif
        }
#if ACOSS_S7_STAT
        if (tid < 12) {
            atomicAdd(&g_s7_stats[4 + tid], (unsigned long long)(unsigned)s_prof[tid]);
            s_prof[tid] = 0;
        }
        if (tid == 0) atomicAdd(&g_s7_stats[3], 1ull);
#endif
    }
}

}  // namespace acoss
