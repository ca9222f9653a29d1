// kernels_sweep7.cuh -- the cross-similarity sweep, seventh generation: the default for 12 / 13
// bins, m = 9 and column songs of up to 1015 frames.
//
// Distances, keys, separators and the float64 repair of ambiguous rows are those of
// kernels_sweep4.cuh / _sweep6.cuh (get_csm + sliding_csm + csm_to_binary[_mutual],
// CRPUtils.py:24-45,67-84,169-219).  What changed after the round-1 captures (profiles/r01_g: one
// row's selection = ~800 dependent warp instructions, 6 200 of the 12 200 cycles of a 16-row
// period, barrier stalls 25-42 %):
//
//   * SELECTION IN LANE GROUPS.  A strip of R = 32..64 rows of keys is kept in shared memory and
//     a row is resolved by a group of LPR = 8 (16) lanes, four (two) rows per warp in ONE
//     instruction stream: every lane streams NC / LPR keys of its row per counting pass
//     (LDS.128 + 2 integer instructions per key, no dependency between keys), the partial counts
//     meet in log2(LPR) shuffles, and the threshold search (bracket + interpolation, the state
//     is replicated in the group's lanes) runs once per pass for four rows.  Groups fetch rows
//     from a shared counter, so a row that needs more passes holds up nobody.  The last pass of
//     a row also tracks the largest key below and the smallest key above the threshold, which
//     are the k-th and (k+1)-th keys: no candidate walk.
//   * 8-ROW PERIODS.  The DMMA slab is 8 rows (one m8n8k4 tile row) and double buffered where
//     shared memory allows, which leaves room for the strip.
//   * PACKED CRP FROM THE SWEEP.  MODE 2 turns a strip into CRP words with one __ballot_sync per
//     (row, 32 columns) -- thread = CRP column -- and writes them straight into the layout the
//     DP kernel reads.  The nibble-byte arena and k_nb_to_packed of gens 4-6 are gone: a CRP
//     leaves the SM once, as 1 bit per cell.
//   * 13 bins (the mfcc half of Serra09.similarity, Serra09.py:177-184) run on the same kernel:
//     K is padded to 16 for the DMMA fragments.
//   * Column capacities 128 / 256 / 512 (one diagonal per thread) and 768 / 1024 (two).
#pragma once
#include <type_traits>

#include "kernels_sweep4.cuh"

namespace acoss {

constexpr int kS7Pad = 9;      // NC >= Lc + kS7Pad zero columns flush the window (m <= 9)
constexpr int kS7Per = 8;      // rows per period (one DMMA tile row)
constexpr int kS7FullW = 6;    // bracket width (ranks) from which the passes also track the boundary keys

#if defined(ACOSS_S7_STATS) && !defined(ACOSS_EMU)
#define ACOSS_S7_STAT 1
__device__ unsigned long long g_s7_stats[8];  // [0] rows, [1] count passes, [2] full passes, [3] unsafe rows
#else
#define ACOSS_S7_STAT 0
#endif

template <int NT_, int CPT_>
struct S7Cfg {
    static constexpr int NT = NT_, CPT = CPT_;
    static constexpr int NC = NT * CPT;                    // padded columns
    static constexpr int NW = NT / 32;
    static constexpr int LPR = NC >= 1024 ? 16 : 8;        // lanes per row in the selection
    static constexpr int NQ = NC / (4 * LPR);              // uint4 per lane and pass
    static constexpr int R = NC <= 512 ? 64 : (NC <= 768 ? 48 : 32);  // rows per strip
    static constexpr bool DBS = NC <= 512;                 // double-buffered slab
    static constexpr int LCP = NC + 8;                     // slab row stride (doubles), = 8 mod 16
    static constexpr int WPR = NC / 32;                    // CRP words per strip row
    static_assert(NT % 32 == 0 && NC % 128 == 0 && R % (2 * kS7Per) == 0, "shape");
};

__host__ __device__ constexpr int s7_capacity_threads(int cap) {  // column capacity -> threads
    return cap <= 512 ? cap : cap / 2;
}
__host__ __device__ constexpr int s7_capacity_of(int Lc) {
    return Lc + kS7Pad <= 128 ? 128
         : Lc + kS7Pad <= 256 ? 256
         : Lc + kS7Pad <= 512 ? 512
         : Lc + kS7Pad <= 768 ? 768
         : Lc + kS7Pad <= 1024 ? 1024 : 0;
}
constexpr int kS7Buckets = 5;
__host__ __device__ constexpr int s7_bucket_of(int cap) {
    return cap == 128 ? 0 : cap == 256 ? 1 : cap == 512 ? 2 : cap == 768 ? 3 : 4;
}
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
    L.rinfo = o; o += 6 * C::R * 4;                      // thr, kwx, vk, vk1, blo, flag
    L.cand = o;  o += C::NW * (kS4MaxFlag + 1) * 4;      // per warp: candidate columns + their count
    L.misc = o;  o += (16 + C::R) * 4;                   // counters + the list of unsafe rows
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

// ---------------------------------------------------------------- selection
// One pass of a lane over its share of a strip row: the number of keys > t and (FULL) the smallest
// t - key >= 0 and the largest t - key < 0 as unsigned numbers, combined over the group's LPR lanes.
template <int LPR, int NQ, bool FULL>
__device__ __forceinline__ void s7_pass(const uint4* __restrict__ rq, int t, unsigned& gt, unsigned& dmin, unsigned& dmax) {
    unsigned s0 = 0, s1 = 0, s2 = 0, s3 = 0;
    unsigned mn = 0xffffffffu, mx = 0u;
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        const uint4 v = rq[q * LPR];
        const unsigned d0 = (unsigned)t - v.x, d1 = (unsigned)t - v.y, d2 = (unsigned)t - v.z, d3 = (unsigned)t - v.w;
        s0 += d0 >> 31;
        s1 += d1 >> 31;
        s2 += d2 >> 31;
        s3 += d3 >> 31;
        if (FULL) {
            mn = s4_umin(mn, s4_umin(s4_umin(d0, d1), s4_umin(d2, d3)));
            mx = s4_umax(mx, s4_umax(s4_umax(d0, d1), s4_umax(d2, d3)));
        }
    }
    unsigned s = (s0 + s1) + (s2 + s3);
#pragma unroll
    for (int o = 1; o < LPR; o <<= 1) {
        s += __shfl_xor_sync(kFullMask, s, o);
        if (FULL) {
            mn = s4_umin(mn, __shfl_xor_sync(kFullMask, mn, o));
            mx = s4_umax(mx, __shfl_xor_sync(kFullMask, mx, o));
        }
    }
    gt = s;
    dmin = mn;
    dmax = mx;
}

// Threshold search of one row; every lane of the group holds the same copy.
struct S7Row {
    int row;       // strip row being resolved, -1: none
    int t;         // threshold of the next pass
    int lo, clo;   // count(lo) = clo < k
    int hi, chi;   // count(hi) = chi >= k
    int tA, cA;    // previous probe (key spacing estimate)
    int iter, gallop;
    float spacing;
    bool rl, rh, haveA, wantfull;
};

__device__ __forceinline__ void s7_row_begin(S7Row& S, int row, int t0) {
    S.row = row;
    S.t = t0 < 0 ? 0 : (t0 > kKeyInf - 2 ? kKeyInf - 2 : t0);
    S.lo = -1;
    S.clo = 0;
    S.hi = kKeyInf - 1;
    S.chi = 0x7fffffff;
    S.tA = S.cA = 0;
    S.iter = S.gallop = 0;
    S.rl = S.rh = S.haveA = S.wantfull = false;
}

// Digest one pass (count c of keys <= S.t; on a full pass `below` = largest key <= t, `above` = smallest
// key > t or kKeyInf).  Returns true when the k-th and (k+1)-th keys are known.
__device__ __forceinline__ bool s7_row_step(S7Row& S, int k, int c, bool full, int below, int above, int& vk, int& vk1) {
    const int t = S.t;
    if (S.haveA && c != S.cA) {
        const float sp = fabsf((float)(t - S.tA)) * __fdividef(1.0f, fabsf((float)(c - S.cA)));
        if (sp >= 1.0f) S.spacing = sp;
    }
    S.tA = t;
    S.cA = c;
    S.haveA = true;
    ++S.iter;
    const int e = k - c;
    if (e > 0) {
        S.lo = t; S.clo = c; S.rl = true;
    } else {
        S.hi = t; S.chi = c; S.rh = true;
    }
    if (e == 0) {
        if (full) {
            vk = below;
            vk1 = above;
            return true;
        }
        S.wantfull = true;  // same threshold again, with the boundary keys
        return false;
    }
    if (full) {
        if (e < 0 && S.rl && S.lo + 1 == t) {
            // the count jumps across k between two adjacent integers: ranks k and k + 1 share key t
            vk = t;
            vk1 = t;
            return true;
        }
        if (e == 1 && above < kKeyInf) {
            // `above` IS the k-th key; one more full pass at it yields the (k+1)-th
            S.t = above;
            S.wantfull = true;
            return false;
        }
        if (e == -1) {
            // `below` is the (k+1)-th key
            if (S.rl && below - 1 <= S.lo) {
                vk = below;
                vk1 = below;
                return true;
            }
            S.t = below - 1;
            S.wantfull = true;
            return false;
        }
    }
    int nt;
    if (S.rl && S.rh) {
        const int width = S.hi - S.lo;
        if (width <= 1) {
            nt = S.hi;  // count(hi) >= k > count(hi - 1): a full pass at hi ends the search
            S.wantfull = true;
        } else {
            if (S.iter > 40 || (S.iter % 3) == 0) {
                nt = S.lo + (width >> 1);
            } else {
                const float fr = __fdividef((float)(k - S.clo), (float)(S.chi - S.clo));
                nt = S.lo + (int)((float)width * fr);
            }
            if (nt <= S.lo) nt = S.lo + 1;
            if (nt >= S.hi) nt = S.hi - 1;
            if (S.chi - S.clo <= kS7FullW) S.wantfull = true;
        }
    } else {
        float stepf = S.spacing * 1.25f * fabsf((float)e) * (float)(1 << (S.gallop < 20 ? S.gallop : 20));
        if (stepf < 1.0f) stepf = 1.0f;
        if (stepf > 1.0e9f) stepf = 1.0e9f;
        const int step = (int)stepf;
        ++S.gallop;
        int64_t n64 = (e > 0) ? (int64_t)t + step : (int64_t)t - step;
        if (n64 < 0) n64 = 0;
        if (n64 > kKeyInf - 2) n64 = kKeyInf - 2;
        if (S.rh && n64 >= S.hi) n64 = S.hi - 1;
        if (S.rl && n64 <= S.lo) n64 = S.lo + 1;
        if ((int)n64 == t) {
            // pinned against the edge of the key domain: the bracket is real on that side
            if (e > 0) { S.hi = kKeyInf - 1; S.chi = 0x7fffffff; S.rh = true; }
            else { S.lo = -1; S.clo = 0; S.rl = true; }
            n64 = S.lo + ((S.hi - S.lo) >> 1);
            if (n64 <= S.lo) n64 = S.lo + 1;
        }
        nt = (int)n64;
        if (fabsf((float)e) <= (float)kS7FullW) S.wantfull = true;
    }
    S.t = nt;
    return false;
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
__global__ void __launch_bounds__(NT, (NT * CPT <= 128 ? 4 : (NT * CPT <= 256 ? 2 : 1))) k_sweep7(Sweep7Args A) {
    using C = S7Cfg<NT, CPT>;
    constexpr int NC = C::NC, NW = C::NW, LPR = C::LPR, NQ = C::NQ, R = C::R, LCP = C::LCP, WPR = C::WPR;
    constexpr bool DBS = C::DBS;
    constexpr int KS = (BINS + 3) / 4;  // DMMA k-steps
    constexpr int PPS = R / kS7Per;     // periods per strip
    static_assert(MW <= kS7Pad && MW <= 15, "the zero pad must flush the window");
    ACOSS_DYNSMEM(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int sub = lane & (LPR - 1), glead = lane & ~(LPR - 1);
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
    int* const rvk = rthr + 2 * R;
    int* const rvk1 = rthr + 3 * R;
    int* const rblo = rthr + 4 * R;
    int* const rflag = rthr + 5 * R;                                     // 1: flagged (float64 separator in rsepx)
    int* const cand = reinterpret_cast<int*>(smem_raw + LY.cand) + warp * (kS4MaxFlag + 1);
    int* const misc = reinterpret_cast<int*>(smem_raw + LY.misc);
    int* const s_item = misc;        // [0] item, [1] next row of the strip, [2] unsafe rows, [3] guess valid,
    int* const s_next = misc + 1;    // [4] guess (absolute key), [5] any flagged column
    int* const s_nunsafe = misc + 2;
    int* const s_unsafe = misc + 16;
    uint64_t* const mbar = reinterpret_cast<uint64_t*>(smem_raw + LY.mbar);

    if (tid == 0) {
        s4_mbar_init(&mbar[0], 1);
        s4_mbar_init(&mbar[1], 1);
        s4_fence_mbar_init();
    }
    __syncthreads();
    unsigned par0 = 0, par1 = 0;
    float spacing_keep = 4096.0f;

    for (;;) {
        if (tid == 0) {
            *s_item = (int)atomicAdd(A.counter, 1u);
            misc[3] = 0;
            misc[5] = 0;
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
                csq[c] = (c < Lc) ? cnorm[c] : 0.0;
                // column sets that hold no column of the song are never written by stage 1: they are the zero pad
                if (32 * (warp + NW * h) >= Lc)
                    for (int r = 0; r < (DBS ? 2 : 1) * kS7Per; ++r) slab[r * LCP + c] = 0.0;
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
        double W[CPT], ring[CPT][16];
#pragma unroll
        for (int h = 0; h < CPT; ++h) {
            cb[h] = tid + NT * h;
            W[h] = OFF;
#pragma unroll
            for (int q = 0; q < 16; ++q) ring[h][q] = 0.0;
        }
        const int nper = (Lr + kS7Per - 1) / kS7Per;
        __syncthreads();
        const bool any_cflag = (MODE == 2) && misc[5] != 0;

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
                        const double2 c2 = cq[4 * bi];
                        double c0 = c2.x, c1 = c2.y;
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
            const double* const dslab = slab + (DBS ? (p & 1) * kS7Per * LCP : 0);
            int* const sl = strip + (a0 % R) * NC;
#pragma unroll
            for (int h = 0; h < CPT; ++h) {
                const int c = cb[h];
                if (c + kS7Per <= NC) {
                    const double* dp = dslab + c;
                    int* sp = sl + c;
                    double v[kS7Per];
#pragma unroll
                    for (int r = 0; r < kS7Per; ++r) v[r] = dp[r * (LCP + 1)];
                    if (a0 >= MW - 1 && c >= MW - 1 && c + kS7Per <= Lc) {
                        // inside the matrix for all 8 rows, every key valid
#pragma unroll
                        for (int r = 0; r < kS7Per; ++r) {
                            W[h] += v[r] - ring[h][(8 * H + r - MW) & 15];
                            ring[h][8 * H + r] = v[r];
                            sp[r * (NC + 1)] = s4_fix(W[h]);
                        }
                    } else {
                        // no wrap, but some keys are invalid (first rows / first columns / the zero pad)
#pragma unroll
                        for (int r = 0; r < kS7Per; ++r) {
                            W[h] += v[r] - ring[h][(8 * H + r - MW) & 15];
                            ring[h][8 * H + r] = v[r];
                            const bool ok = (unsigned)(c + r - (MW - 1)) < (unsigned)No && a0 + r >= MW - 1;
                            sp[r * (NC + 1)] = ok ? s4_fix(W[h]) : kKeyInf;
                        }
                    }
                } else {
                    // the diagonal wraps from the zero pad back to column 0 inside this period
#pragma unroll
                    for (int r = 0; r < kS7Per; ++r) {
                        int col = c + r;
                        if (col >= NC) col -= NC;
                        const double v = dslab[r * LCP + col];
                        W[h] += v - ring[h][(8 * H + r - MW) & 15];
                        ring[h][8 * H + r] = v;
                        const bool ok = (unsigned)(col - (MW - 1)) < (unsigned)No && a0 + r >= MW - 1;
                        sl[r * NC + col] = ok ? s4_fix(W[h]) : kKeyInf;
                    }
                }
                cb[h] = (c + kS7Per >= NC) ? c + kS7Per - NC : c + kS7Per;
            }
        };

        if (DBS) {
            stage1(0);
            __syncthreads();
        }
        bool pre_done = false;  // !DBS: stage 1 of the coming period already ran (behind the selection)
        for (int p = 0; p < nper; ++p) {
            if (!DBS) {
                if (!pre_done) stage1(p);
                pre_done = false;
                __syncthreads();  // the slab of period p is complete
            }
            if (p & 1) stage2(std::integral_constant<int, 1>(), p);
            else stage2(std::integral_constant<int, 0>(), p);
            const bool end_strip = ((p + 1) % PPS == 0) || (p + 1 == nper);
            if (!end_strip) {
                if (DBS) stage1(p + 1);
                __syncthreads();  // keys of period p, (DBS) slab of period p + 1
                continue;
            }

            // ================= the strip is complete: frame rows a_first .. a_last
            const int sp_idx = p / PPS;
            const int a_base = sp_idx * R;
            const int s_first = imax(0, (m - 1) - a_base);                // first strip row that is a window
            const int s_last = imin(R - 1, Lr - 1 - a_base);
            const int nsel = s_last - s_first + 1;
            if (tid == 0) {
                *s_next = 0;
                *s_nunsafe = 0;
            }
            if (tid < R) {
                const int a = a_base + tid;
                rkwx[tid] = (a < Lr) ? s4_fix(rwn[a] + BIGE) : 0;
                rflag[tid] = 0;
                rthr[tid] = (kk <= 0) ? -1 : ((A.flags & (1u << 30)) ? (1 << 30) : (kKeyInf - 1));
            }
            __syncthreads();  // the keys of the strip are complete

            if (do_select && nsel > 0) {
                S7Row S;
                S.row = -1;
                S.spacing = spacing_keep;
                bool exhausted = false;
#if ACOSS_S7_STAT
                unsigned n_cnt = 0, n_full = 0, n_rows = 0;
#endif
                for (;;) {
                    // idle groups fetch the next row of the strip
                    const bool need = S.row < 0 && !exhausted;
                    int r = 0;
                    if (need && sub == 0) r = atomicAdd(s_next, 1);
                    r = __shfl_sync(kFullMask, r, glead);
                    if (need) {
                        if (r < nsel) {
                            const int srow = s_first + r;
                            const volatile int* vm = misc;
                            const int t0 = vm[3] ? (vm[4] - rkwx[srow]) : (1 << 30);
                            s7_row_begin(S, srow, t0);
                        } else {
                            exhausted = true;
                        }
                    }
                    if (__all_sync(kFullMask, S.row < 0)) break;
                    const int srow = S.row < 0 ? s_first : S.row;
                    const uint4* rq = reinterpret_cast<const uint4*>(strip + srow * NC) + sub;
                    const bool full = __any_sync(kFullMask, S.row >= 0 && S.wantfull);
                    unsigned gt, dmin, dmax;
                    if (full) s7_pass<LPR, NQ, true>(rq, S.t, gt, dmin, dmax);
                    else s7_pass<LPR, NQ, false>(rq, S.t, gt, dmin, dmax);
#if ACOSS_S7_STAT
                    if (S.row >= 0) { if (full) ++n_full; else ++n_cnt; }
#endif
                    if (S.row >= 0) {
                        const int c = NC - (int)gt;
                        const int below = S.t - (int)dmin;
                        const int above = (dmax >= 0x80000000u) ? (int)((unsigned)S.t - dmax) : kKeyInf;
                        int vk, vk1;
                        if (s7_row_step(S, kk, c, full, below, above, vk, vk1)) {
                            if (sub == 0) {
                                const int gapk = vk1 - vk;
                                if (gapk >= 2 * G) {
                                    const int thr = vk + (gapk >> 1);
                                    rthr[S.row] = thr;
                                    volatile int* vm = misc;
                                    vm[4] = thr + rkwx[S.row];
                                    vm[3] = 1;
                                } else {
                                    rvk[S.row] = vk;
                                    rvk1[S.row] = vk1;
                                    s_unsafe[atomicAdd(s_nunsafe, 1)] = S.row;
                                }
                            }
#if ACOSS_S7_STAT
                            ++n_rows;
#endif
                            S.row = -1;
                        }
                    }
                }
                spacing_keep = S.spacing;
#if ACOSS_S7_STAT
                if (sub == 0) {
                    atomicAdd(&g_s7_stats[0], (unsigned long long)n_rows);
                    atomicAdd(&g_s7_stats[1], (unsigned long long)n_cnt);
                    atomicAdd(&g_s7_stats[2], (unsigned long long)n_full);
                }
#endif
            }
            // early finishers start the next period's dot products while the others still select
            if (p + 1 < nper) {
                stage1(p + 1);
                pre_done = true;
            }
            __syncthreads();  // every row of the strip has its k-th / (k+1)-th keys

            // ---- unsafe rows: the two keys are closer than 2G, decide in float64 (one warp per row)
            const int nunsafe = *s_nunsafe;
            if (nunsafe > 0) {
                for (int u = warp; u < nunsafe; u += NW) {
                    const int srow = s_unsafe[u];
                    const int vk = rvk[srow], vk1 = rvk1[srow];
                    const int irow = a_base + srow - (m - 1);
                    double sx = 0.0;
                    int ncand = 0;
                    unsigned in_mask = 0;
                    const bool ok = s4_resolve_unsafe(reinterpret_cast<const uint4*>(strip + srow * NC), NC / 128, kk, vk, vk1,
                                                      pp, irow, m, lane, cand, cand + kS4MaxFlag, sx, ncand, in_mask);
                    if (lane == 0) {
                        if (ok) {
                            rflag[srow] = 1;
                            rblo[srow] = vk - 3;
                            rsepx[srow] = sx;
                            rthr[srow] = vk - 3 - 1;
                        } else {
                            rthr[srow] = vk;  // tie cluster: the reference's own choice is undefined here
                        }
#if ACOSS_S7_STAT
                        atomicAdd(&g_s7_stats[3], 1ull);
#endif
                    }
                    __syncwarp();
                }
                __syncthreads();
            }

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
            } else if (nsel > 0) {
                // ---- CRP words: thread = CRP column (window j = frame column j + m - 1), one ballot per row
#pragma unroll
                for (int h = 0; h < CPT; ++h) {
                    const int j = tid + NT * h;
                    const int wi = j >> 5;
                    if (wi < wp) {  // warp-uniform
                        const int cs = csk_s[j];
                        const int* kp = strip + s_first * NC + j + (MW - 1);
                        const bool in_strip = j + (MW - 1) < NC;
                        for (int s = s_first; s <= s_last; ++s, kp += NC) {
                            const int key = in_strip ? *kp : kKeyInf;
                            const int bound = imin(rthr[s], cs - rkwx[s]);
                            const unsigned w = __ballot_sync(kFullMask, key <= bound);
                            if (lane == 0) bits[s * wp + wi] = w;
                        }
                    }
                }
                bool fix = false;
                if (any_cflag) fix = true;
                else
                    for (int s = s_first; s <= s_last; ++s) fix = fix || rflag[s] != 0;  // smem broadcast, uniform
                if (fix) {
                    __syncthreads();
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
                } else {
                    __syncthreads();
                }
                // ---- the strip's rows are one contiguous run of words in the packed CRP
                {
                    uint32_t* dst = A.crp + pwp->crp_off + (int64_t)(a_base + s_first - (m - 1)) * wp;
                    const uint32_t* src = bits + s_first * wp;
                    const int nwords = nsel * wp;
                    for (int e = tid; e < nwords; e += NT) dst[e] = src[e];
                }
            }
            __syncthreads();  // the strip, the row records and the CRP words are free again
        }
    }
}

}  // namespace acoss
