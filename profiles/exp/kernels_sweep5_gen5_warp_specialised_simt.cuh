// kernels_sweep5.cuh -- the cross-similarity sweep, warp-specialised (the default for 12 bins,
// m = 9 and column songs of up to 8 x 120 frames).
//
// Arithmetic, keys, separators and the float64 fallback are those of kernels_sweep4.cuh; what
// changes is WHO does what.  The gen-4 CTA alternates between a compute phase (FP64 bound) and a
// selection phase (a chain of warp-wide reductions, latency bound): ncu showed the FP64 pipe idle
// while every warp of the CTA waited on REDUX results (profiles/r01_e_*).  Here a CTA has
//   * 4 NWG COMPUTE warps: each sweeps its own block of 120 columns (plus an 8-column halo, see
//     s4_halo_lanes) through all rows, never waiting for another compute warp, and drops its keys
//     into a double-buffered strip of 4 NWG rows;
//   * 4 NWG SELECTION warps: each resolves one row of a full strip (s4_resolve_row) while the
//     compute warps are already filling the other strip buffer.
// Hand-over is by named barriers (bar.arrive / bar.sync, FULL[b] and EMPTY[b]).  The two roles need
// very different register counts, so the kernel is launched at 128 registers per thread and the
// warpgroups re-balance with setmaxnreg (compute 168, selection 88): 2 CTAs of 256 threads per SM
// (NWG = 1) or 1 CTA of 512 (NWG = 2).
#pragma once
#include "kernels_sweep4.cuh"

namespace acoss {

constexpr int kS5BarFull = 1;     // named barriers 1, 2: strip buffer b is full
constexpr int kS5BarEmpty = 3;    // 3, 4: strip buffer b may be overwritten
constexpr int kS5BarCompute = 5;  // compute warps only (staging-buffer reuse)

struct S5Layout {
    int xs, ring, fsepx, strip, csk, fcol, fblo, cand, misc, mbar, total;
};
__host__ __device__ inline S5Layout s5_layout(int NC, int m, int bins) {
    S5Layout L;
    int o = 0;
    L.xs = o;    o += 2 * kS4Stage * bins * 8;
    L.ring = o;  o += NC * (m + 1) * 4 * (32 + s4_padt(m)) * 8;
    L.fsepx = o; o += kS4MaxFlag * 8;
    L.strip = o; o += 2 * NC * (32 * NC) * 16;   // 2 buffers x NC rows x 32 NC quads
    L.csk = o;   o += 4 * 32 * NC * 4;
    L.fcol = o;  o += kS4MaxFlag * 4;
    L.fblo = o;  o += kS4MaxFlag * 4;
    L.cand = o;  o += NC * 2 * kS4MaxFlag * 4;
    L.misc = o;  o += (16 + NC) * 4;
    o = (o + 7) & ~7;
    L.mbar = o;  o += 2 * 8;
    L.total = (o + 15) & ~15;
    return L;
}

__device__ __forceinline__ void s5_regs_compute() {
#ifndef ACOSS_EMU
    asm volatile("setmaxnreg.inc.sync.aligned.u32 168;");
#endif
}
__device__ __forceinline__ void s5_regs_select() {
#ifndef ACOSS_EMU
    asm volatile("setmaxnreg.dec.sync.aligned.u32 88;");
#endif
}

template <int BINS, int MODE, int MW, int NWG>
__global__ void __launch_bounds__(256 * NWG, (NWG == 1 ? 2 : 1)) k_sweep5(Sweep4Args A) {
    constexpr int CPL = 4;
    constexpr int NC = 4 * NWG;      // compute warps = selection warps = rows of a strip
    constexpr int TC = 32 * NC;      // compute threads = quads of a strip row
    constexpr int NQS = (NWG == 1) ? NC : 0;  // selection keeps the keys in registers for the small CTA
    ACOSS_DYNSMEM(smem_raw);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool is_compute = warp < NC;
    const int m = (MW > 0) ? MW : A.m;
    const int RD = m + 1;
    const int PADT = s4_padt(m), TWP = 32 + PADT, RS = 4 * TWP;
    const int HL = s4_halo_lanes(m), OWN = 32 - HL;
    const int G = A.gap, BANDW = 2 * A.gap + 6;
    const S5Layout LY = s5_layout(NC, m, BINS);

    double* const xs = reinterpret_cast<double*>(smem_raw + LY.xs);
    double* const ring = reinterpret_cast<double*>(smem_raw + LY.ring);
    double* const fsepx = reinterpret_cast<double*>(smem_raw + LY.fsepx);
    uint4* const strip4 = reinterpret_cast<uint4*>(smem_raw + LY.strip);  // [2][NC][TC]
    int* const csk_s = reinterpret_cast<int*>(smem_raw + LY.csk);         // [4 TC]
    int* const fcol = reinterpret_cast<int*>(smem_raw + LY.fcol);
    int* const fblo = reinterpret_cast<int*>(smem_raw + LY.fblo);
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

// Both roles run the same item loop (the two CTA-wide barriers of the fetch, barrier (A) after the
// set-up and barrier (B) at the end of an item are executed by every warp), but each role's loop
// is its own code region so that it is compiled for the register count set by setmaxnreg.
#define ACOSS_S5_FETCH_ITEM()                                                   \
    if (tid == 0) {                                                             \
        *s_item = (int)atomicAdd(A.counter, 1u);                                \
        *s_nflag = 0;                                                           \
        misc[8] = misc[9] = misc[10] = 0;                                       \
    }                                                                           \
    __syncthreads();                                                            \
    const int it = *s_item;                                                     \
    __syncthreads();                                                            \
    if (it >= A.n_items) break;                                                 \
    const uint32_t item = A.items[it];                                          \
    int pidx, orient;                                                           \
    if (MODE == 0) {                                                            \
        pidx = (int)(item >> 1);                                                \
        orient = (int)(item & 1u);                                              \
    } else {                                                                    \
        pidx = (int)item;                                                       \
        orient = A.pairs[pidx].swap;                                            \
    }                                                                           \
    const PairWork* const pwp = A.pairs + pidx;                                 \
    const int song_i = pwp->song_i, song_j = pwp->song_j;                       \
    const int rsong = orient ? song_j : song_i;                                 \
    const int csong = orient ? song_i : song_j;                                 \
    const int Lr = orient ? pwp->Lj : pwp->Li;                                  \
    const int Lc = orient ? pwp->Li : pwp->Lj;                                  \
    const int Mo = Lr - m + 1, No = Lc - m + 1;                                 \
    if (Mo <= 0 || No <= 0) continue;                                           \
    const int nstrips = (Mo + NC - 1) / NC;                                     \
    const double BIGE = ldexp(1.0, pwp->expo);                                  \
    (void)song_i; (void)song_j; (void)rsong; (void)csong; (void)Lc; (void)nstrips; (void)No

    if (is_compute) {
        // =============================================================== compute warps
        s5_regs_compute();
        unsigned par0 = 0, par1 = 0;  // phase parity of the two staging barriers
        for (;;) {
            ACOSS_S5_FETCH_ITEM();
            const bool own = lane >= HL;
            const int sidx = warp * OWN + lane - HL;
            const bool active = warp * OWN * CPL < Lc + 0;  // this warp's block starts inside the song
            const double* rframes = A.frames + A.offs[rsong] * BINS;
            const double* cframes = A.frames + A.offs[csong] * BINS;
            const double* cwn = A.wnorm + A.offs[csong];
            const double OFF = 1.5 * BIGE;
            const int oti = pwp->oti;
            double* const ring_w = ring + warp * RD * RS;
            double* const ring_t = ring_w + lane + PADT;
            int pooff[CPL];
#pragma unroll
            for (int r = 0; r < CPL; ++r) {
                const int d = r - m;
                const int q = (d >= 0) ? (d >> 2) : -((-d + 3) >> 2);
                pooff[r] = q + (d - 4 * q) * TWP;
            }
            if (tid == 0) {
                const unsigned n0 = (unsigned)imin(kS4Stage, Lr) * BINS * 8u;
                s4_bulk_load(xs, rframes, n0, &mbar[0]);
            }
            for (int p = lane; p < RD * RS; p += 32) ring_w[p] = 0.0;
            const int crot = orient ? (BINS - oti) % BINS : oti;
            double y2[CPL][BINS], Pprev[CPL];
            int kwy[CPL];
            unsigned vmask = 0;
#pragma unroll
            for (int r = 0; r < CPL; ++r) {
                const int b = (warp * OWN + lane - HL) * CPL + r;
                Pprev[r] = 0.0;
                kwy[r] = 0;
                if (b >= 0 && b < Lc) {
#pragma unroll
                    for (int k = 0; k < BINS; ++k) {
                        int src = k + crot;
                        if (src >= BINS) src -= BINS;
                        y2[r][k] = -2.0 * cframes[(int64_t)b * BINS + src];
                    }
                } else {
#pragma unroll
                    for (int k = 0; k < BINS; ++k) y2[r][k] = 0.0;
                }
                const int j = b - (m - 1);
                if (own && j >= 0 && j < No) {
                    vmask |= 1u << r;
                    kwy[r] = s4_fix(cwn[b] + BIGE);
                }
            }
            // quads that nobody writes during the sweep (an idle warp's block, the tail of the row)
            // are "invalid" in both strip buffers
            {
                const uint4 inf4 = make_uint4((unsigned)kKeyInf, (unsigned)kKeyInf, (unsigned)kKeyInf, (unsigned)kKeyInf);
                if (!active && own)
                    for (int rr = 0; rr < 2 * NC; ++rr) strip4[rr * TC + sidx] = inf4;
                for (int p = tid; p < 2 * NC * TC; p += TC)
                    if ((p % TC) >= NC * OWN) strip4[p] = inf4;
            }
            __syncthreads();  // (A) item set up by both roles

            int slot0 = 0, slot1 = RD - 1, slotm = 1 % RD, si = 0, sidx_strip = 0;
            double dot[CPL];
            {
                s4_mbar_wait(&mbar[0], par0);
                par0 ^= 1u;
                if (tid == 0 && kS4Stage < Lr) {
                    const unsigned nn = (unsigned)imin(kS4Stage, Lr - kS4Stage) * BINS * 8u;
                    s4_bulk_load(xs + kS4Stage * BINS, rframes + (int64_t)kS4Stage * BINS, nn, &mbar[1]);
                }
#pragma unroll
                for (int r = 0; r < CPL; ++r) {
                    double acc = 0.0;
                    if (active) {
#pragma unroll
                        for (int k = 0; k < BINS; ++k) acc = fma(xs[k], y2[r][k], acc);
                    }
                    dot[r] = acc;
                }
            }
            for (int a = 0; a < Lr; ++a) {
                const int i = a - (m - 1);
                const double* r1 = ring_t + slot1 * RS;
                const double* rm = ring_t + slotm * RS;
                double* r0 = ring_t + slot0 * RS;
                double P[CPL];
                if (active) {
                    P[0] = r1[3 * TWP - 1] + dot[0];
#pragma unroll
                    for (int r = 1; r < CPL; ++r) P[r] = Pprev[r - 1] + dot[r];
                }
                // ---- start row a + 1 (the last iteration recomputes row Lr - 1; its result is unused)
                const int an = (a + 1 < Lr) ? a + 1 : Lr - 1;
                const int bufn = (an >> 5) & 1;
                if ((an & (kS4Stage - 1)) == 0 && a + 1 < Lr) {
                    named_bar_sync(kS5BarCompute, TC);  // every compute warp is past the other buffer
                    if (tid == 0 && an + kS4Stage < Lr) {
                        const unsigned nn = (unsigned)imin(kS4Stage, Lr - (an + kS4Stage)) * BINS * 8u;
                        s4_bulk_load(xs + (bufn ^ 1) * kS4Stage * BINS, rframes + (int64_t)(an + kS4Stage) * BINS, nn,
                                     &mbar[bufn ^ 1]);
                    }
                    if (bufn == 0) {
                        s4_mbar_wait(&mbar[0], par0);
                        par0 ^= 1u;
                    } else {
                        s4_mbar_wait(&mbar[1], par1);
                        par1 ^= 1u;
                    }
                }
                if (active) {
                    const double* xr = xs + (bufn * kS4Stage + (an & (kS4Stage - 1))) * BINS;
                    double x[BINS];
#pragma unroll
                    for (int k = 0; k + 1 < BINS; k += 2) {
                        const double2 v = *reinterpret_cast<const double2*>(xr + k);
                        x[k] = v.x;
                        x[k + 1] = v.y;
                    }
                    if (BINS & 1) x[BINS - 1] = xr[BINS - 1];
#pragma unroll
                    for (int r = 0; r < CPL; ++r) {
                        double acc = 0.0;
#pragma unroll
                        for (int k = 0; k < BINS; ++k) acc = fma(x[k], y2[r][k], acc);
                        dot[r] = acc;
                    }
                    // ---- rest of row a
                    int key[CPL];
#pragma unroll
                    for (int r = 0; r < CPL; ++r) {
                        const double zz = (P[r] - rm[pooff[r]]) + OFF;
                        r0[r * TWP] = P[r];
                        Pprev[r] = P[r];
                        key[r] = s4_fix(zz) + kwy[r];
                    }
                    if (vmask != 0xfu) {
#pragma unroll
                        for (int r = 0; r < CPL; ++r)
                            if (!(vmask & (1u << r))) key[r] = kKeyInf;
                    }
                    if (i >= 0) {
                        // the buffer of strip s was last read for strip s - 2
                        if (si == 0 && sidx_strip >= 2) named_bar_sync(kS5BarEmpty + (sidx_strip & 1), 2 * TC);
                        if (own)
                            strip4[((sidx_strip & 1) * NC + si) * TC + sidx] =
                                make_uint4((unsigned)key[0], (unsigned)key[1], (unsigned)key[2], (unsigned)key[3]);
                    }
                    __syncwarp();  // the ring is private to the warp
                } else if (i >= 0 && si == 0 && sidx_strip >= 2) {
                    named_bar_sync(kS5BarEmpty + (sidx_strip & 1), 2 * TC);
                }
                slot1 = slot0;
                slot0 = (slot0 + 1 == RD) ? 0 : slot0 + 1;
                slotm = (slotm + 1 == RD) ? 0 : slotm + 1;
                if (i >= 0) {
                    if (si == NC - 1 || a == Lr - 1) {
                        named_bar_arrive(kS5BarFull + (sidx_strip & 1), 2 * TC);  // strip delivered
                        ++sidx_strip;
                        si = 0;
                    } else {
                        ++si;
                    }
                }
            }
            __syncthreads();  // (B) end of item: nobody still reads the item's shared state
        }
    } else {
        // =============================================================== selection warps
        s5_regs_select();
        for (;;) {
            ACOSS_S5_FETCH_ITEM();
            const int sw = warp - NC;  // strip row resolved by this warp
            const int st_tid = tid - TC;
            const bool nonmutual = (A.flags & kFlagNonMutual) != 0;
            const bool use_row = !(nonmutual && orient == 1);
            const bool use_col = !(nonmutual && orient == 0);
            const int64_t sep_off = pwp->sep_off;
            const int64_t rsep_off = sep_off + (orient ? pwp->M : 0);
            const int64_t csep_off = sep_off + (orient ? 0 : pwp->M);
            const int kk = orient ? pwp->k_col : pwp->k_row;
            S4Pair pp;
            pp.fi = A.frames + A.offs[song_i] * BINS;
            pp.fj = A.frames + A.offs[song_j] * BINS;
            pp.bins = BINS;
            pp.oti = pwp->oti;
            pp.m = m;
            pp.orient = orient;
            if (MODE == 2) {
                for (int b = st_tid; b < 4 * TC; b += TC) {
                    const int j = b - (m - 1);
                    int v = -1;
                    if (j >= 0 && j < No) {
                        v = use_col ? A.sepk[csep_off + j] : (kKeyInf - 1);
                        if (v < 0) {
                            const int blo = -v - 1;
                            const int slot = atomicAdd(s_nflag, 1);
                            if (slot < kS4MaxFlag) {
                                fcol[slot] = b;
                                fblo[slot] = blo;
                                fsepx[slot] = A.sepx[csep_off + j];
                                v = blo - 1;
                            } else {
                                // more flagged columns than the patch list holds: the encoded (negative) value stays
                                // and the row code takes its generic per-cell path for them
                                misc[10] = 1;
                            }
                        }
                    }
                    csk_s[b] = v;
                }
            }
            __syncthreads();  // (A) item set up by both roles
            const int nflag = (MODE == 2) ? imin(*s_nflag, kS4MaxFlag) : 0;
            S4Sel st;
            st.spacing = 1.0f;
            int spar = 0;
            for (int s = 0; s < nstrips; ++s) {
                named_bar_sync(kS5BarFull + (s & 1), 2 * TC);  // the compute warps have delivered strip s
                const int nrows = imin(NC, Mo - s * NC);
                if (sw < nrows) {
                    S4RowArgs ra;
                    ra.row4 = strip4 + ((s & 1) * NC + sw) * TC;
                    ra.nq = NC;
                    ra.srow = sw;
                    ra.nrows = nrows;
                    ra.hist_back = 1;
                    ra.use_hist = false;
                    ra.est_rank = -1;
                    ra.irow = s * NC + sw;
                    ra.lane = lane;
                    ra.kk = kk;
                    ra.No = No;
                    ra.m = m;
                    ra.G = G;
                    ra.BANDW = BANDW;
                    ra.use_row = use_row;
                    ra.rwn = A.wnorm + A.offs[rsong];
                    ra.BIGE = BIGE;
                    ra.rsep_off = rsep_off;
                    ra.csep_off = csep_off;
                    ra.misc = misc;
                    ra.spar = spar;
                    ra.cand_s = reinterpret_cast<int*>(smem_raw + LY.cand) + sw * 2 * kS4MaxFlag;
                    ra.s_wcnt = misc + 16 + sw;
                    ra.fcol = fcol;
                    ra.fblo = fblo;
                    ra.fsepx = fsepx;
                    ra.csk_s = csk_s;
                    ra.nflag = nflag;
                    ra.nb_off = pwp->nb_off;
                    ra.nb_stride = pwp->nb_stride;
                    s4_resolve_row<BINS, MODE, NQS>(A, pp, ra, st);
                }
                spar ^= 1;
                if (s + 2 < nstrips) named_bar_arrive(kS5BarEmpty + (s & 1), 2 * TC);  // buffer free for strip s + 2
            }
            __syncthreads();  // (B) end of item: nobody still reads the item's shared state
        }
    }
#undef ACOSS_S5_FETCH_ITEM
}

}  // namespace acoss
