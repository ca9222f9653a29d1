// kernels_sweep4.cuh -- the cross-similarity sweep, fourth generation (the default for songs of
// up to 1408 frames and 12 bins; everything else goes through kernels_sweep.cuh / _sweep2.cuh).
//
// Same mathematics as before (get_csm + sliding_csm + csm_to_binary[_mutual],
// CRPUtils.py:24-45,67-84,169-219): float64 distances, one CTA per (pair, orientation), thread t
// owns columns 4t..4t+3, 12 DFMA + 2 DADD per cell.  What changed after the ncu capture of the
// second generation (profiles/r01_d_*: ~1400 warp instructions per row spent in the float64
// counting-pass selection, FP64 pipe 20 % busy):
//
//   * FIXED-POINT KEYS.  Per pair a power of two 2^e > 4 (max windowed norm of song i + of
//     song j) is chosen (k_oti).  The cross term d = sum_k -2 x.y over the window lies in
//     (-2^e/4, 2^e/4), so d + 1.5*2^e sits in the binade [2^e, 2^(e+1)) and bits 52..21 of that
//     double ARE the integer floor((d + 2^e/2) / u), u = 2^(e-31): one funnel shift, no
//     conversion instruction.  The windowed norms become integers the same way (KW), and
//         key(i, j) = floor((d + 2^e/2) / u) + KW_row(i) + KW_col(j)          (int32 < 2^31)
//     is a monotone function of the windowed squared distance with resolution u (~1e-8
//     relative).  Both orientations of a pair compute the same d up to float64 noise, so their
//     keys differ by at most one unit (a floor boundary).
//   * SELECTION ON INTEGERS, no float64: counting passes (2 integer instructions per key, four
//     keys per LDS.128) bracket the k-th smallest key within kS4EMax ranks, then one pass keeps
//     the two nearest keys per lane and REDUX.MIN extracts the k-th and (k+1)-th exactly.
//   * SAFE SEPARATORS.  thr = mid-point of the k-th and (k+1)-th keys.  If they are >= 2G units
//     apart (G = 2) every key is at least G units from thr and the other orientation's keys
//     (off by <= 1) fall on the same side: the consumer's test is one integer compare.  Rows
//     whose boundary keys are closer than that (~1e-4 of the rows) are resolved in float64 on
//     a canonical direct evaluation of the few cells involved (s4_cell_exact) and exported as
//     FLAGGED separators (band + exact float64 value); consumers re-evaluate only cells whose
//     key falls inside a flagged band.  The CRP is therefore the float64 one.
//   * The row song's frames are staged by TMA bulk copies (cp.async.bulk + mbarrier), double
//     buffered, 32 rows at a time; OTI's rotation is applied to the register-resident song.
//   * MODE 2 emits the CRP as one byte per thread per row (low nibble = the thread's four
//     columns, frame-column space); k_nb_to_packed packs / shifts it for the DP kernel.
#pragma once
#include "common.cuh"

#ifdef ACOSS_EMU
#define ACOSS_MAXNREG4(n)
#else
#define ACOSS_MAXNREG4(n) __maxnreg__(n)
#endif

namespace acoss {

#if defined(ACOSS_S7_STATS) && !defined(ACOSS_EMU)
// selection statistics (profiling builds): [0] s4_select calls, [1] counting passes, [2] walks, [3] walks that failed,
// [4] boundary passes, [5] rows that left the main loop unresolved (bisection), [6] estimator calls, [7] no-guess starts
static __device__ unsigned long long g_s4_sel_stats[8];
#define S4_STAT(k) if (lane == 0) atomicAdd(&g_s4_sel_stats[k], 1ull)
#else
#define S4_STAT(k)
#endif

constexpr int kS4Stage = 32;        // rows of the row song per TMA bulk copy
constexpr int kS4MaxWarps = 11;     // CTAs of up to 352 threads (12 warps fit, measured no faster: profiles/exp/README_r02.md)
#ifndef ACOSS_S4_EMAX
#define ACOSS_S4_EMAX 6
#endif
constexpr int kS4EMax = ACOSS_S4_EMAX;   // ranks walked by the extraction pass (A/B builds: -DACOSS_S4_EMAX=n)
constexpr int kS4MaxFlag = 32;      // flagged columns / patch candidates kept per item / row
constexpr int32_t kKeyInf = 0x7fffffff;

struct Sweep4Args {
    const double* frames;   // [total_frames][BINS]
    const double* wnorm;    // [total_frames] windowed squared norms: sum of the m frames ending here
    const double* norms;    // [total_frames] squared norms of the frames (gen 6, retired)
    const int64_t* offs;
    const PairWork* pairs;
    const uint32_t* items;  // MODE 0: pair*2 + orientation; MODE 2: pair
    int n_items;
    unsigned int* counter;
    int32_t* sepk;          // integer separators (same indexing as the float64 seps of gen 1/2)
    double* sepx;           // exact float64 separators of flagged entries
    uint8_t* nb;            // MODE 2 output: rows x nb_stride bytes per pair
    int m;
    int gap;                // G: keys closer than 2G to a boundary make the row "unsafe"
    uint32_t flags;
};

__host__ __device__ inline int s4_padt(int m) { return (m + 3) / 4 + 1; }
// Every warp sweeps its own block of columns independently of the other warps: its first
// ceil((m-1)/4) lanes recompute the m-1 columns left of the block (a halo), so the prefix along
// the diagonals can start at the block's edge and no row needs a CTA-wide barrier.
__host__ __device__ inline int s4_halo_lanes(int m) { return (m - 1 + 3) / 4; }
__host__ __device__ inline int s4_own_lanes(int m) { return 32 - s4_halo_lanes(m); }
__host__ __device__ inline int s4_own_cols(int m) { return 4 * s4_own_lanes(m); }  // columns a warp contributes
__host__ __device__ inline int s4_max_cols(int m) { return kS4MaxWarps * s4_own_cols(m); }

// shared-memory carve-up (bytes), identical on host and device
struct S4Layout {
    int xs, ring, strip, csk, fcol, fblo, fsepx, cand, misc, mbar, total;
};
__host__ __device__ inline S4Layout s4_layout(int T, int m, int bins) {
    S4Layout L;
    const int NW = T / 32;
    int o = 0;
    L.xs = o;    o += 2 * kS4Stage * bins * 8;                 // double-buffered row frames (16 B aligned)
    L.ring = o;  o += NW * (m + 1) * 4 * (32 + s4_padt(m)) * 8;  // prefix rings, one per warp
    L.fsepx = o; o += kS4MaxFlag * 8;                           // exact separators of flagged columns
    L.strip = o; o += NW * 4 * T * 4;                           // keys of the current strip (uint4 per thread)
    L.csk = o;   o += 4 * T * 4;                                // column separators (certain-in bounds)
    L.fcol = o;  o += kS4MaxFlag * 4;
    L.fblo = o;  o += kS4MaxFlag * 4;
    L.cand = o;  o += NW * 2 * kS4MaxFlag * 4;                  // per warp: candidate columns, then flags
    L.misc = o;  o += 16 * 4 + NW * 4;                          // s_item, s_nflag, per-warp counters
    o = (o + 7) & ~7;
    L.mbar = o;  o += 2 * 8;
    L.total = (o + 15) & ~15;
    return L;
}

// ---------------------------------------------------------------- TMA bulk copy + mbarrier
__device__ __forceinline__ void s4_mbar_init(uint64_t* bar, int count) {
#ifndef ACOSS_EMU
    const unsigned a = (unsigned)__cvta_generic_to_shared(bar);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(a), "r"(count) : "memory");
#else
    *bar = 0;
    (void)count;
#endif
}
__device__ __forceinline__ void s4_fence_mbar_init() {
#ifndef ACOSS_EMU
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
#endif
}
// one thread: arm the barrier with the byte count and start the copy global -> shared
__device__ __forceinline__ void s4_bulk_load(void* dst_smem, const void* src_gmem, unsigned bytes, uint64_t* bar) {
#ifndef ACOSS_EMU
    const unsigned b = (unsigned)__cvta_generic_to_shared(bar);
    const unsigned d = (unsigned)__cvta_generic_to_shared(dst_smem);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(d),
                 "l"(src_gmem), "r"(bytes), "r"(b)
                 : "memory");
#else
    memcpy(dst_smem, src_gmem, bytes);
    (void)bar;
#endif
}
__device__ __forceinline__ void s4_mbar_wait(uint64_t* bar, unsigned parity) {
#ifndef ACOSS_EMU
    const unsigned a = (unsigned)__cvta_generic_to_shared(bar);
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(a),
        "r"(parity)
        : "memory");
#else
    (void)bar;
    (void)parity;
#endif
}

// ---------------------------------------------------------------- fixed point
// bits 52..21 of t, t in [2^e, 2^(e+1)) with e odd (biased exponent even -> bit 52 is 0)
__device__ __forceinline__ int s4_fix(double t) {
    return (int)__funnelshift_r((unsigned)__double2loint(t), (unsigned)__double2hiint(t), 21);
}

__device__ __forceinline__ unsigned s4_umin(unsigned a, unsigned b) { return a < b ? a : b; }
__device__ __forceinline__ unsigned s4_umax(unsigned a, unsigned b) { return a > b ? a : b; }

// Canonical float64 evaluation of one cell of the pair's windowed squared distance:
// sum over the window of |roll(x_i, oti) - y_j|^2 (Serra09.py:166-170), rows = song_i windows.
// Used only for the few cells of "unsafe" rows; never inlined so that producer and consumer
// execute the same instruction sequence.
static __device__ __noinline__ double s4_cell_exact(const double* __restrict__ fi, const double* __restrict__ fj, int bins,
                                             int oti, int m, int wi, int wj) {
    double acc = 0.0;
    for (int q = 0; q < m; ++q) {
        const double* xi = fi + (int64_t)(wi + q) * bins;
        const double* yj = fj + (int64_t)(wj + q) * bins;
        for (int k = 0; k < bins; ++k) {
            int src = k - oti;
            if (src < 0) src += bins;
            const double dlt = xi[src] - yj[k];
            acc = fma(dlt, dlt, acc);
        }
    }
    return acc;
}

// ---------------------------------------------------------------- selection (one warp, one strip row)
// The keys of one strip row as seen by one lane: quads lane, lane+32, ...  NQ > 0 keeps them in
// registers for the whole selection (every pass is then pure integer ALU work); NQ == 0 streams
// them from shared memory (any CTA size).
// ACOSS_S4_PACKED_COUNT=1: counting passes on packed 15-bit keys (DPX).  Measured on the B200: half the
// instructions per pass, no change in time (the selection is latency bound), and the 2 NQ extra
// registers push the 128-register kernels into spills -- off by default.
#ifndef ACOSS_S4_PACKED_COUNT
#define ACOSS_S4_PACKED_COUNT 0
#endif
template <int NQ>
struct S4Keys {
    static constexpr bool kPacked = ACOSS_S4_PACKED_COUNT != 0;
    uint4 k[NQ];
    unsigned nk[kPacked ? 2 * NQ : 1];  // minus the upper 16 bits (15 significant) of every key, two per word
    __device__ __forceinline__ void load(const uint4* __restrict__ row4, int lane, int) {
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            k[q] = row4[lane + 32 * q];
            if (kPacked) {
                nk[kPacked ? 2 * q : 0] = __vneg2(__byte_perm(k[q].x, k[q].y, 0x7632));
                nk[kPacked ? 2 * q + 1 : 0] = __vneg2(__byte_perm(k[q].z, k[q].w, 0x7632));
            }
        }
    }
    // Strided assignment: lane l holds columns l, l + 32, l + 64, ...  Neighbouring columns have similar distances
    // (the songs are smooth), so with four consecutive columns per lane the few keys next to the separator cluster in
    // one lane and the two-deep candidate walk runs dry (measured: 35 % of the walks); far-apart columns do not.
    __device__ __forceinline__ void load_strided(const int* __restrict__ row, int lane) {
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            k[q].x = (unsigned)row[lane + 32 * (4 * q)];
            k[q].y = (unsigned)row[lane + 32 * (4 * q + 1)];
            k[q].z = (unsigned)row[lane + 32 * (4 * q + 2)];
            k[q].w = (unsigned)row[lane + 32 * (4 * q + 3)];
            if (kPacked) {
                nk[kPacked ? 2 * q : 0] = __vneg2(__byte_perm(k[q].x, k[q].y, 0x7632));
                nk[kPacked ? 2 * q + 1 : 0] = __vneg2(__byte_perm(k[q].z, k[q].w, 0x7632));
            }
        }
    }
    __device__ __forceinline__ uint4 get(int q) const { return k[q]; }
    __device__ __forceinline__ int nq() const { return NQ; }
    // number of keys whose upper half is <= t15 (0 <= t15 <= 32766): one DPX op + one packed add per TWO keys
    __device__ __forceinline__ int count15(int t15) const {
        const unsigned tt = (unsigned)(t15 + 1) * 0x00010001u, one = 0x00010001u;
        unsigned a0 = 0, a1 = 0;
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            a0 = __vadd2(a0, __viaddmin_s16x2_relu(nk[kPacked ? 2 * q : 0], tt, one));
            a1 = __vadd2(a1, __viaddmin_s16x2_relu(nk[kPacked ? 2 * q + 1 : 0], tt, one));
        }
        const unsigned a = __vadd2(a0, a1);
        return __reduce_add_sync(kFullMask, (int)((a & 0xffffu) + (a >> 16)));
    }
};
template <>
struct S4Keys<0> {
    static constexpr bool kPacked = false;
    __device__ __forceinline__ int count15(int) const { return 0; }
    const uint4* p;
    int n;
    __device__ __forceinline__ void load(const uint4* __restrict__ row4, int lane, int nq_) {
        p = row4 + lane;
        n = nq_;
    }
    __device__ __forceinline__ uint4 get(int q) const { return p[32 * q]; }
    __device__ __forceinline__ int nq() const { return n; }
};

// number of keys <= t (keys and t in [-1, 2^31): the sign of t - key says key > t)
template <class K>
__device__ __forceinline__ int s4_count(const K& keys, int t, int lane) {
    unsigned s0 = 0, s1 = 0, s2 = 0, s3 = 0;  // four independent chains
    const int nq = keys.nq();
#pragma unroll
    for (int q = 0; q < nq; ++q) {
        const uint4 v = keys.get(q);
        s0 += (unsigned)(t - (int)v.x) >> 31;
        s1 += (unsigned)(t - (int)v.y) >> 31;
        s2 += (unsigned)(t - (int)v.z) >> 31;
        s3 += (unsigned)(t - (int)v.w) >> 31;
    }
    (void)lane;
    return 128 * nq - __reduce_add_sync(kFullMask, (int)((s0 + s1) + (s2 + s3)));
}

// The `need` smallest values of f(key) where f(key) = key - base (UP) or base - key (!UP); keys on
// the wrong side of base wrap to huge unsigned values.  Every lane keeps its two smallest,
// then `need` rounds of REDUX.MIN pop them in order.  o_prev = the (need-1)-th, o_last = the
// need-th smallest (1-based), as offsets.  Returns false when a lane ran out of kept values
// (its third candidate might have been skipped): the caller keeps bracketing instead.
template <bool UP, class K>
__device__ __forceinline__ bool s4_walk(const K& keys, unsigned base, int need, int lane, unsigned& o_prev,
                                        unsigned& o_last) {
    // two independent (smallest, second smallest) chains, merged at the end
    unsigned a1 = 0xffffffffu, a2 = 0xffffffffu, b1 = 0xffffffffu, b2 = 0xffffffffu;
    const int nq = keys.nq();
#pragma unroll
    for (int q = 0; q < nq; ++q) {
        const uint4 v = keys.get(q);
        unsigned x, y;
        x = UP ? (v.x - base) : (base - v.x);
        y = UP ? (v.y - base) : (base - v.y);
        a2 = s4_umin(a2, s4_umax(a1, x));
        a1 = s4_umin(a1, x);
        b2 = s4_umin(b2, s4_umax(b1, y));
        b1 = s4_umin(b1, y);
        x = UP ? (v.z - base) : (base - v.z);
        y = UP ? (v.w - base) : (base - v.w);
        a2 = s4_umin(a2, s4_umax(a1, x));
        a1 = s4_umin(a1, x);
        b2 = s4_umin(b2, s4_umax(b1, y));
        b1 = s4_umin(b1, y);
    }
    unsigned m1 = s4_umin(a1, b1);
    unsigned m2 = s4_umin(s4_umax(a1, b1), s4_umin(a2, b2));
    int popped = 0;
    bool exhausted_early = false;
    o_prev = 0xffffffffu;
    o_last = 0xffffffffu;
    for (int i = 0; i < need; ++i) {
        const unsigned g = __reduce_min_sync(kFullMask, m1);
        const unsigned owners = __ballot_sync(kFullMask, m1 == g);
        const int owner = __ffs((int)owners) - 1;
        if (lane == owner) {
            m1 = m2;
            m2 = 0xffffffffu;
            ++popped;
            if (popped >= 2 && i + 1 < need) exhausted_early = true;
        }
        o_prev = o_last;
        o_last = g;
    }
    return !__any_sync(kFullMask, exhausted_early);
}

// largest key <= t and smallest key > t (kKeyInf when there is none)
template <class K>
__device__ __forceinline__ void s4_boundary(const K& keys, int t, int& below, int& above) {
    unsigned a1 = 0xffffffffu, b1 = 0xffffffffu;
    const unsigned ut = (unsigned)t, up = (unsigned)t + 1u;
    const int nq = keys.nq();
#pragma unroll
    for (int q = 0; q < nq; ++q) {
        const uint4 v = keys.get(q);
        a1 = s4_umin(a1, s4_umin(s4_umin(ut - v.x, ut - v.y), s4_umin(ut - v.z, ut - v.w)));
        b1 = s4_umin(b1, s4_umin(s4_umin(v.x - up, v.y - up), s4_umin(v.z - up, v.w - up)));
    }
    a1 = __reduce_min_sync(kFullMask, a1);
    b1 = __reduce_min_sync(kFullMask, b1);
    below = t - (int)a1;
    above = (b1 >= 0x80000000u) ? kKeyInf : (int)(up + b1);
}

struct S4Sel {
    float spacing;  // key units per rank around the separator
    float scale;    // running estimate of (k-th key - row minimum), 0 = none yet (histogram selection)
};

// k-th and (k+1)-th smallest key of the strip row (1 <= k < nvalid; invalid positions hold
// kKeyInf).  t0 = first threshold tried when have_guess.  All lanes return the same values.
template <class K>
__device__ __forceinline__ void s4_select(const K& keys, int k, int nvalid, int t0, bool have_guess, S4Sel& st, int lane,
                                          int& vk, int& vk1) {
    int lo = -1, clo = 0, hi = kKeyInf - 1, chi = nvalid;
    bool real_lo = false, real_hi = false;
    int t = t0;
    const int nq = keys.nq();
    S4_STAT(0);
    if (!have_guess) {
        S4_STAT(7);
        // no guess yet: bracket by the row's extremes
        unsigned mn = 0xffffffffu, mx = 0u;
#pragma unroll
        for (int q = 0; q < nq; ++q) {
            const uint4 v = keys.get(q);
            mn = s4_umin(mn, s4_umin(s4_umin(v.x, v.y), s4_umin(v.z, v.w)));
            if (v.x != (unsigned)kKeyInf) mx = s4_umax(mx, v.x);
            if (v.y != (unsigned)kKeyInf) mx = s4_umax(mx, v.y);
            if (v.z != (unsigned)kKeyInf) mx = s4_umax(mx, v.z);
            if (v.w != (unsigned)kKeyInf) mx = s4_umax(mx, v.w);
        }
        mn = __reduce_min_sync(kFullMask, mn);
        mx = __reduce_max_sync(kFullMask, mx);
        lo = (int)mn - 1;
        hi = (int)mx;
        real_lo = real_hi = true;
        t = lo + (int)((float)(hi - lo) * __fdividef((float)k, (float)nvalid));
        if (t <= lo) t = lo + 1;
        if (t >= hi) t = hi - 1;
        if (hi - lo <= 1) t = hi;  // all keys equal
    }
    if (t < 0) t = 0;
    if (t > kKeyInf - 2) t = kKeyInf - 2;
    int tA = 0, cA = 0;
    bool haveA = false;
    float spacing = st.spacing > 1.0f ? st.spacing : 1.0f;
    int gallop = 0;
    bool done = false;
    for (int iter = 0; iter < 96 && !done; ++iter) {
        int c;
        if (K::kPacked) {
            // thresholds of the form x * 2^16 + 0xffff are counted on the packed upper halves (half the
            // instructions); any threshold inside the bracket is as good as another for the search
            const int tr = t | 0xffff;
            if (tr < 0x7fff0000 && !(real_hi && tr >= hi)) t = tr;
            c = ((t & 0xffff) == 0xffff && t < 0x7fff0000) ? keys.count15(t >> 16) : s4_count(keys, t, lane);
        } else {
            c = s4_count(keys, t, lane);
        }
        S4_STAT(1);
        if (haveA && c != cA) {
            const float sp = fabsf((float)(t - tA)) * __fdividef(1.0f, fabsf((float)(c - cA)));
            if (sp >= 1.0f) spacing = sp;
        }
        tA = t;
        cA = c;
        haveA = true;
        const int e = k - c;
        if (e == 0) {
            S4_STAT(4);
            s4_boundary(keys, t, vk, vk1);
            done = true;
            break;
        }
        if (e > 0) {
            lo = t; clo = c; real_lo = true;
        } else {
            hi = t; chi = c; real_hi = true;
        }
        if (e > 0 && e <= kS4EMax) {
            unsigned oa, ob;
            const bool ok = s4_walk<true>(keys, (unsigned)t + 1u, e + 1, lane, oa, ob);
            S4_STAT(2);
            if (!(ok && oa < 0x80000000u)) { S4_STAT(3); }
            if (ok && oa < 0x80000000u) {
                vk = (int)((unsigned)t + 1u + oa);
                vk1 = (ob >= 0x80000000u) ? kKeyInf : (int)((unsigned)t + 1u + ob);
                done = true;
                break;
            }
            // otherwise keep bracketing (the bracket [lo, hi] still holds)
        } else if (e < 0 && -e <= kS4EMax) {
            unsigned oa, ob;
            const bool ok = s4_walk<false>(keys, (unsigned)t, -e + 1, lane, oa, ob);
            S4_STAT(2);
            if (!(ok && ob < 0x80000000u)) { S4_STAT(3); }
            if (ok && ob < 0x80000000u) {
                vk1 = t - (int)oa;
                vk = t - (int)ob;
                done = true;
                break;
            }
        }
        // next threshold
        if (real_lo && real_hi) {
            if (hi - lo <= 1) {
                // count jumps across k between two adjacent integers: ranks k and k+1 share key hi
                vk = hi;
                vk1 = hi;
                done = true;
                break;
            }
            int nt;
            if ((iter % 3) == 2) {
                nt = lo + ((hi - lo) >> 1);
            } else {
                const float fr = __fdividef((float)(k - clo), (float)(chi - clo));
                nt = lo + (int)((float)(hi - lo) * fr);
            }
            if (nt <= lo) nt = lo + 1;
            if (nt >= hi) nt = hi - 1;
            t = nt;
        } else {
            float stepf = spacing * 1.25f * fabsf((float)e) * (float)(1 << (gallop < 20 ? gallop : 20));
            if (stepf < 1.0f) stepf = 1.0f;
            if (stepf > 1.0e9f) stepf = 1.0e9f;
            const int step = (int)stepf;
            ++gallop;
            int64_t nt = (e > 0) ? (int64_t)t + step : (int64_t)t - step;
            if (nt < 0) nt = 0;
            if (nt > kKeyInf - 2) nt = kKeyInf - 2;
            if (real_hi && nt >= hi) nt = hi - 1;
            if (real_lo && nt <= lo) nt = lo + 1;
            if ((int)nt == t) {
                // pinned against the domain edge: the bracket is now real on that side
                if (e > 0) { hi = kKeyInf - 1; chi = nvalid; real_hi = true; }
                else { lo = -1; clo = 0; real_lo = true; }
                nt = lo + ((hi - lo) >> 1);
                if (nt <= lo) nt = lo + 1;
            }
            t = (int)nt;
        }
    }
    if (!done) {
        S4_STAT(5);
        // plain bisection for the smallest t with count(<= t) >= k, then its successor
        int l = -1, h = kKeyInf - 1;
        if (real_lo) l = lo;
        if (real_hi) h = hi;
        while (h - l > 1) {
            const int mid = l + ((h - l) >> 1);
            if (s4_count(keys, mid, lane) >= k) h = mid; else l = mid;
        }
        vk = h;
        if (s4_count(keys, h, lane) >= k + 1) {
            vk1 = h;
        } else {
            int dummy;
            s4_boundary(keys, h, dummy, vk1);
        }
    }
    st.spacing = spacing;
}

// A first threshold from the row's own keys (no history needed): every lane finds the second
// smallest of its keys; the k-th key of the row is close to the jrank-th smallest of those 32
// values, where jrank = 32 P(a lane holds at least two of the k smallest) - 1 (s4_est_rank).
// On the benchmark's rows the count at this threshold is within 6 ranks of k for ~85 % of the
// rows, against ~15 % for the separator of a row 8-16 rows earlier.  Returns kKeyInf when there
// is no such value.
template <class K>
__device__ __forceinline__ int s4_estimate(const K& keys, int jrank, int lane) {
    S4_STAT(6);
    unsigned a1 = 0xffffffffu, a2 = 0xffffffffu, b1 = 0xffffffffu, b2 = 0xffffffffu;
    const int nq = keys.nq();
#pragma unroll
    for (int q = 0; q < nq; ++q) {
        const uint4 v = keys.get(q);
        a2 = s4_umin(a2, s4_umax(a1, v.x));
        a1 = s4_umin(a1, v.x);
        b2 = s4_umin(b2, s4_umax(b1, v.y));
        b1 = s4_umin(b1, v.y);
        a2 = s4_umin(a2, s4_umax(a1, v.z));
        a1 = s4_umin(a1, v.z);
        b2 = s4_umin(b2, s4_umax(b1, v.w));
        b1 = s4_umin(b1, v.w);
    }
    // lanes with fewer than two keys hold 0xffffffff: clamp so that differences of two values keep their sign
    const unsigned m2 = s4_umin(s4_umin(s4_umax(a1, b1), s4_umin(a2, b2)), (unsigned)kKeyInf);
    unsigned r0 = 0, r1 = 0;  // strict rank: equal values (improbable) share a rank and the estimate is skipped
#pragma unroll
    for (int l = 0; l < 32; l += 2) {
        r0 += (__shfl_sync(kFullMask, m2, l) - m2) >> 31;
        r1 += (__shfl_sync(kFullMask, m2, l + 1) - m2) >> 31;
    }
    (void)lane;
    const unsigned hit = __ballot_sync(kFullMask, (int)(r0 + r1) == jrank);
    if (hit == 0u) return kKeyInf;
    const unsigned g = __shfl_sync(kFullMask, m2, __ffs((int)hit) - 1);
    return (g >= (unsigned)kKeyInf) ? kKeyInf : (int)g;
}
// 0-based rank to use with s4_estimate for the k-th of nvalid keys spread over 32 lanes
__device__ __forceinline__ int s4_est_rank(int k, int nvalid) {
    const float n = (float)nvalid * (1.0f / 32.0f), p = (float)k / (float)nvalid;
    const float q1 = exp2f((n - 1.0f) * log2f(1.0f - p));  // (1-p)^(n-1)
    const float P2 = 1.0f - q1 * (1.0f - p) - n * p * q1;  // P(Binomial(n, p) >= 2)
    int j = (int)(32.0f * P2 + 0.5f) - 1;
    return j < 0 ? 0 : (j > 31 ? 31 : j);
}

// Histogram selection (no guess needed): 64 bins of width 2^sh above the row minimum, sized from
// the running estimate `scale` of (k-th key - minimum) so that the k-th key is expected around bin
// 16..32; an inclusive scan over the bin counts finds the bin holding rank k and the exact count
// below it, and a candidate walk from that bin's lower edge pops the few keys up to rank k + 1.
// Over-full bins are refined with a second (third) histogram inside the bin.  Returns false when
// the k-th key lies beyond the 64 bins or the walk ran out of per-lane candidates: the caller
// then uses s4_select, which is exact for any input.  hist_s: 64 ints of per-row scratch.
template <class K>
__device__ __forceinline__ bool s4_select_hist(const K& keys, int k, float scale, int* hist_s, int lane, int& vk,
                                               int& vk1, unsigned& row_min) {
    const int nq = keys.nq();
    unsigned mn = 0xffffffffu;
#pragma unroll
    for (int q = 0; q < nq; ++q) {
        const uint4 v = keys.get(q);
        mn = s4_umin(mn, s4_umin(s4_umin(v.x, v.y), s4_umin(v.z, v.w)));
    }
    mn = __reduce_min_sync(kFullMask, mn);
    row_min = mn;
    if (!(scale >= 1.0f)) return false;
    int sci = (scale < 1.0e9f) ? (int)scale : 1000000000;
    if (sci < 16) sci = 16;
    int sh = 27 - __clz(sci);  // sci >> sh in [16, 32)
    unsigned base = mn;
    int kk = k;
    for (int level = 0; level < 3; ++level) {
        hist_s[lane] = 0;
        hist_s[lane + 32] = 0;
        __syncwarp();
#pragma unroll
        for (int q = 0; q < nq; ++q) {
            const uint4 v = keys.get(q);
            const unsigned b0 = (v.x - base) >> sh, b1 = (v.y - base) >> sh, b2 = (v.z - base) >> sh, b3 = (v.w - base) >> sh;
            if (b0 < 64u) atomicAdd(&hist_s[b0], 1);
            if (b1 < 64u) atomicAdd(&hist_s[b1], 1);
            if (b2 < 64u) atomicAdd(&hist_s[b2], 1);
            if (b3 < 64u) atomicAdd(&hist_s[b3], 1);
        }
        __syncwarp();
        const int h0 = hist_s[2 * lane], h1 = hist_s[2 * lane + 1];
        int c = h0 + h1;  // inclusive scan over lanes
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int o = __shfl_up_sync(kFullMask, c, d);
            if (lane >= d) c += o;
        }
        __syncwarp();
        const unsigned hit = __ballot_sync(kFullMask, c >= kk);
        if (hit == 0u) return false;  // rank k lies beyond the bins
        const int ls = __ffs((int)hit) - 1;
        const int excl = c - h0 - h1;
        const int first = (excl + h0 >= kk) ? 0 : 1;
        const int binl = 2 * lane + first;
        const int belowl = first ? excl + h0 : excl;
        const int popl = first ? h1 : h0;
        const int bin = __shfl_sync(kFullMask, binl, ls);
        const int below = __shfl_sync(kFullMask, belowl, ls);
        const int pop = __shfl_sync(kFullMask, popl, ls);
        const int need = kk - below;  // 1 .. pop
        const unsigned edge = base + ((unsigned)bin << sh);
        if (need + 1 <= kS4EMax + 1) {
            unsigned oa, ob;
            const bool ok = s4_walk<true>(keys, edge, need + 1, lane, oa, ob);
            if (!ok || oa >= 0x80000000u) return false;
            vk = (int)(edge + oa);
            vk1 = (ob >= 0x80000000u) ? kKeyInf : (int)(edge + ob);
            return true;
        }
        // too many keys between the bin's edge and rank k: refine inside the bin
        if (sh == 0 || pop < need) return false;
        base = edge;
        kk = need;
        sh = (sh > 6) ? sh - 6 : 0;
    }
    return false;
}

// Unsafe row: the k-th and (k+1)-th keys are closer than 2G.  Collect the cells whose key lies
// within 2 units of them, evaluate those in float64 (canonical orientation) and rank them.
// Returns true and (sepx, n_cand, in-mask over candidate slots) when resolved; false when the
// row is a genuine tie cluster (more than kS4MaxFlag candidates, or equal float64 values):
// the caller then keeps the key-level separator vk.
struct S4Pair {
    const double* fi;  // frames of song_i / song_j (canonical orientation)
    const double* fj;
    int bins, oti, m, orient;
};
static __device__ __noinline__ bool s4_resolve_unsafe(const uint4* __restrict__ row4, int nq, int k, int vk, int vk1,
                                               const S4Pair& pp, int irow, int mwin, int lane, int* cand_cols,
                                               int* cand_cnt, double& sepx, int& n_cand, unsigned& in_mask) {
    const int blo = vk - 2, bhi = vk1 + 2;
    S4Keys<0> ks;
    ks.load(row4, lane, nq);
    const int n_below = s4_count(ks, blo - 1, lane);
    if (lane == 0) *cand_cnt = 0;
    __syncwarp();
    for (int q = 0; q < nq; ++q) {
        const uint4 v = row4[lane + 32 * q];
        const int kv[4] = {(int)v.x, (int)v.y, (int)v.z, (int)v.w};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if (kv[c] >= blo && kv[c] <= bhi) {
                const int pos = atomicAdd(cand_cnt, 1);
                if (pos < kS4MaxFlag) cand_cols[pos] = 4 * (lane + 32 * q) + c;
            }
        }
    }
    __syncwarp();
    n_cand = *cand_cnt;
    const int r = k - n_below;  // how many candidates belong to the k smallest
    if (n_cand > kS4MaxFlag || r < 1 || r >= n_cand) return false;
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    double z = INF;
    int col = 0x7fffffff;
    if (lane < n_cand) {
        col = cand_cols[lane];
        const int jwin = col - (mwin - 1);
        z = pp.orient ? s4_cell_exact(pp.fi, pp.fj, pp.bins, pp.oti, pp.m, jwin, irow)
                      : s4_cell_exact(pp.fi, pp.fj, pp.bins, pp.oti, pp.m, irow, jwin);
    }
    int rank = 0;
    for (int l = 0; l < 32; ++l) {
        const double zo = __shfl_sync(kFullMask, z, l);
        const int co = __shfl_sync(kFullMask, col, l);
        if (l < n_cand && (zo < z || (zo == z && co < col))) ++rank;
    }
    const bool is_in = (lane < n_cand) && (rank < r);
    const bool is_out = (lane < n_cand) && (rank >= r);
    const double zin = warp_max(is_in ? z : -INF);
    const double zout = warp_min(is_out ? z : INF);
    in_mask = __ballot_sync(kFullMask, is_in);
    sepx = zin + 0.5 * (zout - zin);
    if (!(sepx < zout)) sepx = zin;
    return zin < zout;
}

// ---------------------------------------------------------------- one strip row, one warp
// Selects the row's separator and, MODE 0: exports it (absolute key units, flagged rows with
// their float64 value); MODE 2: turns the row into CRP nibbles using the column separators.
struct S4RowArgs {
    const uint4* row4;   // the strip row (nq * 32 quads)
    int nq, srow, nrows; // quads per lane, row within the strip, rows in the strip
    int est_rank;        // >= 0: first threshold from s4_estimate with this rank; < 0: from the separator history
    bool use_hist;       // histogram selection first (cand_s doubles as its 64-bin scratch)
    int hist_back;       // the guess extrapolates from the strip's last row and the one hist_back rows before it
    int irow, lane;      // window index within the row song
    int kk, No, m, G, BANDW;
    bool use_row;
    const double* rwn;   // windowed norms of the row song
    double BIGE;
    int64_t rsep_off;
    int64_t csep_off;    // column separators of the item (MODE 2)
    int* misc;           // misc[2..5] separator history, misc[8..9] its counts, misc[10] flag-list overflow
    int spar;
    int* cand_s;         // per-warp scratch: 2 * kS4MaxFlag ints
    int* s_wcnt;
    const int* fcol;     // flagged columns of the item
    const int* fblo;
    const double* fsepx;
    const int* csk_s;    // column separators (certainly-in bounds)
    int nflag;
    int64_t nb_off;
    int nb_stride;
};

template <int BINS, int MODE, int NQ>
__device__ __forceinline__ void s4_resolve_row(const Sweep4Args& A, const S4Pair& pp, const S4RowArgs& R, S4Sel& st) {
    const int irow = R.irow, lane = R.lane, warp = R.srow, nrows = R.nrows, m = R.m, kk = R.kk, No = R.No;
    const int G = R.G, BANDW = R.BANDW, nflag = R.nflag, orient = pp.orient, spar = R.spar, NW = R.nq;
    int* const misc = R.misc;
    int* const hist = misc + 2;
    int* const cand_s = R.cand_s;
    const int* const fcol = R.fcol;
    const int* const fblo = R.fblo;
    const double* const fsepx = R.fsepx;
    const int* const csk_s = R.csk_s;
    const bool use_row = R.use_row;
    const int64_t rsep_off = R.rsep_off;
    const uint4* row4 = R.row4;
    const int kwx = s4_fix(R.rwn[irow + m - 1] + R.BIGE);
    int thr = kKeyInf - 1;  // row-relative certainly-in bound
    bool rflag = false;
    double rsepx = 0.0;
    int rblo = 0, ncand = 0;
    unsigned cand_in = 0;
    if (kk <= 0) {
        thr = -1;
    } else if (A.flags & (1u << 30)) {
        thr = 1 << 30;  // profiling knob: no selection (results are meaningless)
    } else if (kk < No && (MODE == 0 || use_row)) {
        int vk, vk1;
        S4Keys<NQ> keys;
        keys.load(row4, lane, NW);
        // guess: the previous strip's last separator, extrapolated by half its last step
        const int nh = misc[8 + (spar ^ 1)];
        int g = 0;
        if (nh >= 1) {
            const int tl = hist[2 * (spar ^ 1)];
            g = tl;
            if (nh >= 2) g += ((tl - hist[2 * (spar ^ 1) + 1]) * (warp + 1)) / (2 * R.hist_back);
        }
        bool have = false;
        bool have_g = nh >= 1;
        int t0 = g - kwx;
        if (R.est_rank >= 0) {
            const int ge = s4_estimate(keys, R.est_rank, lane);
            if (ge < kKeyInf - 1) {
                t0 = ge;
                have_g = true;
            }
        }
        if (R.use_hist) {
            unsigned rmin;
            have = s4_select_hist(keys, kk, st.scale, cand_s, lane, vk, vk1, rmin);
            if (!have) s4_select(keys, kk, No, t0, have_g, st, lane, vk, vk1);
            const float sc = (float)(unsigned)(vk - (int)rmin);
            st.scale = (st.scale >= 1.0f) ? 0.5f * (st.scale + sc) : sc;
            if (st.scale < 1.0f) st.scale = 1.0f;
        } else {
            s4_select(keys, kk, No, t0, have_g, st, lane, vk, vk1);
        }
        if (lane == 0) {
            if (warp == nrows - 1) hist[2 * spar] = vk + kwx;
            if (warp == nrows - 1 - R.hist_back) hist[2 * spar + 1] = vk + kwx;
            if (warp == 0) misc[8 + spar] = nrows > R.hist_back ? 2 : 1;
        }
        const int gap = vk1 - vk;
        if (gap >= 2 * G) {
            thr = vk + (gap >> 1);
        } else {
            const bool ok = s4_resolve_unsafe(row4, NW, kk, vk, vk1, pp, irow, m, lane, cand_s, R.s_wcnt, rsepx,
                                              ncand, cand_in);
            if (ok) {
                rflag = true;
                rblo = vk - 3;
                thr = rblo - 1;
            } else {
                // tie cluster (the reference's own choice is undefined here): equal keys across the k-th position are
                // all left out, like the float64 kernels do (warp_select_sep); fewer than k neighbours
                thr = (vk1 == vk) ? vk - 1 : vk;
                ncand = 0;
            }
        }
    }
    if (MODE == 0) {
        if (lane == 0) {
            if (rflag) {
                A.sepk[rsep_off + irow] = -(rblo + kwx) - 1;
                A.sepx[rsep_off + irow] = rsepx;
            } else {
                // kk <= 0 -> -1 ... never in; kk >= No -> everything valid
                A.sepk[rsep_off + irow] = (thr < 0) ? 0 : ((thr >= kKeyInf - 1) ? kKeyInf - 1 : thr + kwx);
                if (thr < 0) A.sepk[rsep_off + irow] = 0;
            }
        }
    } else {
        // ---- patch list: cells of flagged bands decided in float64
        int npatch = 0;
        int* patch_cols = cand_s + kS4MaxFlag;
        if (rflag || nflag > 0) {
            // candidates: the unsafe row's band cells (slots 0..ncand-1 of cand_s), then
            // this row's cell of every flagged column
            const int ntot = ncand + nflag;
            for (int base = 0; base < ntot; base += 32) {
                const int c = base + lane;
                bool setbit = false;
                int col = 0;
                if (c < ntot) {
                    const bool from_row = c < ncand;
                    col = from_row ? cand_s[c] : fcol[c - ncand];
                    const int key = ((const int*)row4)[col];
                    const int kabs = key + kwx;
                    // column test
                    int f = -1;
                    for (int q = 0; q < nflag; ++q)
                        if (fcol[q] == col) f = q;
                    double zx = 0.0;
                    bool have_z = false;
                    bool colin, rowin;
                    if (f >= 0) {
                        if (kabs < fblo[f]) colin = true;
                        else if (kabs > fblo[f] + BANDW) colin = false;
                        else {
                            const int jwin = col - (m - 1);
                            zx = orient ? s4_cell_exact(pp.fi, pp.fj, BINS, pp.oti, m, jwin, irow)
                                        : s4_cell_exact(pp.fi, pp.fj, BINS, pp.oti, m, irow, jwin);
                            have_z = true;
                            colin = zx <= fsepx[f];
                        }
                    } else {
                        colin = kabs <= csk_s[col];
                    }
                    if (rflag) {
                        if (key < rblo) rowin = true;
                        else if (key > rblo + BANDW) rowin = false;
                        else {
                            if (!have_z) {
                                const int jwin = col - (m - 1);
                                zx = orient ? s4_cell_exact(pp.fi, pp.fj, BINS, pp.oti, m, jwin, irow)
                                            : s4_cell_exact(pp.fi, pp.fj, BINS, pp.oti, m, irow, jwin);
                            }
                            rowin = zx <= rsepx;
                        }
                    } else {
                        rowin = key <= thr;
                    }
                    setbit = colin && rowin && key != kKeyInf;
                }
                const unsigned sb = __ballot_sync(kFullMask, setbit);
                if (setbit) {
                    const int pos = npatch + __popc(sb & ((1u << lane) - 1u));
                    if (pos < kS4MaxFlag) patch_cols[pos] = col;
                }
                npatch += __popc(sb);
            }
            if (npatch > kS4MaxFlag) npatch = kS4MaxFlag;
            __syncwarp();
        }
        // ---- bits: nibble per (lane, quad), one byte per thread-quad of the row
        uint8_t* out = A.nb + R.nb_off + (int64_t)irow * R.nb_stride + lane;
        const uint4* csk4 = reinterpret_cast<const uint4*>(csk_s);
        const bool generic = misc[10] != 0;
        const int nqq = (NQ > 0) ? NQ : NW;  // compile-time trip count when the CTA size is
#pragma unroll
        for (int q = 0; q < nqq; ++q) {
            const int tq = lane + 32 * q;
            const uint4 kv = row4[tq];
            const uint4 cv = csk4[tq];
            unsigned nib = 0;
            nib |= ((int)kv.x <= imin(thr, (int)cv.x - kwx)) ? 1u : 0u;
            nib |= ((int)kv.y <= imin(thr, (int)cv.y - kwx)) ? 2u : 0u;
            nib |= ((int)kv.z <= imin(thr, (int)cv.z - kwx)) ? 4u : 0u;
            nib |= ((int)kv.w <= imin(thr, (int)cv.w - kwx)) ? 8u : 0u;
            if (generic && (int)(cv.x | cv.y | cv.z | cv.w) < 0) {
                // generic path (only when the item has more flagged columns than the patch list
                // holds): a negative entry is a flagged column that is not in the list
                const int kvv[4] = {(int)kv.x, (int)kv.y, (int)kv.z, (int)kv.w};
                const int cvv[4] = {(int)cv.x, (int)cv.y, (int)cv.z, (int)cv.w};
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    if (cvv[c] >= -1 || kvv[c] == kKeyInf) continue;  // -1 marks a column outside the plot
                    const int blo = -cvv[c] - 1, kabs = kvv[c] + kwx, col = 4 * tq + c;
                    const int jwin = col - (m - 1);
                    bool colin;
                    if (kabs < blo) colin = true;
                    else if (kabs > blo + BANDW) colin = false;
                    else {
                        const double zx = orient ? s4_cell_exact(pp.fi, pp.fj, BINS, pp.oti, m, jwin, irow)
                                                 : s4_cell_exact(pp.fi, pp.fj, BINS, pp.oti, m, irow, jwin);
                        colin = zx <= A.sepx[R.csep_off + jwin];
                    }
                    bool rowin;
                    if (rflag && kvv[c] >= rblo) {
                        if (kvv[c] > rblo + BANDW) rowin = false;
                        else {
                            const double zx = orient ? s4_cell_exact(pp.fi, pp.fj, BINS, pp.oti, m, jwin, irow)
                                                     : s4_cell_exact(pp.fi, pp.fj, BINS, pp.oti, m, irow, jwin);
                            rowin = zx <= rsepx;
                        }
                    } else {
                        rowin = kvv[c] <= thr;
                    }
                    if (colin && rowin) nib |= 1u << c;
                }
            }
            if (npatch > 0) {
                for (int p = 0; p < npatch; ++p) {
                    const int col = patch_cols[p];
                    if ((col >> 2) == tq) nib |= 1u << (col & 3);
                }
            }
            out[32 * q] = (uint8_t)nib;
        }
    }

}

// ---------------------------------------------------------------- the kernel
// MW > 0: compile-time embedding window; MW == 0: read from the arguments.
// NQ > 0: the CTA has exactly 32 NQ threads and a selecting warp keeps its row's keys in
// registers; NQ == 0: any CTA size, keys streamed from shared memory.
template <int BINS, int MODE, int MW, int NQ>
__global__ void ACOSS_MAXNREG4(168) k_sweep4(Sweep4Args A) {
    constexpr int CPL = 4;
    ACOSS_DYNSMEM(smem_raw);

    const int T = (int)blockDim.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, NW = T >> 5;
    const int m = (MW > 0) ? MW : A.m;
    const int RD = m + 1;
    const int PADT = s4_padt(m), TWP = 32 + PADT, RS = 4 * TWP;
    const int HL = s4_halo_lanes(m), OWN = 32 - HL;
    const bool own = lane >= HL;              // lanes below HL recompute the halo columns
    const int sidx = warp * OWN + lane - HL;  // this thread's quad in a strip row / CRP row
    const int G = A.gap, BANDW = 2 * A.gap + 6;
    const S4Layout LY = s4_layout(T, m, BINS);

    double* const xs = reinterpret_cast<double*>(smem_raw + LY.xs);
    double* const ring = reinterpret_cast<double*>(smem_raw + LY.ring);
    double* const fsepx = reinterpret_cast<double*>(smem_raw + LY.fsepx);
    uint4* const strip4 = reinterpret_cast<uint4*>(smem_raw + LY.strip);  // [NW][T]
    int* const csk_s = reinterpret_cast<int*>(smem_raw + LY.csk);         // [4T]
    int* const fcol = reinterpret_cast<int*>(smem_raw + LY.fcol);
    int* const fblo = reinterpret_cast<int*>(smem_raw + LY.fblo);
    int* const cand_s = reinterpret_cast<int*>(smem_raw + LY.cand) + warp * 2 * kS4MaxFlag;
    int* const misc = reinterpret_cast<int*>(smem_raw + LY.misc);
    int* const s_item = misc;
    int* const s_nflag = misc + 1;
    int* const s_wcnt = misc + 16 + warp;
    uint64_t* const mbar = reinterpret_cast<uint64_t*>(smem_raw + LY.mbar);

    if (tid == 0) {
        s4_mbar_init(&mbar[0], 1);
        s4_mbar_init(&mbar[1], 1);
        s4_fence_mbar_init();
    }
    __syncthreads();
    unsigned par0 = 0, par1 = 0;  // phase parity of the two staging barriers

    double* const ring_w = ring + warp * RD * RS;  // this warp's private ring
    double* const ring_t = ring_w + lane + PADT;
    // quads past the last warp's columns are never written: they stay "invalid"
    for (int p = tid; p < NW * T; p += T)
        if ((p % T) >= NW * OWN) strip4[p] = make_uint4((unsigned)kKeyInf, (unsigned)kKeyInf, (unsigned)kKeyInf, (unsigned)kKeyInf);
    int pooff[CPL];
#pragma unroll
    for (int r = 0; r < CPL; ++r) {
        const int d = r - m;
        const int q = (d >= 0) ? (d >> 2) : -((-d + 3) >> 2);
        pooff[r] = q + (d - 4 * q) * TWP;
    }

    for (;;) {
        if (tid == 0) {
            *s_item = (int)atomicAdd(A.counter, 1u);
            *s_nflag = 0;
            misc[10] = 0;
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
        const PairWork pw = A.pairs[pidx];
        const int rsong = orient ? pw.song_j : pw.song_i;
        const int csong = orient ? pw.song_i : pw.song_j;
        const int Lr = orient ? pw.Lj : pw.Li;
        const int Lc = orient ? pw.Li : pw.Lj;
        const int Mo = Lr - m + 1, No = Lc - m + 1;
        if (Mo <= 0 || No <= 0) continue;
        const int kk = orient ? pw.k_col : pw.k_row;
        const double* rframes = A.frames + A.offs[rsong] * BINS;
        const double* cframes = A.frames + A.offs[csong] * BINS;
        const double* rwn = A.wnorm + A.offs[rsong];
        const double* cwn = A.wnorm + A.offs[csong];
        const int64_t rsep_off = pw.sep_off + (orient ? pw.M : 0);
        const int64_t csep_off = pw.sep_off + (orient ? 0 : pw.M);
        const bool nonmutual = (A.flags & kFlagNonMutual) != 0;
        const bool use_row = !(nonmutual && orient == 1);
        const bool use_col = !(nonmutual && orient == 0);
        const double BIGE = ldexp(1.0, pw.expo);
        const double OFF = 1.5 * BIGE;
        S4Pair pp;
        pp.fi = A.frames + A.offs[pw.song_i] * BINS;
        pp.fj = A.frames + A.offs[pw.song_j] * BINS;
        pp.bins = BINS;
        pp.oti = pw.oti;
        pp.m = m;
        pp.orient = orient;

        // ---- per-item setup
        if (tid == 0) {
            const unsigned n0 = (unsigned)imin(kS4Stage, Lr) * BINS * 8u;
            s4_bulk_load(xs, rframes, n0, &mbar[0]);
        }
        for (int p = tid; p < NW * RD * RS; p += T) ring[p] = 0.0;
        // The rolled song is song_i: roll(x_i, oti)[k] = x_i[(k - oti) mod BINS].  The rotation
        // is always applied to the register-resident (column) song:
        //   rows = song_i: sum_k x_i[k] y_j[(k + oti) mod BINS]
        //   rows = song_j: sum_k y_j[k] x_i[(k - oti) mod BINS]
        const int crot = orient ? (BINS - pw.oti) % BINS : pw.oti;
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
        if (MODE == 2) {
            for (int b = tid; b < 4 * T; b += T) {
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
                            v = blo - 1;  // certainly-in bound; the band is patched per row
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
        S4Sel st;
        st.spacing = 1.0f;
        // absolute separators of the last two rows of the previous strip, double buffered by strip
        // parity: hist[par][0] = last row, [1] = the one before, misc[8 + par] = how many are valid
        int* const hist = misc + 2;
        if (tid == 0) misc[8] = misc[9] = 0;
        int spar = 0;
        __syncthreads();
        const int nflag = (MODE == 2) ? imin(*s_nflag, kS4MaxFlag) : 0;

        int slot0 = 0, slot1 = RD - 1, slotm = 1 % RD, si = 0;
        // The 12-DFMA dot products start from zero, so they do not depend on the previous row, and
        // are issued ONE ROW AHEAD: while they occupy the FP64 pipe, the rest of the current row
        // (diagonal carry, window difference, keys, ring / strip stores) fills the idle issue slots.
        // The carry along the diagonal is one DADD per cell (measured: profiles/exp/ubench/row_loop.cu,
        // 150 -> 102 cycles per row per SM).
        double dot[CPL];
        {
            s4_mbar_wait(&mbar[0], par0);  // block 0 (started at item setup)
            par0 ^= 1u;
            if (tid == 0 && kS4Stage < Lr) {
                const unsigned nn = (unsigned)imin(kS4Stage, Lr - kS4Stage) * BINS * 8u;
                s4_bulk_load(xs + kS4Stage * BINS, rframes + (int64_t)kS4Stage * BINS, nn, &mbar[1]);
            }
#pragma unroll
            for (int r = 0; r < CPL; ++r) {
                double acc = 0.0;
#pragma unroll
                for (int k = 0; k < BINS; ++k) acc = fma(xs[k], y2[r][k], acc);
                dot[r] = acc;
            }
        }
        for (int a = 0; a < Lr; ++a) {
            const int i = a - (m - 1);
            const double* r1 = ring_t + slot1 * RS;
            const double* rm = ring_t + slotm * RS;
            double* r0 = ring_t + slot0 * RS;
            // ---- finish row a: diagonal carry
            double P[CPL];
            P[0] = r1[3 * TWP - 1] + dot[0];
#pragma unroll
            for (int r = 1; r < CPL; ++r) P[r] = Pprev[r - 1] + dot[r];
            // ---- start row a + 1 (the last iteration recomputes row Lr - 1; its result is unused)
            const int an = (a + 1 < Lr) ? a + 1 : Lr - 1;
            const int bufn = (an >> 5) & 1;
            if ((an & (kS4Stage - 1)) == 0 && a + 1 < Lr) {
                // block boundary: all warps are past the other buffer's last reader (barrier), then
                // prefetch the block after this one into it and wait for this block
                __syncthreads();
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
            {
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
            if (i >= 0 && own)
                strip4[si * T + sidx] = make_uint4((unsigned)key[0], (unsigned)key[1], (unsigned)key[2], (unsigned)key[3]);
            __syncwarp();  // the ring is private to the warp
            slot1 = slot0;
            slot0 = (slot0 + 1 == RD) ? 0 : slot0 + 1;
            slotm = (slotm + 1 == RD) ? 0 : slotm + 1;

            if (i < 0) continue;
            if (si != NW - 1 && a != Lr - 1) {
                ++si;
                continue;
            }
            // ---- selection phase: warp w resolves strip row w
            const int nrows = si + 1;
            si = 0;
            __syncthreads();  // every warp has delivered its part of the strip
            if (warp < nrows) {
                S4RowArgs ra;
                ra.row4 = strip4 + warp * T;
                ra.nq = NW;
                ra.srow = warp;
                ra.nrows = nrows;
                ra.hist_back = 1;
                ra.use_hist = false;
                ra.est_rank = -1;
                ra.irow = i - (nrows - 1) + warp;  // window index within the row song
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
                ra.cand_s = cand_s;
                ra.s_wcnt = s_wcnt;
                ra.fcol = fcol;
                ra.fblo = fblo;
                ra.fsepx = fsepx;
                ra.csk_s = csk_s;
                ra.nflag = nflag;
                ra.nb_off = pw.nb_off;
                ra.nb_stride = pw.nb_stride;
                s4_resolve_row<BINS, MODE, NQ>(A, pp, ra, st);
            }
            spar ^= 1;
            __syncthreads();
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------- helpers around the sweep
// One CTA per song: windowed squared norms wnorm[f] = sum of norms of the m frames ending at f
// (0 for the first m-1 frames of a song) and the song's maximum.
static __global__ void __launch_bounds__(128) k_wnorm(const double* __restrict__ norms, const int64_t* __restrict__ offs,
                                               const int32_t* __restrict__ lens, int n_songs, int m,
                                               double* __restrict__ wnorm, double* __restrict__ wnmax) {
    __shared__ double s_max[4];
    const int s = blockIdx.x;
    if (s >= n_songs) return;
    const int L = lens[s];
    const double* nr = norms + offs[s];
    double* wr = wnorm + offs[s];
    double mx = 0.0;
    for (int f = threadIdx.x; f < L; f += blockDim.x) {
        double w = 0.0;
        if (f >= m - 1)
            for (int q = m - 1; q >= 0; --q) w += nr[f - q];
        wr[f] = w;
        mx = fmax(mx, w);
    }
    mx = warp_max(mx);
    if ((threadIdx.x & 31) == 0) s_max[threadIdx.x >> 5] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) mx = fmax(mx, s_max[w]);
        wnmax[s] = mx;
    }
}

// One thread per pair: the fixed-point exponent e (odd, 2^e > 4 (wmax_i + wmax_j)).
static __global__ void __launch_bounds__(128) k_expo(PairWork* __restrict__ pairs, int n_pairs,
                                              const double* __restrict__ wnmax) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pairs) return;
    double s = wnmax[pairs[p].song_i] + wnmax[pairs[p].song_j];
    if (!(s > 1e-280)) s = 1e-280;
    if (!(s < 1e280)) s = 1e280;
    int e = ilogb(s) + 3;
    if ((e & 1) == 0) ++e;
    pairs[p].expo = e;
}

// One CTA per pair: nibble bytes (frame-column space) -> packed CRP words shifted by m-1
// (bit t of word w of row r = CRP[r][32 w + t], the layout of the DP kernel and the debug dump).
static __global__ void __launch_bounds__(128) k_nb_to_packed(const PairWork* __restrict__ pairs, const uint32_t* __restrict__ items,
                                                      int n_items, const uint8_t* __restrict__ nb, uint32_t* __restrict__ crp,
                                                      int m) {
    const int it = blockIdx.x;
    if (it >= n_items) return;
    const PairWork pw = pairs[items[it]];
    const int rows = pw.crp_rows, wp = pw.crp_wp, cols = pw.crp_cols, stride = pw.nb_stride;
    const uint8_t* src = nb + pw.nb_off;
    uint32_t* dst = crp + pw.crp_off;
    const int total = rows * wp;
    for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
        const int r = idx / wp, w = idx - r * wp;
        const uint8_t* row = src + (int64_t)r * stride;
        const int b0 = 32 * w + (m - 1);  // first frame column of this word
        const int q0 = b0 >> 2, sh = b0 & 3;
        uint64_t acc = 0;
#pragma unroll
        for (int t = 0; t < 9; ++t) {
            const int q = q0 + t;
            const uint64_t nib = (q < stride) ? (uint64_t)(row[q] & 0xf) : 0ull;
            acc |= nib << (4 * t);
        }
        uint32_t word = (uint32_t)(acc >> sh);
        const int valid = cols - 32 * w;
        if (valid < 32) word &= (valid <= 0) ? 0u : ((1u << valid) - 1u);
        dst[idx] = word;
    }
}

}  // namespace acoss
