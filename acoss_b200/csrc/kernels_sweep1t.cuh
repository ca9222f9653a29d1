// kernels_sweep1t.cuh -- the general (float64) sweep for LONG column songs, cut into column tiles.
//
// kernels_sweep.cuh handles a column song of any length inside ONE CTA: beyond 1024 columns a thread owns
// several column groups, reloads their 48 frame values for every row, and keeps the prefix ring and the
// strip in a global scratch.  At the stress configuration (BASELINE configs[4]: 4096 frames, m = 20) that
// costs 18 cycles per cell per SM (profiles/r01_j_stress.json).  Here the column song is cut into tiles of
// 1024 frame columns, one CTA per (pair, orientation, tile):
//   * a thread owns 4 columns for the whole sweep (registers), the (m+1)-deep prefix ring of the tile lives
//     in shared memory;
//   * tiles overlap by m - 1 columns: the prefix sum along a diagonal restarts at the tile's left edge, so
//     every window that ends at local column >= m - 1 lies inside the tile (same sum as the untiled
//     sweep up to float64 rounding of the prefix, which no comparison depends on: separators sit in the
//     middle of the gap between the k-th and (k+1)-th neighbour);
//   * MODE 3 writes the windowed squared distances T (float64) of the tile's own columns to a per-pair
//     matrix in global memory, k_select_rows then selects every row's separator (csm_to_binary's
//     argpartition, CRPUtils.py:186-199) from a shared-memory copy of the row;
//   * MODE 1 compares T with the row and column separators and writes the tile's words of the packed CRP.
// Arithmetic, separators and CRP semantics are those of kernels_sweep.cuh (float64 throughout).
#pragma once
#include "kernels_sweep.cuh"

namespace acoss {

constexpr int kTileThreads = 256;
constexpr int kTileCols = kTileThreads * kSweepCPL;  // frame columns per tile
constexpr int kTileItemBits = 22;                    // item = (tile << 22) | (pair * 2 + orientation)  or  | pair

// output columns a tile owns (a multiple of 32 so that tiles write whole CRP words)
__host__ __device__ inline int tile_stride(int m) { return ((kTileCols - (m - 1)) / 32) * 32; }
__host__ __device__ inline int tile_count(int Lc, int m) {
    const int No = Lc - m + 1, S = tile_stride(m);
    return (No <= 0 || S <= 0) ? 0 : (No + S - 1) / S;
}
__host__ __device__ inline size_t tile_smem_bytes(int m, int bins) {
    size_t bitrows = ((size_t)2 * (kTileCols / 32 + 2) * sizeof(uint32_t) + 15) & ~(size_t)15;
    return bitrows + (size_t)kStageRows * (bins + 1) * sizeof(double) + (size_t)(m + 1) * kTileCols * sizeof(double);
}

struct TileArgs {
    const double* frames;
    const double* norms;
    const int64_t* offs;
    const PairWork* pairs;
    const uint32_t* items;
    int n_items;
    unsigned int* counter;
    double* seps;
    uint32_t* crp;
    double* tmat;  // per (pair, orientation): [Mo][No] windowed squared distances, at PairWork::tm_off[orientation]
    int m;
    uint32_t flags;
};

// ring layout: column c of a ring row sits at (c & 3) * 256 + (c >> 2), so that the four columns of a thread
// are 256 doubles apart and a warp's accesses to one of them are contiguous (no bank conflicts)
__device__ __forceinline__ int tile_ridx(int c) { return (c & 3) * kTileThreads + (c >> 2); }

template <int BINS, int MODE>
__global__ void __launch_bounds__(kTileThreads) k_sweep_tile(TileArgs A) {
    constexpr int CPL = kSweepCPL, THREADS = kTileThreads, W = kTileCols;
    constexpr int XS = BINS + 1;
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    ACOSS_DYNSMEM(smem_raw);
    __shared__ int s_item;
    const int tid = threadIdx.x, lane = tid & 31;
    const int m = A.m, RD = m + 1, S = tile_stride(m);

    uint32_t* bitrow = reinterpret_cast<uint32_t*>(smem_raw);
    constexpr int bitrow_words = W / 32 + 2;
    size_t off = ((size_t)2 * bitrow_words * sizeof(uint32_t) + 15) & ~(size_t)15;
    double* xs = reinterpret_cast<double*>(smem_raw + off);
    off += (size_t)kStageRows * XS * sizeof(double);
    double* ring = reinterpret_cast<double*>(smem_raw + off);

    for (;;) {
        if (tid == 0) s_item = (int)atomicAdd(A.counter, 1u);
        __syncthreads();
        const int it = s_item;
        __syncthreads();
        if (it >= A.n_items) break;

        const uint32_t item = A.items[it];
        const int tile = (int)(item >> kTileItemBits);
        const uint32_t low = item & ((1u << kTileItemBits) - 1u);
        int pidx, orient;
        if (MODE == 3) {
            pidx = (int)(low >> 1);
            orient = (int)(low & 1u);
        } else {
            pidx = (int)low;
            orient = A.pairs[pidx].swap;
        }
        const PairWork pw = A.pairs[pidx];
        const int rsong = orient ? pw.song_j : pw.song_i;
        const int csong = orient ? pw.song_i : pw.song_j;
        const int Lr = orient ? pw.Lj : pw.Li;
        const int Lc = orient ? pw.Li : pw.Lj;
        const int rshift = orient ? 0 : pw.oti;  // OTI rolls the first song of the pair
        const int cshift = orient ? pw.oti : 0;
        const int Mo = Lr - m + 1, No = Lc - m + 1;
        if (Mo <= 0 || No <= 0) continue;
        const int c_lo = tile * S;  // first frame column of the tile = first output column it owns
        if (c_lo >= No) continue;
        const double* rframes = A.frames + A.offs[rsong] * BINS;
        const double* cframes = A.frames + A.offs[csong] * BINS;
        const double* rnorms = A.norms + A.offs[rsong];
        const double* cnorms = A.norms + A.offs[csong];
        const int64_t rsep_off = pw.sep_off + (orient ? pw.M : 0);
        const int64_t csep_off = pw.sep_off + (orient ? 0 : pw.M);
        const bool nonmutual = (A.flags & kFlagNonMutual) != 0;
        double* tm = (MODE == 3) ? A.tmat + pw.tm_off[orient] : nullptr;

        // this thread's 4 columns: frames pre-scaled by -2 and rolled, norms, column separators
        const int lb0 = tid * CPL;  // local column
        double y2[CPL][BINS], ny[CPL], cs[CPL], Pprev[CPL];
        bool own[CPL];
#pragma unroll
        for (int r = 0; r < CPL; ++r) {
            const int b = c_lo + lb0 + r;
            Pprev[r] = 0.0;
            if (b < Lc) {
#pragma unroll
                for (int k = 0; k < BINS; ++k) {
                    int src = k - cshift;
                    if (src < 0) src += BINS;
                    y2[r][k] = -2.0 * cframes[(int64_t)b * BINS + src];
                }
                ny[r] = cnorms[b];
            } else {
#pragma unroll
                for (int k = 0; k < BINS; ++k) y2[r][k] = 0.0;
                ny[r] = 0.0;
            }
            const int jl = lb0 + r - (m - 1);  // output column within the tile
            own[r] = jl >= 0 && jl < S && c_lo + jl < No;
            cs[r] = -INF;
            if (MODE == 1 && own[r]) cs[r] = (nonmutual && !orient) ? INF : A.seps[csep_off + c_lo + jl];
        }

        int slot = 0;  // a % RD, kept by increment (RD is a runtime value: no integer division in the row loop)
        for (int a = 0; a < Lr; ++a) {
            if ((a & (kStageRows - 1)) == 0) {
                for (int idx = tid; idx < kStageRows * XS; idx += THREADS) {
                    const int rr = idx / XS, k = idx - rr * XS;
                    const int a2 = a + rr;
                    double v = 0.0;
                    if (a2 < Lr) {
                        if (k < BINS) {
                            int src = k - rshift;
                            if (src < 0) src += BINS;
                            v = rframes[(int64_t)a2 * BINS + src];
                        } else {
                            v = rnorms[a2];
                        }
                    }
                    xs[idx] = v;
                }
                __syncthreads();
            }
            const double* xr = xs + (a & (kStageRows - 1)) * XS;
            double x[BINS];
#pragma unroll
            for (int k = 0; k < BINS; ++k) x[k] = xr[k];
            const double nx = xr[BINS];

            const int i = a - (m - 1);  // output row finished by this row
            double rs = -INF;
            if (MODE == 1 && i >= 0) rs = (nonmutual && orient) ? INF : A.seps[rsep_off + i];

            const int slot1 = slot == 0 ? RD - 1 : slot - 1, slotm = slot + 1 == RD ? 0 : slot + 1;
            const double* ring1 = ring + (size_t)slot1 * W;  // row a-1
            const double* ringm = ring + (size_t)slotm * W;  // row a-m
            double* ring0 = ring + (size_t)slot * W;
            slot = slotm;

            double P[CPL], T[CPL];
#pragma unroll
            for (int r = 0; r < CPL; ++r) {
                double acc = nx + ny[r];
#pragma unroll
                for (int k = 0; k < BINS; ++k) acc = fma(x[k], y2[r][k], acc);
                const double C = fmax(acc, 0.0);
                const int lb = lb0 + r;
                double pd;
                if (r == 0)
                    pd = (a > 0 && lb > 0) ? ring1[tile_ridx(lb - 1)] : 0.0;
                else
                    pd = Pprev[r - 1];
                P[r] = pd + C;
                const double po = (a >= m && lb >= m) ? ringm[tile_ridx(lb - m)] : 0.0;
                T[r] = P[r] - po;
            }
#pragma unroll
            for (int r = 0; r < CPL; ++r) {
                ring0[r * kTileThreads + tid] = P[r];
                Pprev[r] = P[r];
            }
            if (MODE == 3) {
                if (i >= 0) {
                    double* trow = tm + (int64_t)i * No + c_lo + (lb0 - (m - 1));
#pragma unroll
                    for (int r = 0; r < CPL; ++r)
                        if (own[r]) trow[r] = T[r];
                }
            } else {
                unsigned nib = 0;
#pragma unroll
                for (int r = 0; r < CPL; ++r) nib |= ((T[r] <= rs) && (T[r] <= cs[r])) ? (1u << r) : 0u;
                unsigned v = nib << (4 * (lane & 7));
                v |= __shfl_xor_sync(kFullMask, v, 1);
                v |= __shfl_xor_sync(kFullMask, v, 2);
                v |= __shfl_xor_sync(kFullMask, v, 4);
                if ((lane & 7) == 0) bitrow[(a & 1) * bitrow_words + tid / 8] = v;
            }
            __syncthreads();

            if (MODE == 1 && i >= 0) {
                const int wp = pw.crp_wp;
                const unsigned sh = (unsigned)(m - 1) & 31u;
                const int wo = (m - 1) >> 5;
                const uint32_t* br = bitrow + (a & 1) * bitrow_words;
                uint32_t* out = A.crp + pw.crp_off + (int64_t)i * wp + c_lo / 32;
                const int nw = imin(S / 32, wp - c_lo / 32);  // words this tile owns in the row
                for (int w = tid; w < nw; w += THREADS) {
                    const uint32_t lo = (w + wo < W / 32) ? br[w + wo] : 0u;
                    const uint32_t hi = (w + wo + 1 < W / 32) ? br[w + wo + 1] : 0u;
                    out[w] = __funnelshift_r(lo, hi, sh);
                }
            }
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------- separators from the T matrix
struct SelArgs {
    const PairWork* pairs;
    const uint32_t* items;  // pair * 2 + orientation
    int n_items;
    int row_blocks;         // CTAs per item (blockIdx.x = item * row_blocks + block)
    const double* tmat;
    double* seps;
    int m;
    int stage_cols;         // > 0: dynamic shared memory holds kSelWarps rows of this many doubles
    uint32_t flags;         // debug: bit 23 = no secant steps in k_select_rows_cta (A/B)
};
constexpr int kSelWarps = 4;
constexpr int kSelRowsPerWarp = 8;

__global__ void __launch_bounds__(32 * kSelWarps) k_select_rows(SelArgs A) {
    ACOSS_DYNSMEM(smem_raw);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int it = (int)(blockIdx.x / (unsigned)A.row_blocks), rb = (int)(blockIdx.x % (unsigned)A.row_blocks);
    if (it >= A.n_items) return;
    const uint32_t item = A.items[it];
    const int pidx = (int)(item >> 1), orient = (int)(item & 1u);
    const PairWork* pw = A.pairs + pidx;
    const int Lr = orient ? pw->Lj : pw->Li, Lc = orient ? pw->Li : pw->Lj;
    const int Mo = Lr - A.m + 1, No = Lc - A.m + 1;
    if (Mo <= 0 || No <= 0) return;
    const int kk = orient ? pw->k_col : pw->k_row;
    const int64_t rsep_off = pw->sep_off + (orient ? pw->M : 0);
    const double* tm = A.tmat + pw->tm_off[orient];
    double* srow = reinterpret_cast<double*>(smem_raw) + (size_t)warp * A.stage_cols;
    const bool staged = A.stage_cols >= No;
    double guess = __longlong_as_double(0x7ff8000000000000LL);  // NaN: no guess yet
    const int i0 = (rb * kSelWarps + warp) * kSelRowsPerWarp;
    for (int i = i0; i < imin(i0 + kSelRowsPerWarp, Mo); ++i) {
        const double* row = tm + (int64_t)i * No;
        if (staged) {
            __syncwarp();
            for (int j = lane; j < No; j += 32) srow[j] = row[j];
            __syncwarp();
            row = srow;
        }
        const double sep = warp_select_sep(row, No, 1, kk, guess, lane);
        guess = sep;
        if (lane == 0) A.seps[rsep_off + i] = sep;
    }
}

// CTA-per-row selection: the row's values live in the registers of 256 threads (EPT each), every counting pass
// is register compares + one REDUX + one barrier.  Same separator as warp_select_sep: the midpoint between the
// k-th and (k+1)-th smallest, exact float64 ties across the k-th position left out.
constexpr int kSelThreads = 256;
constexpr int kSelRowsPerCta = 32;

__device__ __forceinline__ int cta_sum(int v, int* slots, int& par, int warp, int lane) {
    v = warp_sum(v);
    int* sl = slots + par * (kSelThreads / 32);
    if (lane == 0) sl[warp] = v;
    __syncthreads();
    int t = 0;
#pragma unroll
    for (int w = 0; w < kSelThreads / 32; ++w) t += sl[w];
    par ^= 1;  // the next reduction uses the other slot set: one barrier per reduction
    return t;
}
__device__ __forceinline__ void cta_minmax(double& lo, double& hi, double* slots, int& par, int warp, int lane) {
    lo = warp_min(lo);
    hi = warp_max(hi);
    double* sl = slots + par * 2 * (kSelThreads / 32);
    if (lane == 0) {
        sl[2 * warp] = lo;
        sl[2 * warp + 1] = hi;
    }
    __syncthreads();
#pragma unroll
    for (int w = 0; w < kSelThreads / 32; ++w) {
        lo = fmin(lo, sl[2 * w]);
        hi = fmax(hi, sl[2 * w + 1]);
    }
    par ^= 1;
}

template <int EPT>
__global__ void __launch_bounds__(kSelThreads) k_select_rows_cta(SelArgs A) {
    __shared__ int s_cnt[2 * (kSelThreads / 32)];
    __shared__ double s_mm[2 * 2 * (kSelThreads / 32)];
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int it = (int)(blockIdx.x / (unsigned)A.row_blocks), rb = (int)(blockIdx.x % (unsigned)A.row_blocks);
    if (it >= A.n_items) return;
    const uint32_t item = A.items[it];
    const int pidx = (int)(item >> 1), orient = (int)(item & 1u);
    const PairWork* pw = A.pairs + pidx;
    const int Lr = orient ? pw->Lj : pw->Li, Lc = orient ? pw->Li : pw->Lj;
    const int Mo = Lr - A.m + 1, No = Lc - A.m + 1;
    if (Mo <= 0 || No <= 0) return;
    const int k = orient ? pw->k_col : pw->k_row;
    const int64_t rsep_off = pw->sep_off + (orient ? pw->M : 0);
    const double* tm = A.tmat + pw->tm_off[orient];
    int pc = 0, pm = 0;
    const double NANV = __longlong_as_double(0x7ff8000000000000LL);
    double guess = NANV;  // no guess yet
    double slope = 0.0;   // distance per rank near the separator, carried from row to row
    const int i0 = rb * kSelRowsPerCta;
    for (int i = i0; i < imin(i0 + kSelRowsPerCta, Mo); ++i) {
        const double* row = tm + (int64_t)i * No;
        double v[EPT];
#pragma unroll
        for (int e = 0; e < EPT; ++e) {
            const int j = tid + e * kSelThreads;
            v[e] = (j < No) ? row[j] : NANV;  // NaN pads: never <= a threshold, ignored by fmin / fmax
        }
        double sep;
        if (k <= 0) {
            sep = -INF;
        } else if (k >= No) {
            sep = INF;
        } else {
            double lo = INF, hi = -INF;
#pragma unroll
            for (int e = 0; e < EPT; ++e) {
                lo = fmin(lo, v[e]);
                hi = fmax(hi, v[e]);
            }
            cta_minmax(lo, hi, s_mm, pm, warp, lane);
            auto count_le = [&](double t) {
                int c = 0;
#pragma unroll
                for (int e = 0; e < EPT; ++e) c += (v[e] <= t) ? 1 : 0;
                return cta_sum(c, s_cnt, pc, warp, lane);
            };
            int clo = count_le(lo), chi = No;
            double t = lo;
            if (clo > k) {
                sep = -INF;  // ties at the minimum straddle k
            } else {
                bool found = (clo == k);
                if (!found) {
                    const bool use_guess = (guess > lo) && (guess < hi);
                    // Secant steps on the LOCAL slope (distance per rank around the separator: from the last two
                    // counts of this row, else carried over from the previous row) before the bracket-wide
                    // interpolation: the values of a row are far from uniform between its minimum and maximum, so
                    // interpolating over the whole bracket needs ~7 counting passes where the local slope needs ~3.
                    double tp = 0.0, tpp = 0.0;   // thresholds of the last two counting passes of this row
                    int cp = -1, cpp = -1;        // and their counts
                    for (int iter = 0; iter < 400 && !found; ++iter) {
                        if (iter == 0 && use_guess) {
                            t = guess;
                        } else {
                            double tn = NANV;
                            if (!(A.flags & (1u << 23)) && iter < 6 && cp >= 0) {
                                if (cpp >= 0 && cp != cpp) slope = (tp - tpp) / (double)(cp - cpp);
                                if (slope > 0.0) tn = tp + ((double)(k - cp) + 0.5) * slope;
                            }
                            if (tn > lo && tn < hi) {
                                t = tn;
                            } else {
                                if (iter & 1)
                                    t = 0.5 * lo + 0.5 * hi;
                                else
                                    // the fraction only steers the search: single precision is plenty
                                    t = lo + (hi - lo) * (double)((float)(k - clo) * __frcp_rn((float)(chi - clo)));
                                if (!(t > lo && t < hi)) {
                                    t = 0.5 * lo + 0.5 * hi;
                                    if (!(t > lo && t < hi)) break;  // bracket collapsed: exact ties
                                }
                            }
                        }
                        const int c = count_le(t);
                        tpp = tp;
                        cpp = cp;
                        tp = t;
                        cp = c;
                        if (c == k) {
                            found = true;
                        } else if (c < k) {
                            lo = t;
                            clo = c;
                        } else {
                            hi = t;
                            chi = c;
                        }
                    }
                }
                if (!found) t = lo;
                double vk = INF, vk1 = INF;  // vk holds -max via min of negatives
                double nvk = INF;
#pragma unroll
                for (int e = 0; e < EPT; ++e) {
                    if (v[e] <= t)
                        nvk = fmin(nvk, -v[e]);
                    else
                        vk1 = fmin(vk1, v[e]);
                }
                double a = nvk, b = -vk1;  // reduce min(nvk) and min(vk1) = -max(-vk1) in one call
                cta_minmax(a, b, s_mm, pm, warp, lane);
                vk = -a;
                vk1 = -b;
                sep = vk + 0.5 * (vk1 - vk);
                if (!(sep < vk1)) sep = vk;
            }
        }
        guess = sep;
        if (tid == 0) A.seps[rsep_off + i] = sep;
    }
}

// ---------------------------------------------------------------- CRP straight from the stored T matrix
// When both orientations of a pair went through MODE 3 in the same wave, the CRP orientation's T matrix is
// still in memory when its separators arrive: comparing it with them is one streaming pass (8 bytes in, 1 bit
// out per cell) instead of a third sweep.  The stored T is exactly what a MODE 1 tile would recompute.
struct EmitArgs {
    const PairWork* pairs;
    const uint32_t* items;  // pair
    int n_items;
    int row_blocks;
    const double* tmat;
    const double* seps;
    uint32_t* crp;
    int m;
    uint32_t flags;
};
constexpr int kEmitThreads = 256;
constexpr int kEmitRowsPerWarp = 4;

__global__ void __launch_bounds__(kEmitThreads) k_emit_crp(EmitArgs A) {
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int it = (int)(blockIdx.x / (unsigned)A.row_blocks), rb = (int)(blockIdx.x % (unsigned)A.row_blocks);
    if (it >= A.n_items) return;
    const int pidx = (int)A.items[it];
    const PairWork* pw = A.pairs + pidx;
    const int orient = pw->swap;
    const int Lr = orient ? pw->Lj : pw->Li, Lc = orient ? pw->Li : pw->Lj;
    const int Mo = Lr - A.m + 1, No = Lc - A.m + 1;
    if (Mo <= 0 || No <= 0) return;
    const bool nonmutual = (A.flags & kFlagNonMutual) != 0;
    const double* tm = A.tmat + pw->tm_off[orient];
    const double* rsep = A.seps + pw->sep_off + (orient ? pw->M : 0);
    const double* csep = A.seps + pw->sep_off + (orient ? 0 : pw->M);
    const int wp = pw->crp_wp;
    const int i0 = (rb * (kEmitThreads / 32) + warp) * kEmitRowsPerWarp;
    for (int i = i0; i < imin(i0 + kEmitRowsPerWarp, Mo); ++i) {
        const double rs = (nonmutual && orient) ? INF : rsep[i];
        const double* row = tm + (int64_t)i * No;
        uint32_t* out = A.crp + pw->crp_off + (int64_t)i * wp;
        uint32_t mine = 0;
        for (int w = 0; w < wp; ++w) {
            const int j = 32 * w + lane;
            bool bit = false;
            if (j < No) {
                const double t = row[j];
                const double cs = (nonmutual && !orient) ? INF : csep[j];
                bit = (t <= rs) && (t <= cs);
            }
            const uint32_t word = __ballot_sync(kFullMask, bit);
            if ((w & 31) == lane) mine = word;
            if ((w & 31) == 31 || w == wp - 1) {
                const int wb = w & ~31;
                if (wb + lane <= w) out[wb + lane] = mine;  // 32 words at a time, coalesced
            }
        }
    }
}

}  // namespace acoss
