// api.cu -- host side of libacoss_b200.so: the C ABI declared in include/acoss_b200.h.
// Owns device memory and the stream, cuts the pair list into chunks whose scratch
// (separators, packed CRPs) fits a fixed budget, and launches the kernels:
//   k_oti -> k_sweep<MODE 0> (both orientations) -> k_sweep<MODE 1> -> k_dp.
// There is no CPU path: every entry point fails with ACOSS_ERR_CUDA when the device is
// missing or a launch fails.
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

#include "../../include/acoss_b200.h"
#include "common.cuh"
#include "kernels_dense.cuh"
#include "kernels_dp.cuh"
#include "kernels_dp16.cuh"
#include "kernels_prep.cuh"
#include "kernels_sweep.cuh"
#include "kernels_sweep1t.cuh"
#include "kernels_sweep2.cuh"
#include "kernels_sweep4.cuh"
#include "kernels_feat.cuh"
#include "kernels_eval.cuh"

#include "ctx.h"
#include "sweep7.h"

using namespace acoss;

namespace {

int neighbours(double kappa, int n) {
    // CRPUtils.py:188-193: int(np.round(kappa*n)) (round-half-even on the double product)
    if (kappa < 1.0) return (int)std::nearbyint(kappa * (double)n);
    return (int)kappa;
}


constexpr int64_t kMaxChunkSepDoubles = (int64_t)1 << 28;  // 2 GiB
constexpr int64_t kMaxChunkCrpWords = (int64_t)1 << 30;    // 4 GiB
constexpr int64_t kMaxChunkNbBytes = (int64_t)6 << 30;     // 6 GiB of nibble-byte CRPs (gen 4)
constexpr int64_t kMaxWaveTmatDoubles = (int64_t)4 << 30;  // 32 GiB of T matrices per wave of the tiled long-song sweep
constexpr int kCounterSlots = 96;
constexpr int kG4Buckets = kS4MaxWarps;            // gen-4 sweep: bucket b = b + 1 warps
constexpr int kFastBuckets = kFastMaxCols / 128;  // bucket b: T = 32 * (b + 1) threads, up to 128 * (b + 1) columns

struct ItemLists {
    std::vector<uint32_t> fast0[kFastBuckets], fast2[kFastBuckets];  // fast sweep: separators / separators + CRP
    std::vector<uint32_t> wide0, wide1;                              // general sweep (long songs): MODE 0 / MODE 1
    std::vector<uint32_t> t3, tsel, t1;                              // tiled long-song sweep: T tiles, row selection, CRP tiles
    int tsel_rows = 0, tsel_cols = 0;                                // largest Mo / No among tsel
    // The T matrices are scratch between a tile sweep and its row selection only, so the chunk's tiled items run
    // in WAVES that reuse one buffer (the DP still sees the whole chunk: one warp per pair needs many pairs).
    std::vector<size_t> wave_t3, wave_tsel, wave_te;                 // first item of every wave in t3 / tsel / te
    std::vector<uint32_t> te;                                        // pairs whose CRP comes from the stored T matrix
    int te_rows = 0;
    std::vector<uint32_t> g4_0[kG4Buckets], g4_2[kG4Buckets];    // gen-4 sweep: separators / CRP
    std::vector<uint32_t> g4_pack;                                   // pairs whose nibble CRP needs packing
    std::vector<uint32_t> g7_0[kSweep7Buckets], g7_2[kSweep7Buckets];  // gen-7 sweep by column capacity
    bool have_keys = false;                                          // some pair uses the integer-key sweeps
    std::vector<uint32_t> dp4, dp8, dp16;
    bool dp_packed_ok = true;                                        // every plot fits the 16-bit lanes of k_dp16
    int wide0_w = 0, wide1_w = 0;
};

// One chunk of a pair list, prepared on the host: work records, item lists, scratch totals and the pinned staging
// image [PairWork...][items...].  Preparing chunk k + 1 overlaps the kernels of chunk k (acoss_serra09_pairs).
struct ChunkPlan {
    std::vector<PairWork> W;
    ItemLists L;
    int64_t p0 = 0;
    int n = 0;
    int64_t sep_total = 0, crp_total = 0, bnd_total = 0, nb_total = 0, tmat_peak = 0;
    size_t n_items = 0, pw_bytes = 0, res_off = 0;   // res_off: float results [q | d | sw] of n each in the staging image
    PinBuf* stage = nullptr;
    size_t o_fast0[kFastBuckets], o_fast2[kFastBuckets], o_wide0, o_wide1, o_dp4, o_dp8, o_dp16;
    size_t o_g40[kG4Buckets], o_g42[kG4Buckets], o_g4pack;
    size_t o_g70[kSweep7Buckets], o_g72[kSweep7Buckets], o_t3, o_tsel, o_t1, o_te;
};

template <int BINS, int MODE>
int launch_sweep4(acoss_ctx* c, const uint32_t* d_items, int n_items, int bucket, int m, unsigned int* d_counter,
                  uint32_t flags) {
    if (n_items <= 0) return ACOSS_OK;
    const int T = 32 * (bucket + 1);
    auto kfn = k_sweep4<BINS, MODE, 0, 0>;
    if (m == 9) {
        switch (bucket + 1) {  // register-resident keys for CTAs of up to 5 warps
            case 1: kfn = k_sweep4<BINS, MODE, 9, 1>; break;
            case 2: kfn = k_sweep4<BINS, MODE, 9, 2>; break;
            case 3: kfn = k_sweep4<BINS, MODE, 9, 3>; break;
            case 4: kfn = k_sweep4<BINS, MODE, 9, 4>; break;
            case 5: kfn = k_sweep4<BINS, MODE, 9, 5>; break;
            default: kfn = k_sweep4<BINS, MODE, 9, 0>; break;
        }
    }
    const size_t smem = (size_t)s4_layout(T, m, BINS).total;
    if (smem > (size_t)c->max_smem - 1024) return fail(ACOSS_ERR_ARG, "window m=%d too large for the gen-4 sweep", m);
    CU(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kfn, T, smem));
    if (occ < 1) return fail(ACOSS_ERR_CUDA, "k_sweep4 cannot be resident (T %d, smem %zu B)", T, smem);
    const int grid = (int)std::min<int64_t>((int64_t)n_items, (int64_t)c->sm_count * occ);
    Sweep4Args A;
    memset(&A, 0, sizeof(A));
    A.frames = c->frames.as<double>();
    A.wnorm = c->wnorm.as<double>();
    A.offs = c->offs.as<int64_t>();
    A.pairs = c->pairs.as<PairWork>();
    A.items = d_items;
    A.n_items = n_items;
    A.counter = d_counter;
    A.sepk = c->sepk.as<int32_t>();
    A.sepx = c->sepx.as<double>();
    A.nb = c->nb.as<uint8_t>();
    A.m = m;
    A.gap = c->opt_gap;
    A.flags = flags;
    ACOSS_LAUNCH(kfn, grid, T, smem, c->stream, A);
    c->launches++;
    CU(cudaGetLastError());
    return ACOSS_OK;
}

template <int BINS, int MODE>
int launch_sweep_fast(acoss_ctx* c, const uint32_t* d_items, int n_items, int bucket, int m, unsigned int* d_counter,
                      uint32_t flags) {
    if (n_items <= 0) return ACOSS_OK;
    // windows with a compiled-in specialisation: the reference default (Serra09.py:87) and the
    // long-song stress configuration; anything else takes the runtime-m instantiation
    auto kfn = (m == 9) ? k_sweep_fast<BINS, MODE, 9> : (m == 20 ? k_sweep_fast<BINS, MODE, 20> : k_sweep_fast<BINS, MODE, 0>);
    const int T = 32 * (bucket + 1);
    const int R = std::min(T / 32, kFastMaxStripRows);
    const size_t smem = fast_smem_bytes(T, m, R);
    if (smem > (size_t)c->max_smem - 1024) return fail(ACOSS_ERR_ARG, "window m=%d too large for the fast sweep", m);
    CU(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kfn, T, smem));
    if (occ < 1) return fail(ACOSS_ERR_CUDA, "k_sweep_fast cannot be resident (T %d, smem %zu B)", T, smem);
    const int grid = (int)std::min<int64_t>((int64_t)n_items, (int64_t)c->sm_count * occ);
    SweepArgs A;
    memset(&A, 0, sizeof(A));
    A.frames = c->frames.as<double>();
    A.norms = c->norms.as<double>();
    A.offs = c->offs.as<int64_t>();
    A.pairs = c->pairs.as<PairWork>();
    A.items = d_items;
    A.n_items = n_items;
    A.counter = d_counter;
    A.seps = c->seps.as<double>();
    A.crp = c->crp.as<uint32_t>();
    A.m = m;
    A.R = R;
    A.Wmax = 4 * T;
    A.flags = flags;
    ACOSS_LAUNCH(kfn, grid, T, smem, c->stream, A);
    c->launches++;
    CU(cudaGetLastError());
    return ACOSS_OK;
}

template <int BINS, int THREADS, int MODE>
int launch_sweep(acoss_ctx* c, const uint32_t* d_items, int n_items, int Wmax, int m, unsigned int* d_counter,
                 uint32_t flags) {
    if (n_items <= 0) return ACOSS_OK;
    auto kfn = k_sweep<BINS, THREADS, MODE>;
    const int R = 8;
    const int RD = m + 1;
    const size_t fixed = sweep_smem_fixed_bytes(Wmax, BINS);
    const size_t var = (size_t)(RD + (MODE == 0 ? R : 0)) * Wmax * sizeof(double);
    const size_t budget = (size_t)c->max_smem - 1024;  // static shared + launch reserve
    const bool use_g = (fixed + var > budget) || c->opt_force_gscratch;
    const size_t smem = use_g ? fixed : fixed + var;
    CU(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kfn, THREADS, smem));
    if (occ < 1) return fail(ACOSS_ERR_CUDA, "k_sweep cannot be resident (smem %zu B)", smem);
    const int grid = std::min<int64_t>((int64_t)n_items, (int64_t)c->sm_count * occ);
    SweepArgs A;
    A.frames = c->frames.as<double>();
    A.norms = c->norms.as<double>();
    A.offs = c->offs.as<int64_t>();
    A.pairs = c->pairs.as<PairWork>();
    A.items = d_items;
    A.n_items = n_items;
    A.counter = d_counter;
    A.seps = c->seps.as<double>();
    A.crp = c->crp.as<uint32_t>();
    A.m = m;
    A.R = R;
    A.Wmax = Wmax;
    A.use_gscratch = use_g ? 1 : 0;
    A.gscratch = nullptr;
    A.gscratch_stride = 0;
    A.flags = flags;
    if (use_g) {
        A.gscratch_stride = (int64_t)(RD + R) * Wmax;
        CU(c->gscratch.reserve((size_t)grid * A.gscratch_stride * sizeof(double)));
        A.gscratch = c->gscratch.as<double>();
    }
    ACOSS_LAUNCH(kfn, grid, THREADS, smem, c->stream, A);
    c->launches++;
    CU(cudaGetLastError());
    return ACOSS_OK;
}

template <int BINS, int MODE>
int launch_sweep_tile(acoss_ctx* c, const uint32_t* d_items, int n_items, int m, unsigned int* d_counter, uint32_t flags) {
    if (n_items <= 0) return ACOSS_OK;
    auto kfn = k_sweep_tile<BINS, MODE>;
    const size_t smem = tile_smem_bytes(m, BINS);
    CU(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kfn, kTileThreads, smem));
    if (occ < 1) return fail(ACOSS_ERR_CUDA, "k_sweep_tile cannot be resident (smem %zu B)", smem);
    const int grid = (int)std::min<int64_t>((int64_t)n_items, (int64_t)c->sm_count * occ);
    TileArgs A;
    memset(&A, 0, sizeof(A));
    A.frames = c->frames.as<double>();
    A.norms = c->norms.as<double>();
    A.offs = c->offs.as<int64_t>();
    A.pairs = c->pairs.as<PairWork>();
    A.items = d_items;
    A.n_items = n_items;
    A.counter = d_counter;
    A.seps = c->seps.as<double>();
    A.crp = c->crp.as<uint32_t>();
    A.tmat = c->tmat.as<double>();
    A.m = m;
    A.flags = flags;
    ACOSS_LAUNCH(kfn, grid, kTileThreads, smem, c->stream, A);
    c->launches++;
    CU(cudaGetLastError());
    return ACOSS_OK;
}

int launch_select_rows(acoss_ctx* c, const uint32_t* d_items, int n_items, int max_rows, int max_cols, int m) {
    if (n_items <= 0) return ACOSS_OK;
    SelArgs A;
    memset(&A, 0, sizeof(A));
    A.pairs = c->pairs.as<PairWork>();
    A.items = d_items;
    A.n_items = n_items;
    A.row_blocks = (max_rows + kSelWarps * kSelRowsPerWarp - 1) / (kSelWarps * kSelRowsPerWarp);
    A.tmat = c->tmat.as<double>();
    A.seps = c->seps.as<double>();
    A.m = m;
    A.flags = c->opt_debug_flags;
    if (max_cols <= 32 * kSelThreads) {
        // the row fits the registers of one CTA: CTA-per-row selection
        A.row_blocks = (max_rows + kSelRowsPerCta - 1) / kSelRowsPerCta;
        const int64_t grid = (int64_t)n_items * A.row_blocks;
        if (grid > 0x7fffffffLL) return fail(ACOSS_ERR_ARG, "too many rows for one selection launch");
        if (max_cols <= 4 * kSelThreads) { ACOSS_LAUNCH(k_select_rows_cta<4>, (int)grid, kSelThreads, 0, c->stream, A); }
        else if (max_cols <= 16 * kSelThreads) { ACOSS_LAUNCH(k_select_rows_cta<16>, (int)grid, kSelThreads, 0, c->stream, A); }
        else { ACOSS_LAUNCH(k_select_rows_cta<32>, (int)grid, kSelThreads, 0, c->stream, A); }
        c->launches++;
        CU(cudaGetLastError());
        return ACOSS_OK;
    }
    // longer rows: one warp per row; rows are staged in shared memory when four of them fit (otherwise the passes
    // read global memory)
    size_t smem = (size_t)kSelWarps * (size_t)max_cols * sizeof(double);
    A.stage_cols = max_cols;
    if (smem > (size_t)c->max_smem - 1024) {
        smem = 0;
        A.stage_cols = 0;
    }
    auto kfn = k_select_rows;
    CU(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t grid = (int64_t)n_items * A.row_blocks;
    if (grid > 0x7fffffffLL) return fail(ACOSS_ERR_ARG, "too many rows for one selection launch");
    ACOSS_LAUNCH(kfn, (int)grid, 32 * kSelWarps, smem, c->stream, A);
    c->launches++;
    CU(cudaGetLastError());
    return ACOSS_OK;
}

int launch_emit_crp(acoss_ctx* c, const uint32_t* d_items, int n_items, int max_rows, int m, uint32_t flags) {
    if (n_items <= 0) return ACOSS_OK;
    EmitArgs A;
    memset(&A, 0, sizeof(A));
    A.pairs = c->pairs.as<PairWork>();
    A.items = d_items;
    A.n_items = n_items;
    const int rows_per_cta = (kEmitThreads / 32) * kEmitRowsPerWarp;
    A.row_blocks = (max_rows + rows_per_cta - 1) / rows_per_cta;
    A.tmat = c->tmat.as<double>();
    A.seps = c->seps.as<double>();
    A.crp = c->crp.as<uint32_t>();
    A.m = m;
    A.flags = flags;
    const int64_t grid = (int64_t)n_items * A.row_blocks;
    if (grid > 0x7fffffffLL) return fail(ACOSS_ERR_ARG, "too many rows for one emission launch");
    ACOSS_LAUNCH(k_emit_crp, (int)grid, kEmitThreads, 0, c->stream, A);
    c->launches++;
    CU(cudaGetLastError());
    return ACOSS_OK;
}

template <int CPL>
int launch_dp(acoss_ctx* c, const uint32_t* d_items, int n_items, uint32_t flags, float* q, float* d, float* sw,
              bool packed_ok) {
    if (n_items <= 0) return ACOSS_OK;
    DpArgs A;
    A.pairs = c->pairs.as<PairWork>();
    A.items = d_items;
    A.n_items = n_items;
    A.crp = c->crp.as<uint32_t>();
    A.bnd = c->bnd.as<int4>();
    A.qmax_out = q;
    A.dmax_out = d;
    A.sw_out = sw;
    A.flags = flags;
    const int grid = (n_items + 3) / 4;
    if (flags & kFlagWantSw) {
        auto kfn = k_dp<CPL, true>;
        ACOSS_LAUNCH(kfn, grid, 128, 0, c->stream, A);
    } else if (packed_ok && c->opt_dp_packed) {
        auto kfn = k_dp16<CPL>;   // Qmax and Dmax of a cell in one register, DPX s16x2 instructions
        ACOSS_LAUNCH(kfn, grid, 128, 0, c->stream, A);
    } else {
        auto kfn = k_dp<CPL, false>;
        ACOSS_LAUNCH(kfn, grid, 128, 0, c->stream, A);
    }
    c->launches++;
    CU(cudaGetLastError());
    return ACOSS_OK;
}

// Upload item lists, run the DP kernels over the PairWork records already on the device.
int run_dp_lists(acoss_ctx* c, const ItemLists& L, const uint32_t* d_items_base, size_t off4, size_t off8,
                 size_t off16, uint32_t flags) {
    float* q = c->qout.as<float>();
    float* d = c->dout.as<float>();
    float* sw = c->swout.as<float>();
    int rc;
    if ((rc = launch_dp<4>(c, d_items_base + off4, (int)L.dp4.size(), flags, q, d, sw, L.dp_packed_ok)) != ACOSS_OK) return rc;
    if ((rc = launch_dp<8>(c, d_items_base + off8, (int)L.dp8.size(), flags, q, d, sw, L.dp_packed_ok)) != ACOSS_OK) return rc;
    if ((rc = launch_dp<16>(c, d_items_base + off16, (int)L.dp16.size(), flags, q, d, sw, L.dp_packed_ok)) != ACOSS_OK) return rc;
    return ACOSS_OK;
}

int dp_cpl(const acoss_ctx* c, int cols) {
    int cpl = cols <= 128 ? 4 : (cols <= 256 ? 8 : 16);
    return std::min(cpl, c->opt_dp_max_cpl);
}

void classify_dp(const acoss_ctx* c, ItemLists& L, uint32_t p, int rows, int cols) {
    if (rows + cols > kDp16MaxSide || cols > kDp16MaxCols) L.dp_packed_ok = false;
    const int cpl = dp_cpl(c, cols);
    (cpl == 4 ? L.dp4 : (cpl == 8 ? L.dp8 : L.dp16)).push_back(p);
}

int64_t bnd_records(const acoss_ctx* c, int rows, int cols) {
    // plots wider than one 32*CPL-column block need the hand-over scratch (2 buffers x 2 int4 per row)
    return cols > 32 * dp_cpl(c, cols) ? (int64_t)rows * 2 * 2 : 0;
}

// Host side of a chunk: work records + item lists + the staging image.  Touches no stream.
template <int BINS>
int prep_chunk(acoss_ctx* c, const int32_t* pairs, int64_t p0, int n, int m, double kappa, uint32_t flags,
               PinBuf* stage, ChunkPlan& P) {
    P = ChunkPlan();
    P.p0 = p0;
    P.n = n;
    P.stage = stage;
    P.W.resize((size_t)n);
    std::vector<PairWork>& W = P.W;
    ItemLists& L = P.L;
    int64_t sep_total = 0, crp_total = 0, bnd_total = 0, nb_total = 0, tmat_total = 0, tmat_peak = 0;
    const bool tiled_ok = c->opt_tiled_wide && !c->opt_force_gscratch && tile_stride(m) >= 32 &&
                          tile_smem_bytes(m, BINS) <= (size_t)c->max_smem - 1024 && n < (1 << (kTileItemBits - 1));
    const bool gen4_ok = c->opt_sweep_gen >= 4 && BINS == 12 && !c->opt_force_wide && s4_own_lanes(m) >= 16 &&
                         (size_t)s4_layout(32 * kG4Buckets, m, BINS).total <= (size_t)c->max_smem - 1024;
    const bool g7_ok = c->opt_sweep_gen >= 7 && m == 9 && (BINS == 12 || BINS == 13) && !c->opt_force_wide;
    const bool noswap = (flags & kFlagNoSwap) != 0;
    const bool nonmutual = (flags & kFlagNonMutual) != 0;
    for (int p = 0; p < n; ++p) {
        PairWork& w = W[p];
        w.song_i = pairs[2 * (p0 + p)];
        w.song_j = pairs[2 * (p0 + p) + 1];
        w.Li = c->h_len[w.song_i];
        w.Lj = c->h_len[w.song_j];
        w.M = w.Li - m + 1;
        w.N = w.Lj - m + 1;
        const bool ok = w.M > 0 && w.N > 0;
        w.k_row = ok ? neighbours(kappa, w.N) : 0;
        w.k_col = ok ? neighbours(kappa, w.M) : 0;
        w.oti = 0;
        w.swap = (!noswap && w.Lj > w.Li) ? 1 : 0;
        w.crp_rows = ok ? (w.swap ? w.N : w.M) : 0;
        w.crp_cols = ok ? (w.swap ? w.M : w.N) : 0;
        w.crp_wp = (w.crp_cols + 31) / 32;
        w.expo = 1;
        w.nb_off = 0;
        w.nb_stride = 0;
        w.pad1 = 0;
        w.tm_off[0] = w.tm_off[1] = -1;
        w.sep_off = sep_total;
        w.crp_off = crp_total;
        const int64_t br = bnd_records(c, w.crp_rows, w.crp_cols);
        if (br > 0 && std::min(w.crp_rows, w.crp_cols) > 32767)
            return fail(ACOSS_ERR_ARG, "plot %d x %d: the DP hand-over packs Qmax values in 16 bits (min side <= 32767)", w.crp_rows, w.crp_cols);
        w.bnd_off = br > 0 ? bnd_total : -1;
        bnd_total += br;
        if (ok) {
            sep_total += (int64_t)w.M + w.N;
            crp_total += (int64_t)w.crp_rows * w.crp_wp;
            // Two sweeps per pair.  The CRP / DP orientation dpo has rows = the longer song; the
            // other orientation supplies the separators of dpo's columns.
            const int dpo = w.swap, other = 1 - dpo;
            const int wB = 256 * kSweepCPL;
            auto Lc_of = [&](int o) { return o ? w.Li : w.Lj; };
            // gen 7 per orientation where the column song fits; both orientations of a pair must be served by
            // integer-key kernels (gens 4-7 share the separator format), which for 13 bins means gen 7 twice
            auto b7 = [&](int Lc) {
                int b = sweep7_bucket_for(Lc);
                if ((c->opt_debug_flags & (1u << 21)) && (b == 2 || b == 4 || b == 6)) ++b;  // A/B: capacities 128, 256, 512, 768, 1024 only
                return (g7_ok && b >= 0 && acoss_sweep7_smem(b, BINS, 2) <= (size_t)c->max_smem) ? b : -1;
            };
            const bool keys12 = gen4_ok && std::max(w.Li, w.Lj) <= s4_max_cols(m);
            const bool keys7 = b7(w.Li) >= 0 && b7(w.Lj) >= 0;
            if (keys12 || keys7) {
                L.have_keys = true;
                const int own4 = s4_own_cols(m);
                const int b0 = (Lc_of(other) - 1) / own4, b2 = (Lc_of(dpo) - 1) / own4;  // compute warps needed - 1
                if (!(nonmutual && dpo == 0)) {
                    if (b7(Lc_of(other)) >= 0)
                        L.g7_0[b7(Lc_of(other))].push_back((uint32_t)p * 2 + other);
                    else
                        L.g4_0[b0].push_back((uint32_t)p * 2 + other);
                }
                if (b7(Lc_of(dpo)) >= 0) {
                    L.g7_2[b7(Lc_of(dpo))].push_back((uint32_t)p);  // writes packed CRP words itself
                } else {
                    L.g4_2[b2].push_back((uint32_t)p);
                    w.nb_stride = 32 * (b2 + 1);
                    L.g4_pack.push_back((uint32_t)p);
                    w.nb_off = nb_total;
                    nb_total += ((int64_t)w.crp_rows * w.nb_stride + 15) & ~(int64_t)15;
                }
                classify_dp(c, L, (uint32_t)p, w.crp_rows, w.crp_cols);
                continue;
            }
            const bool fast_ok = !c->opt_force_wide && (size_t)fast_smem_bytes(32 * kFastBuckets, m, kFastMaxStripRows) <= (size_t)c->max_smem - 1024;
            // separators of orientation o through the general sweep: tiled (T matrix + row selection) or one CTA
            auto push_wide0 = [&](int o) {
                const int Lc = Lc_of(o);
                const int Mo = o ? w.N : w.M, No = o ? w.M : w.N;
                if (tiled_ok && (int64_t)Mo * No <= kMaxWaveTmatDoubles) {
                    if (L.wave_t3.empty() || tmat_total + (int64_t)Mo * No > kMaxWaveTmatDoubles) {
                        L.wave_t3.push_back(L.t3.size());
                        L.wave_tsel.push_back(L.tsel.size());
                        L.wave_te.push_back(L.te.size());
                        tmat_total = 0;
                    }
                    tmat_peak = std::max(tmat_peak, tmat_total + (int64_t)Mo * No);
                    for (int t = 0; t < tile_count(Lc, m); ++t)
                        L.t3.push_back(((uint32_t)t << kTileItemBits) | ((uint32_t)p * 2 + o));
                    L.tsel.push_back((uint32_t)p * 2 + o);
                    L.tsel_rows = std::max(L.tsel_rows, Mo);
                    L.tsel_cols = std::max(L.tsel_cols, No);
                    w.tm_off[o] = tmat_total;
                    tmat_total += (int64_t)Mo * No;
                } else {
                    L.wide0.push_back((uint32_t)p * 2 + o);
                    L.wide0_w = std::max(L.wide0_w, (Lc + wB - 1) / wB * wB);
                }
            };
            size_t wave_other = 0;  // wave of the other orientation's tiled item
            if (!(nonmutual && dpo == 0)) {  // column separators of the CRP
                const int Lc = Lc_of(other);
                if (fast_ok && Lc <= kFastMaxCols)
                    L.fast0[(Lc - 1) / 128].push_back((uint32_t)p * 2 + other);
                else
                    push_wide0(other);
                if (w.tm_off[other] >= 0) wave_other = L.wave_t3.size();
            }
            {
                const int Lc = Lc_of(dpo);
                if (fast_ok && Lc <= kFastMaxCols) {
                    L.fast2[(Lc - 1) / 128].push_back((uint32_t)p);
                } else {
                    if (!(nonmutual && dpo == 1)) push_wide0(dpo);
                    // The CRP straight from the stored T matrix when its separators are complete inside the same
                    // wave: the column separators come from a tiled item of this wave, or from a fast / one-CTA
                    // sweep (those run before the waves); the row separators from the selection just pushed.
                    const bool cols_ready = w.tm_off[other] < 0 || wave_other == L.wave_t3.size();
                    if (tiled_ok && w.tm_off[dpo] >= 0 && cols_ready) {
                        L.te.push_back((uint32_t)p);
                        L.te_rows = std::max(L.te_rows, w.crp_rows);
                    } else if (tiled_ok) {
                        for (int t = 0; t < tile_count(Lc, m); ++t) L.t1.push_back(((uint32_t)t << kTileItemBits) | (uint32_t)p);
                    } else {
                        L.wide1.push_back((uint32_t)p);
                        L.wide1_w = std::max(L.wide1_w, (Lc + wB - 1) / wB * wB);
                    }
                }
            }
        }
        classify_dp(c, L, (uint32_t)p, w.crp_rows, w.crp_cols);
    }
    size_t n_items = L.wide0.size() + L.wide1.size() + L.dp4.size() + L.dp8.size() + L.dp16.size() + L.g4_pack.size() +
                     L.t3.size() + L.tsel.size() + L.t1.size() + L.te.size();
    for (int b = 0; b < kFastBuckets; ++b) n_items += L.fast0[b].size() + L.fast2[b].size();
    for (int b = 0; b < kG4Buckets; ++b) n_items += L.g4_0[b].size() + L.g4_2[b].size();
    for (int b = 0; b < kSweep7Buckets; ++b) n_items += L.g7_0[b].size() + L.g7_2[b].size();
    // stage: [PairWork...][items...]
    const size_t pw_bytes = (size_t)n * sizeof(PairWork);
    P.pw_bytes = pw_bytes;
    // (results come back into the same pinned image: a device-to-host copy into the caller's pageable arrays would
    // block the issuing thread until the chunk's kernels are done, and with it the preparation of the next chunk)
    P.res_off = (pw_bytes + n_items * sizeof(uint32_t) + 63) & ~(size_t)63;
    CU(stage->reserve(P.res_off + 3 * (size_t)n * sizeof(float) + 64));
    memcpy(stage->p, W.data(), pw_bytes);
    uint32_t* hi = reinterpret_cast<uint32_t*>(stage->as<char>() + pw_bytes);
    size_t pos = 0;
    auto put = [&](const std::vector<uint32_t>& v) {
        size_t at = pos;
        if (!v.empty()) memcpy(hi + pos, v.data(), v.size() * sizeof(uint32_t));
        pos += v.size();
        return at;
    };
    for (int b = 0; b < kFastBuckets; ++b) P.o_fast0[b] = put(L.fast0[b]);
    for (int b = 0; b < kFastBuckets; ++b) P.o_fast2[b] = put(L.fast2[b]);
    P.o_wide0 = put(L.wide0);
    P.o_wide1 = put(L.wide1);
    P.o_t3 = put(L.t3);
    P.o_tsel = put(L.tsel);
    P.o_t1 = put(L.t1);
    P.o_te = put(L.te);
    P.o_dp4 = put(L.dp4);
    P.o_dp8 = put(L.dp8);
    P.o_dp16 = put(L.dp16);
    for (int b = 0; b < kG4Buckets; ++b) P.o_g40[b] = put(L.g4_0[b]);
    for (int b = 0; b < kG4Buckets; ++b) P.o_g42[b] = put(L.g4_2[b]);
    P.o_g4pack = put(L.g4_pack);
    for (int b = 0; b < kSweep7Buckets; ++b) P.o_g70[b] = put(L.g7_0[b]);
    for (int b = 0; b < kSweep7Buckets; ++b) P.o_g72[b] = put(L.g7_2[b]);
    P.n_items = n_items;
    P.sep_total = sep_total;
    P.crp_total = crp_total;
    P.bnd_total = bnd_total;
    P.nb_total = nb_total;
    P.tmat_peak = tmat_peak;
    return ACOSS_OK;
}

// Device side of a prepared chunk: scratch, staging copies, kernels, result copies.  Asynchronous on c->stream.
template <int BINS>
int exec_chunk(acoss_ctx* c, ChunkPlan& P, int m, uint32_t flags, const float* qmax_raw, const float* dmax_raw,
               const float* sw_raw, bool want_oti, bool want_crp) {
    const ItemLists& L = P.L;
    const int n = P.n;
    const int64_t sep_total = P.sep_total, crp_total = P.crp_total, bnd_total = P.bnd_total,
                  nb_total = P.nb_total, tmat_peak = P.tmat_peak;
    const size_t n_items = P.n_items, pw_bytes = P.pw_bytes;
    const uint32_t* hi = reinterpret_cast<const uint32_t*>(P.stage->as<char>() + pw_bytes);
    // ---- device buffers
    CU(c->pairs.reserve((size_t)n * sizeof(PairWork)));
    CU(c->seps.reserve((size_t)std::max<int64_t>(sep_total, 1) * sizeof(double)));
    CU(c->crp.reserve((size_t)std::max<int64_t>(crp_total, 1) * sizeof(uint32_t)));
    CU(c->bnd.reserve((size_t)std::max<int64_t>(bnd_total, 1) * sizeof(int4)));
    CU(c->qout.reserve((size_t)n * sizeof(float)));
    CU(c->dout.reserve((size_t)n * sizeof(float)));
    CU(c->swout.reserve((size_t)n * sizeof(float)));
    CU(c->otiout.reserve((size_t)n * sizeof(int32_t)));
    CU(c->counters.reserve(kCounterSlots * sizeof(unsigned int)));
    if (L.have_keys) {
        CU(c->sepk.reserve((size_t)std::max<int64_t>(sep_total, 1) * sizeof(int32_t)));
        CU(c->sepx.reserve((size_t)std::max<int64_t>(sep_total, 1) * sizeof(double)));
        CU(c->nb.reserve((size_t)std::max<int64_t>(nb_total, 16)));
    }
    if (tmat_peak > 0) CU(c->tmat.reserve((size_t)tmat_peak * sizeof(double)));
    CU(c->items.reserve(std::max<size_t>(n_items, 1) * sizeof(uint32_t)));
    CU(cudaMemcpyAsync(c->pairs.p, P.stage->p, pw_bytes, cudaMemcpyHostToDevice, c->stream));
    if (n_items)
        CU(cudaMemcpyAsync(c->items.p, hi, n_items * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemsetAsync(c->counters.p, 0, kCounterSlots * sizeof(unsigned int), c->stream));
    const uint32_t* d_items = c->items.as<uint32_t>();
    unsigned int* d_cnt = c->counters.as<unsigned int>();

    // ---- kernels
    CU(cudaEventRecord(c->ev[0], c->stream));
    {
        auto kfn = k_oti;
        const int no_oti = ((flags & kFlagNoOti) || !c->has_gchroma) ? 1 : 0;
        ACOSS_LAUNCH(kfn, (n + 127) / 128, 128, 0, c->stream, c->pairs.as<PairWork>(), n,
                     c->gchroma.as<double>(), c->bins, no_oti);
        c->launches++;
        CU(cudaGetLastError());
    }
    int rc;
    if (L.have_keys) {
        // windowed norms for this m (once per upload / window) and the pairs' fixed-point exponents
        if (c->wnorm_m != m) {
            CU(c->wnorm.reserve((size_t)std::max<int64_t>(c->total_frames, 1) * sizeof(double)));
            CU(c->wnmax.reserve((size_t)c->n_songs * sizeof(double)));
            auto kw = k_wnorm;
            ACOSS_LAUNCH(kw, c->n_songs, 128, 0, c->stream, c->norms.as<double>(), c->offs.as<int64_t>(),
                         c->lens.as<int32_t>(), c->n_songs, m, c->wnorm.as<double>(), c->wnmax.as<double>());
            c->launches++;
            CU(cudaGetLastError());
            c->wnorm_m = m;
        }
        auto ke = k_expo;
        ACOSS_LAUNCH(ke, (n + 127) / 128, 128, 0, c->stream, c->pairs.as<PairWork>(), n, c->wnmax.as<double>());
        c->launches++;
        CU(cudaGetLastError());
    }
    CU(cudaEventRecord(c->ev[1], c->stream));
    // Pairs with a song beyond the integer-key sweeps' reach go through the general kernels: few items,
    // each a long CTA.  Launched alone they leave most SMs idle (profiles/r01_g: 27 % of a step), so
    // they run on a second, higher-priority stream, first, and the persistent CTAs of the main sweeps
    // take the SMs that are left.  The two families never share a pair.
    bool side_work = !L.wide0.empty() || !L.wide1.empty() || !L.t3.empty() || !L.t1.empty(), main_work = L.have_keys;
    for (int b = 0; b < kFastBuckets; ++b) side_work = side_work || !L.fast0[b].empty() || !L.fast2[b].empty();
    const bool side = side_work && main_work && !c->opt_single_stream;
    {
        struct StreamSwap {
            acoss_ctx* c;
            cudaStream_t keep;
            ~StreamSwap() { c->stream = keep; }
        } swap{c, c->stream};
        if (side) {
            CU(cudaEventRecord(c->ev_prep, c->stream));
            CU(cudaStreamWaitEvent(c->stream2, c->ev_prep, 0));
            c->stream = c->stream2;
        }
        // largest buckets first: their CTAs are the long ones
        if ((rc = launch_sweep<BINS, 256, 0>(c, d_items + P.o_wide0, (int)L.wide0.size(), L.wide0_w, m, d_cnt + kFastBuckets, flags))) return rc;
        for (int b = kFastBuckets - 1; b >= 0; --b)
            if ((rc = launch_sweep_fast<BINS, 0>(c, d_items + P.o_fast0[b], (int)L.fast0[b].size(), b, m, d_cnt + b, flags))) return rc;
        for (size_t wv = 0; wv < L.wave_t3.size(); ++wv) {
            const size_t a3 = L.wave_t3[wv], b3 = wv + 1 < L.wave_t3.size() ? L.wave_t3[wv + 1] : L.t3.size();
            const size_t as = L.wave_tsel[wv], bs = wv + 1 < L.wave_tsel.size() ? L.wave_tsel[wv + 1] : L.tsel.size();
            const size_t ae = L.wave_te[wv], be = wv + 1 < L.wave_te.size() ? L.wave_te[wv + 1] : L.te.size();
            if (wv > 0) CU(cudaMemsetAsync(d_cnt + 72, 0, sizeof(unsigned int), c->stream));  // the wave's work-queue head
            if ((rc = launch_sweep_tile<BINS, 3>(c, d_items + P.o_t3 + a3, (int)(b3 - a3), m, d_cnt + 72, flags))) return rc;
            if ((rc = launch_select_rows(c, d_items + P.o_tsel + as, (int)(bs - as), L.tsel_rows, L.tsel_cols, m))) return rc;
            if ((rc = launch_emit_crp(c, d_items + P.o_te + ae, (int)(be - ae), L.te_rows, m, flags))) return rc;
        }
        if ((rc = launch_sweep_tile<BINS, 1>(c, d_items + P.o_t1, (int)L.t1.size(), m, d_cnt + 73, flags))) return rc;
        if ((rc = launch_sweep<BINS, 256, 1>(c, d_items + P.o_wide1, (int)L.wide1.size(), L.wide1_w, m, d_cnt + 2 * kFastBuckets + 1, flags))) return rc;
        for (int b = kFastBuckets - 1; b >= 0; --b)
            if ((rc = launch_sweep_fast<BINS, 2>(c, d_items + P.o_fast2[b], (int)L.fast2[b].size(), b, m, d_cnt + kFastBuckets + 1 + b, flags))) return rc;
        if (side) CU(cudaEventRecord(c->ev_side, c->stream2));
    }
    for (int b = kSweep7Buckets - 1; b >= 0; --b)
        if ((rc = acoss_launch_sweep7(c, 0, BINS, d_items + P.o_g70[b], (int)L.g7_0[b].size(), b, d_cnt + 80 + b, flags))) return rc;
    for (int b = kG4Buckets - 1; b >= 0; --b)
        if (BINS == 12 && (rc = launch_sweep4<12, 0>(c, d_items + P.o_g40[b], (int)L.g4_0[b].size(), b, m, d_cnt + 32 + b, flags))) return rc;
    CU(cudaEventRecord(c->ev[2], c->stream));
    for (int b = kSweep7Buckets - 1; b >= 0; --b)
        if ((rc = acoss_launch_sweep7(c, 2, BINS, d_items + P.o_g72[b], (int)L.g7_2[b].size(), b, d_cnt + 80 + kSweep7Buckets + b, flags))) return rc;
    for (int b = kG4Buckets - 1; b >= 0; --b)
        if (BINS == 12 && (rc = launch_sweep4<12, 2>(c, d_items + P.o_g42[b], (int)L.g4_2[b].size(), b, m, d_cnt + 32 + kG4Buckets + b, flags))) return rc;
    if (!L.g4_pack.empty()) {
        auto kp = k_nb_to_packed;
        ACOSS_LAUNCH(kp, (int)L.g4_pack.size(), 128, 0, c->stream, c->pairs.as<PairWork>(), d_items + P.o_g4pack,
                     (int)L.g4_pack.size(), c->nb.as<uint8_t>(), c->crp.as<uint32_t>(), m);
        c->launches++;
        CU(cudaGetLastError());
    }
    if (side) CU(cudaStreamWaitEvent(c->stream, c->ev_side, 0));
    CU(cudaEventRecord(c->ev[3], c->stream));
    if ((rc = run_dp_lists(c, L, d_items, P.o_dp4, P.o_dp8, P.o_dp16, flags))) return rc;
    CU(cudaEventRecord(c->ev[4], c->stream));

    if (c->bind_q || c->bind_d) {
        auto ks = k_scatter_scores;
        ACOSS_LAUNCH(ks, (n + 127) / 128, 128, 0, c->stream, c->pairs.as<PairWork>(), n, c->qout.as<float>(),
                     c->dout.as<float>(), c->bind_map.p ? c->bind_map.as<int32_t>() : nullptr, c->bind_q, c->bind_d,
                     c->bind_n);
        c->launches++;
        CU(cudaGetLastError());
    }
    // ---- results
    float* res = reinterpret_cast<float*>(P.stage->as<char>() + P.res_off);
    if (qmax_raw) CU(cudaMemcpyAsync(res, c->qout.p, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    if (dmax_raw) CU(cudaMemcpyAsync(res + n, c->dout.p, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    if (sw_raw && (flags & kFlagWantSw))
        CU(cudaMemcpyAsync(res + 2 * (size_t)n, c->swout.p, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    if (want_oti) {
        // the oti lives in the PairWork records
        CU(cudaMemcpyAsync(P.stage->p, c->pairs.p, pw_bytes, cudaMemcpyDeviceToHost, c->stream));
    }
    if (want_crp) {
        c->h_crp.resize((size_t)std::max<int64_t>(crp_total, 1));
        CU(cudaMemcpyAsync(c->h_crp.data(), c->crp.p, (size_t)crp_total * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    }
    return ACOSS_OK;
}

// Wait for a chunk and add its stage times.
int wait_chunk(acoss_ctx* c) {
    auto t0 = std::chrono::steady_clock::now();
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaGetLastError());
    c->host_ms[4] += std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    for (int s = 0; s < 4; ++s) {
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, c->ev[s], c->ev[s + 1]));
        c->stage_ms[s] += ms;
    }
    {
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, c->ev[0], c->ev[4]));
        c->stage_ms[4] += ms;
    }
    return ACOSS_OK;
}

// Hand a finished chunk's results (pinned staging image, host records) to the caller's arrays.
int deliver_chunk(acoss_ctx* c, const ChunkPlan& P, uint32_t flags, float* qmax_raw, float* dmax_raw, float* sw_raw,
                  int32_t* oti_out, int32_t* shape_out, uint32_t* crp_out, const int64_t* crp_offsets,
                  int32_t* crp_swapped_out) {
    const std::vector<PairWork>& W = P.W;
    const int64_t p0 = P.p0;
    const int n = P.n;
    const auto t0 = std::chrono::steady_clock::now();
    struct Lap {
        acoss_ctx* c;
        std::chrono::steady_clock::time_point t0;
        ~Lap() { c->host_ms[5] += std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count(); }
    } lap{c, t0};
    const float* res = reinterpret_cast<const float*>(P.stage->as<char>() + P.res_off);
    if (qmax_raw) memcpy(qmax_raw + p0, res, (size_t)n * sizeof(float));
    if (dmax_raw) memcpy(dmax_raw + p0, res + n, (size_t)n * sizeof(float));
    if (sw_raw && (flags & kFlagWantSw)) memcpy(sw_raw + p0, res + 2 * (size_t)n, (size_t)n * sizeof(float));
    const PairWork* back = P.stage->as<PairWork>();
    for (int p = 0; p < n; ++p) {
        if (oti_out) oti_out[p0 + p] = back[p].oti;
        if (crp_swapped_out) crp_swapped_out[p0 + p] = W[p].swap;
        if (shape_out) {
            shape_out[2 * (p0 + p)] = W[p].M;
            shape_out[2 * (p0 + p) + 1] = W[p].N;
        }
        if (sw_raw && !(flags & kFlagWantSw)) sw_raw[p0 + p] = 0.0f;
        if (crp_out && crp_offsets) {
            const size_t words = (size_t)W[p].crp_rows * W[p].crp_wp;
            if (words) memcpy(crp_out + crp_offsets[p0 + p], c->h_crp.data() + W[p].crp_off, words * sizeof(uint32_t));
        }
    }
    return ACOSS_OK;
}

int check_ctx(acoss_ctx* c) {
    if (!c) return fail(ACOSS_ERR_ARG, "null context");
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) return fail(ACOSS_ERR_CUDA, "cudaSetDevice(%d): %s", c->device, cudaGetErrorString(e));
    return ACOSS_OK;
}

}  // namespace

std::string& acoss_error_slot() {
    thread_local std::string g_last_error;
    return g_last_error;
}

extern "C" {

const char* acoss_last_error(void) { return acoss_error_slot().c_str(); }
int acoss_version(void) { return 100; }

int acoss_create(acoss_ctx** out, int device) {
    if (!out) return fail(ACOSS_ERR_ARG, "out is null");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0)
        return fail(ACOSS_ERR_CUDA, "no CUDA device available (%s); this library has no CPU path",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
    if (device < 0 || device >= ndev) return fail(ACOSS_ERR_ARG, "device %d out of range (0..%d)", device, ndev - 1);
    CU(cudaSetDevice(device));
    acoss_ctx* c = new acoss_ctx();
    c->device = device;
    CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    {
        int prio_lo = 0, prio_hi = 0;
        CU(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
        CU(cudaStreamCreateWithPriority(&c->stream2, cudaStreamNonBlocking, prio_hi));
        CU(cudaEventCreateWithFlags(&c->ev_prep, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&c->ev_side, cudaEventDisableTiming));
    }
    CU(cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device));
    CU(cudaDeviceGetAttribute(&c->max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
    for (int i = 0; i < 5; ++i) CU(cudaEventCreate(&c->ev[i]));
    *out = c;
    return ACOSS_OK;
}

int acoss_destroy(acoss_ctx* c) {
    if (!c) return ACOSS_OK;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    if (c->stream2) cudaStreamSynchronize(c->stream2);
    DevBuf* bufs[] = {&c->frames, &c->norms, &c->offs, &c->gchroma, &c->pairs, &c->items, &c->counters,
                      &c->seps, &c->crp, &c->bnd, &c->qout, &c->dout, &c->swout, &c->otiout, &c->gscratch,
                      &c->dense_in, &c->dense_desc, &c->dense_starts,
                      &c->lens, &c->wnorm, &c->wnmax, &c->sepk, &c->sepx, &c->nb, &c->eval_D, &c->tmat,
                      &c->frames2, &c->snf_desc, &c->snf_mats, &c->snf_md, &c->snf_sidx, &c->snf_sval, &c->bind_map};
    for (DevBuf* b : bufs) b->release();
    c->h_stage.release();
    c->h_stage2.release();
    for (int i = 0; i < 5; ++i)
        if (c->ev[i]) cudaEventDestroy(c->ev[i]);
    if (c->ev_prep) cudaEventDestroy(c->ev_prep);
    if (c->ev_side) cudaEventDestroy(c->ev_side);
    if (c->stream2) cudaStreamDestroy(c->stream2);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return ACOSS_OK;
}

int acoss_upload_songs(acoss_ctx* c, const void* frames, int dtype, const int64_t* offsets, const int32_t* lengths,
                       int32_t n_songs, int32_t bins, const double* gchroma) {
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!frames || !offsets || !lengths) return fail(ACOSS_ERR_ARG, "null input");
    if (n_songs <= 0) return fail(ACOSS_ERR_ARG, "n_songs must be positive");
    if (bins != 12 && bins != 13) return fail(ACOSS_ERR_ARG, "bins must be 12 (chroma) or 13 (mfcc), got %d", bins);
    if (dtype != ACOSS_F32 && dtype != ACOSS_F64) return fail(ACOSS_ERR_ARG, "unknown dtype %d", dtype);
    int64_t total = 0;
    for (int s = 0; s < n_songs; ++s) {
        if (lengths[s] < 0 || offsets[s] < 0) return fail(ACOSS_ERR_ARG, "negative length/offset for song %d", s);
        total = std::max<int64_t>(total, offsets[s] + lengths[s]);
    }
    c->n_songs = n_songs;
    c->bins = bins;
    c->total_frames = total;
    c->h_offs.assign(offsets, offsets + n_songs);
    c->h_len.assign(lengths, lengths + n_songs);
    const size_t n_el = (size_t)total * bins;
    CU(c->frames.reserve(std::max<size_t>(n_el, 1) * sizeof(double)));
    CU(c->norms.reserve(std::max<size_t>((size_t)total, 1) * sizeof(double)));
    CU(c->offs.reserve((size_t)n_songs * sizeof(int64_t)));
    if (dtype == ACOSS_F64) {
        CU(cudaMemcpyAsync(c->frames.p, frames, n_el * sizeof(double), cudaMemcpyDefault, c->stream));  // host or device source
    } else {
        CU(c->dense_in.reserve(std::max<size_t>(n_el, 1) * sizeof(float)));
        CU(cudaMemcpyAsync(c->dense_in.p, frames, n_el * sizeof(float), cudaMemcpyDefault, c->stream));
        if (n_el) {
            auto kfn = k_f32_to_f64;
            ACOSS_LAUNCH(kfn, (n_el + 255) / 256, 256, 0, c->stream, c->dense_in.as<float>(), c->frames.as<double>(),
                         (int64_t)n_el);
            c->launches++;
        }
    }
    CU(cudaMemcpyAsync(c->offs.p, offsets, (size_t)n_songs * sizeof(int64_t), cudaMemcpyHostToDevice, c->stream));
    CU(c->lens.reserve((size_t)n_songs * sizeof(int32_t)));
    CU(cudaMemcpyAsync(c->lens.p, lengths, (size_t)n_songs * sizeof(int32_t), cudaMemcpyHostToDevice, c->stream));
    c->wnorm_m = -1;
    if (total) {
        auto kfn = k_norms<double>;
        ACOSS_LAUNCH(kfn, (total + 255) / 256, 256, 0, c->stream, c->frames.as<double>(), c->norms.as<double>(), total,
                     (int)bins);
        c->launches++;
    }
    c->has_gchroma = gchroma != nullptr;
    CU(c->gchroma.reserve((size_t)n_songs * bins * sizeof(double)));
    if (gchroma)
        CU(cudaMemcpyAsync(c->gchroma.p, gchroma, (size_t)n_songs * bins * sizeof(double), cudaMemcpyDefault, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaGetLastError());
    c->h_gchroma.clear();
    c->bins2 = 0;   // a new song set invalidates the auxiliary arena
    if (gchroma) {
        c->h_gchroma.resize((size_t)n_songs * bins);
        CU(cudaMemcpy(c->h_gchroma.data(), c->gchroma.p, (size_t)n_songs * bins * sizeof(double), cudaMemcpyDeviceToHost));
    }
    return ACOSS_OK;
}

int acoss_serra09_pairs(acoss_ctx* c, const int32_t* pairs, int64_t P, int32_t m, double kappa, uint32_t flags,
                        float* qmax_raw, float* dmax_raw, float* sw_raw, int32_t* oti_out, int32_t* shape_out,
                        uint32_t* crp_out, const int64_t* crp_offsets, int32_t* crp_swapped_out) {
    const auto t_call = std::chrono::steady_clock::now();
    int rc = check_ctx(c);
    if (rc) return rc;
    if (c->n_songs <= 0) return fail(ACOSS_ERR_STATE, "acoss_upload_songs has not been called");
    if (P < 0 || (P > 0 && !pairs)) return fail(ACOSS_ERR_ARG, "bad pair list");
    if (m < 1) return fail(ACOSS_ERR_ARG, "m must be >= 1");
    if (!(kappa > 0.0)) return fail(ACOSS_ERR_ARG, "kappa must be > 0");
    if (crp_out && !crp_offsets) return fail(ACOSS_ERR_ARG, "crp_out needs crp_offsets");
    if (!(flags & kFlagNoOti) && !c->has_gchroma) return fail(ACOSS_ERR_STATE, "OTI requested but no gchroma uploaded");
    for (int64_t p = 0; p < P; ++p) {
        const int32_t i = pairs[2 * p], j = pairs[2 * p + 1];
        if (i < 0 || i >= c->n_songs || j < 0 || j >= c->n_songs)
            return fail(ACOSS_ERR_ARG, "pair %lld = (%d, %d) out of range (n_songs %d)", (long long)p, i, j, c->n_songs);
    }
    for (int s = 0; s < 5; ++s) c->stage_ms[s] = 0.f;
    c->have_stage_ms = true;
    flags |= c->opt_debug_flags;
    // Chunks whose scratch fits the budget.  The host side of chunk k + 1 (work records, item lists, staging image)
    // is prepared while the kernels of chunk k run; two pinned staging buffers alternate.
    auto cut = [&](int64_t p0) {
        int n = 0;
        int64_t sep = 0, crp = 0, nbb = 0;
        while (p0 + n < P && n < c->opt_chunk_pairs) {
            const int Li = c->h_len[pairs[2 * (p0 + n)]], Lj = c->h_len[pairs[2 * (p0 + n) + 1]];
            const int64_t M = std::max(Li - m + 1, 0), N = std::max(Lj - m + 1, 0);
            const int64_t s2 = M + N, c2 = std::max(M, N) * ((std::min(M, N) + 31) / 32 + 1);
            const int64_t n2 = std::max(M, N) * (((std::min(Li, Lj) + 99) / 100) * 32 + 128) + 16;
            if (n > 0 && (sep + s2 > kMaxChunkSepDoubles || crp + c2 > kMaxChunkCrpWords || nbb + n2 > kMaxChunkNbBytes)) break;
            sep += s2;
            crp += c2;
            nbb += n2;
            ++n;
        }
        return n;
    };
    auto prep = [&](int64_t p0, int n, PinBuf* stage, ChunkPlan& plan) {
        return c->bins == 12 ? prep_chunk<12>(c, pairs, p0, n, m, kappa, flags, stage, plan)
                             : prep_chunk<13>(c, pairs, p0, n, m, kappa, flags, stage, plan);
    };
    const bool want_oti = oti_out || crp_swapped_out;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto ms_since = [&](std::chrono::steady_clock::time_point t) {
        return std::chrono::duration<float, std::milli>(now() - t).count();
    };
    for (float& v : c->host_ms) v = 0.f;
    ChunkPlan plans[2];
    PinBuf* stages[2] = {&c->h_stage, &c->h_stage2};
    int cur = 0;
    int64_t p0 = 0;
    auto t_prep = now();
    if (P > 0 && (rc = prep(0, cut(0), stages[0], plans[0]))) return rc;
    c->host_ms[1] = ms_since(t_prep);
    int pending = -1;   // chunk waited for but not yet delivered: its results go out under the next chunk's kernels
    auto deliver = [&](int k) {
        return deliver_chunk(c, plans[k], flags, qmax_raw, dmax_raw, sw_raw, oti_out, shape_out, crp_out, crp_offsets,
                             crp_swapped_out);
    };
    while (p0 < P) {
        ChunkPlan& plan = plans[cur];
        auto t_exec = now();
        rc = c->bins == 12 ? exec_chunk<12>(c, plan, m, flags, qmax_raw, dmax_raw, sw_raw, want_oti, crp_out != nullptr)
                           : exec_chunk<13>(c, plan, m, flags, qmax_raw, dmax_raw, sw_raw, want_oti, crp_out != nullptr);
        c->host_ms[2] += ms_since(t_exec);
        const int64_t next = p0 + plan.n;
        int rc2 = ACOSS_OK;
        t_prep = now();
        if (!rc && pending >= 0) {
            rc2 = deliver(pending);
            pending = -1;
        }
        if (!rc && !rc2 && next < P) rc2 = prep(next, cut(next), stages[cur ^ 1], plans[cur ^ 1]);
        c->host_ms[3] += ms_since(t_prep);
        if (rc || rc2) {
            const std::string msg = acoss_error_slot();
            cudaStreamSynchronize(c->stream);  // nothing of this call stays in flight
            acoss_error_slot() = msg;
            return rc ? rc : rc2;
        }
        if ((rc = wait_chunk(c))) return rc;
        if (crp_out) {   // the plots come back through one host buffer: hand them out before the next chunk reuses it
            if ((rc = deliver(cur))) return rc;
        } else {
            pending = cur;
        }
        p0 = next;
        cur ^= 1;
    }
    if (pending >= 0 && (rc = deliver(pending))) return rc;
    c->host_ms[0] = ms_since(t_call);
    return ACOSS_OK;
}

int acoss_bind_score_matrices(acoss_ctx* c, float* qmax_dev, float* dmax_dev, int64_t n, const int32_t* song_of_slot,
                              int32_t n_slots) {
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!qmax_dev && !dmax_dev) {   // unbind
        c->bind_q = c->bind_d = nullptr;
        c->bind_n = 0;
        c->bind_map.release();
        return ACOSS_OK;
    }
    if (n <= 0) return fail(ACOSS_ERR_ARG, "score matrices need n > 0");
#ifndef ACOSS_EMU
    for (float* p : {qmax_dev, dmax_dev}) {
        if (!p) continue;
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, p) != cudaSuccess || at.type != cudaMemoryTypeDevice || at.device != c->device) {
            cudaGetLastError();
            return fail(ACOSS_ERR_ARG, "score matrices must be device memory of device %d", c->device);
        }
    }
#endif
    if (song_of_slot) {
        if (n_slots <= 0) return fail(ACOSS_ERR_ARG, "song_of_slot needs n_slots > 0");
        for (int32_t s = 0; s < n_slots; ++s)
            if (song_of_slot[s] >= n) return fail(ACOSS_ERR_ARG, "song_of_slot[%d] = %d outside the %lld x %lld matrices", s, song_of_slot[s], (long long)n, (long long)n);
        CU(c->bind_map.reserve((size_t)n_slots * sizeof(int32_t)));
        CU(cudaMemcpyAsync(c->bind_map.p, song_of_slot, (size_t)n_slots * sizeof(int32_t), cudaMemcpyHostToDevice, c->stream));
        CU(cudaStreamSynchronize(c->stream));
    } else {
        c->bind_map.release();
    }
    c->bind_q = qmax_dev;
    c->bind_d = dmax_dev;
    c->bind_n = n;
    return ACOSS_OK;
}

int acoss_dp_packed(acoss_ctx* c, const uint32_t* crp, const int64_t* crp_offsets, const int32_t* rows,
                    const int32_t* cols, int32_t B, uint32_t flags, float* out) {
    int rc = check_ctx(c);
    if (rc) return rc;
    if (B < 0 || (B > 0 && (!crp || !crp_offsets || !rows || !cols || !out))) return fail(ACOSS_ERR_ARG, "null input");
    if (B == 0) return ACOSS_OK;
    std::vector<PairWork> W((size_t)B);
    ItemLists L;
    int64_t crp_total = 0, bnd_total = 0;
    for (int b = 0; b < B; ++b) {
        if (rows[b] < 0 || cols[b] < 0) return fail(ACOSS_ERR_ARG, "negative shape for plot %d", b);
        PairWork& w = W[b];
        memset(&w, 0, sizeof(w));
        w.crp_rows = rows[b];
        w.crp_cols = cols[b];
        w.crp_wp = (cols[b] + 31) / 32;
        w.crp_off = crp_total;
        crp_total += (int64_t)w.crp_rows * w.crp_wp;
        const int64_t br = bnd_records(c, w.crp_rows, w.crp_cols);
        if (br > 0 && std::min(w.crp_rows, w.crp_cols) > 32767)
            return fail(ACOSS_ERR_ARG, "plot %d x %d: the DP hand-over packs Qmax values in 16 bits (min side <= 32767)", w.crp_rows, w.crp_cols);
        w.bnd_off = br > 0 ? bnd_total : -1;
        bnd_total += br;
        classify_dp(c, L, (uint32_t)b, w.crp_rows, w.crp_cols);
    }
    CU(c->pairs.reserve((size_t)B * sizeof(PairWork)));
    CU(c->crp.reserve((size_t)std::max<int64_t>(crp_total, 1) * sizeof(uint32_t)));
    CU(c->bnd.reserve((size_t)std::max<int64_t>(bnd_total, 1) * sizeof(int4)));
    CU(c->qout.reserve((size_t)B * sizeof(float)));
    CU(c->dout.reserve((size_t)B * sizeof(float)));
    CU(c->swout.reserve((size_t)B * sizeof(float)));
    CU(c->items.reserve((size_t)B * sizeof(uint32_t)));
    // gather the plots into the device arena
    c->h_crp.resize((size_t)std::max<int64_t>(crp_total, 1));
    for (int b = 0; b < B; ++b) {
        const size_t words = (size_t)W[b].crp_rows * W[b].crp_wp;
        if (words) memcpy(c->h_crp.data() + W[b].crp_off, crp + crp_offsets[b], words * sizeof(uint32_t));
    }
    std::vector<uint32_t> items;
    items.insert(items.end(), L.dp4.begin(), L.dp4.end());
    items.insert(items.end(), L.dp8.begin(), L.dp8.end());
    items.insert(items.end(), L.dp16.begin(), L.dp16.end());
    CU(cudaMemcpyAsync(c->pairs.p, W.data(), (size_t)B * sizeof(PairWork), cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->crp.p, c->h_crp.data(), (size_t)crp_total * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->items.p, items.data(), items.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemsetAsync(c->swout.p, 0, (size_t)B * sizeof(float), c->stream));
    rc = run_dp_lists(c, L, c->items.as<uint32_t>(), 0, L.dp4.size(), L.dp4.size() + L.dp8.size(), flags);
    if (rc) return rc;
    std::vector<float> q((size_t)B), d((size_t)B), s((size_t)B);
    CU(cudaMemcpyAsync(q.data(), c->qout.p, (size_t)B * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(d.data(), c->dout.p, (size_t)B * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(s.data(), c->swout.p, (size_t)B * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaGetLastError());
    for (int b = 0; b < B; ++b) {
        out[3 * b] = q[b];
        out[3 * b + 1] = d[b];
        out[3 * b + 2] = (flags & kFlagWantSw) ? s[b] : 0.0f;
    }
    return ACOSS_OK;
}

}  // extern "C"

// csm == nullptr: the matrices are already in c->dense_in (acoss_align_dense_resident)
static int align_dense_impl(acoss_ctx* c, const void* csm, int dtype, const int64_t* csm_offsets, const int32_t* rows,
                            const int32_t* cols, int32_t B, double kappa, uint32_t flags, float* out, uint32_t* crp_out,
                            const int64_t* crp_offsets) {
    int rc = check_ctx(c);
    if (rc) return rc;
    if (B < 0 || (B > 0 && (!rows || !cols || !out))) return fail(ACOSS_ERR_ARG, "null input");
    if (dtype != ACOSS_F32 && dtype != ACOSS_F64) return fail(ACOSS_ERR_ARG, "unknown dtype %d", dtype);
    if (!(kappa > 0.0)) return fail(ACOSS_ERR_ARG, "kappa must be > 0");
    if (crp_out && !crp_offsets) return fail(ACOSS_ERR_ARG, "crp_out needs crp_offsets");
    if (B == 0) return ACOSS_OK;
    const size_t esz = dtype == ACOSS_F64 ? sizeof(double) : sizeof(float);
    std::vector<PairWork> W((size_t)B);
    std::vector<DenseDesc> D((size_t)B);
    std::vector<int64_t> starts((size_t)2 * B);  // [line_start | word_start]
    ItemLists L;
    int64_t el_total = 0, sep_total = 0, crp_total = 0, bnd_total = 0, lines = 0;
    for (int b = 0; b < B; ++b) {
        if (rows[b] <= 0 || cols[b] <= 0) return fail(ACOSS_ERR_ARG, "empty matrix %d", b);
        DenseDesc& d = D[b];
        d.rows = rows[b];
        d.cols = cols[b];
        d.k_row = neighbours(kappa, cols[b]);
        d.k_col = neighbours(kappa, rows[b]);
        d.wp = (cols[b] + 31) / 32;
        d.pad = 0;
        d.csm_off = el_total;
        d.sep_off = sep_total;
        d.crp_off = crp_total;
        starts[b] = lines;
        starts[B + b] = crp_total;
        PairWork& w = W[b];
        memset(&w, 0, sizeof(w));
        w.crp_rows = d.rows;
        w.crp_cols = d.cols;
        w.crp_wp = d.wp;
        w.crp_off = crp_total;
        const int64_t br = bnd_records(c, w.crp_rows, w.crp_cols);
        if (br > 0 && std::min(w.crp_rows, w.crp_cols) > 32767)
            return fail(ACOSS_ERR_ARG, "plot %d x %d: the DP hand-over packs Qmax values in 16 bits (min side <= 32767)", w.crp_rows, w.crp_cols);
        w.bnd_off = br > 0 ? bnd_total : -1;
        bnd_total += br;
        classify_dp(c, L, (uint32_t)b, w.crp_rows, w.crp_cols);
        el_total += (int64_t)d.rows * d.cols;
        sep_total += (int64_t)d.rows + d.cols;
        crp_total += (int64_t)d.rows * d.wp;
        lines += (int64_t)d.rows + d.cols;
    }
    if (csm) CU(c->dense_in.reserve((size_t)el_total * esz));
    else if (c->dense_in.cap < (size_t)el_total * esz) return fail(ACOSS_ERR_STATE, "resident dense input is smaller than the shapes say");
    CU(c->dense_desc.reserve((size_t)B * sizeof(DenseDesc)));
    CU(c->dense_starts.reserve((size_t)2 * B * sizeof(int64_t)));
    CU(c->pairs.reserve((size_t)B * sizeof(PairWork)));
    CU(c->seps.reserve((size_t)sep_total * sizeof(double)));
    CU(c->crp.reserve((size_t)crp_total * sizeof(uint32_t)));
    CU(c->bnd.reserve((size_t)std::max<int64_t>(bnd_total, 1) * sizeof(int4)));
    CU(c->qout.reserve((size_t)B * sizeof(float)));
    CU(c->dout.reserve((size_t)B * sizeof(float)));
    CU(c->swout.reserve((size_t)B * sizeof(float)));
    CU(c->items.reserve((size_t)B * sizeof(uint32_t)));
    for (int b = 0; b < B && csm; ++b)
        CU(cudaMemcpyAsync(c->dense_in.as<char>() + D[b].csm_off * esz,
                           reinterpret_cast<const char*>(csm) + csm_offsets[b] * esz,
                           (size_t)D[b].rows * D[b].cols * esz, cudaMemcpyHostToDevice, c->stream));
    std::vector<uint32_t> items;
    items.insert(items.end(), L.dp4.begin(), L.dp4.end());
    items.insert(items.end(), L.dp8.begin(), L.dp8.end());
    items.insert(items.end(), L.dp16.begin(), L.dp16.end());
    CU(cudaMemcpyAsync(c->dense_desc.p, D.data(), (size_t)B * sizeof(DenseDesc), cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->dense_starts.p, starts.data(), (size_t)2 * B * sizeof(int64_t), cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->pairs.p, W.data(), (size_t)B * sizeof(PairWork), cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->items.p, items.data(), items.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemsetAsync(c->swout.p, 0, (size_t)B * sizeof(float), c->stream));
    const int mutual = (flags & kFlagNonMutual) ? 0 : 1;
    const int64_t* d_line_start = c->dense_starts.as<int64_t>();
    const int64_t* d_word_start = d_line_start + B;
    if (dtype == ACOSS_F64) {
        auto k1 = k_dense_seps<double>;
        ACOSS_LAUNCH(k1, (lines + 3) / 4, 128, 0, c->stream, c->dense_in.as<double>(), c->dense_desc.as<DenseDesc>(),
                     d_line_start, (int)B, lines, c->seps.as<double>());
        auto k2 = k_dense_bits<double>;
        ACOSS_LAUNCH(k2, (crp_total + 255) / 256, 256, 0, c->stream, c->dense_in.as<double>(),
                     c->dense_desc.as<DenseDesc>(), d_word_start, (int)B, crp_total, c->seps.as<double>(),
                     c->crp.as<uint32_t>(), mutual);
    } else {
        auto k1 = k_dense_seps<float>;
        ACOSS_LAUNCH(k1, (lines + 3) / 4, 128, 0, c->stream, c->dense_in.as<float>(), c->dense_desc.as<DenseDesc>(),
                     d_line_start, (int)B, lines, c->seps.as<double>());
        auto k2 = k_dense_bits<float>;
        ACOSS_LAUNCH(k2, (crp_total + 255) / 256, 256, 0, c->stream, c->dense_in.as<float>(),
                     c->dense_desc.as<DenseDesc>(), d_word_start, (int)B, crp_total, c->seps.as<double>(),
                     c->crp.as<uint32_t>(), mutual);
    }
    c->launches += 2;
    CU(cudaGetLastError());
    rc = run_dp_lists(c, L, c->items.as<uint32_t>(), 0, L.dp4.size(), L.dp4.size() + L.dp8.size(), flags);
    if (rc) return rc;
    std::vector<float> q((size_t)B), d((size_t)B), s((size_t)B);
    CU(cudaMemcpyAsync(q.data(), c->qout.p, (size_t)B * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(d.data(), c->dout.p, (size_t)B * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(s.data(), c->swout.p, (size_t)B * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    if (crp_out) {
        c->h_crp.resize((size_t)crp_total);
        CU(cudaMemcpyAsync(c->h_crp.data(), c->crp.p, (size_t)crp_total * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    }
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaGetLastError());
    for (int b = 0; b < B; ++b) {
        out[3 * b] = q[b];
        out[3 * b + 1] = d[b];
        out[3 * b + 2] = (flags & kFlagWantSw) ? s[b] : 0.0f;
        if (crp_out)
            memcpy(crp_out + crp_offsets[b], c->h_crp.data() + D[b].crp_off, (size_t)D[b].rows * D[b].wp * sizeof(uint32_t));
    }
    return ACOSS_OK;
}

int acoss_align_dense_resident(acoss_ctx* c, int dtype, const int32_t* rows, const int32_t* cols, int32_t B, double kappa,
                               uint32_t flags, float* out, uint32_t* crp_out, const int64_t* crp_offsets) {
    return align_dense_impl(c, nullptr, dtype, nullptr, rows, cols, B, kappa, flags, out, crp_out, crp_offsets);
}

extern "C" {

int acoss_align_dense(acoss_ctx* c, const void* csm, int dtype, const int64_t* csm_offsets, const int32_t* rows,
                      const int32_t* cols, int32_t B, double kappa, uint32_t flags, float* out, uint32_t* crp_out,
                      const int64_t* crp_offsets) {
    if (B > 0 && (!csm || !csm_offsets)) return fail(ACOSS_ERR_ARG, "null input");
    return align_dense_impl(c, csm, dtype, csm_offsets, rows, cols, B, kappa, flags, out, crp_out, crp_offsets);
}

static int dp_unpacked(acoss_ctx* c, const uint8_t* S, int32_t rows, int32_t cols, uint32_t flags, float out[3]) {
    if (rows < 0 || cols < 0) return fail(ACOSS_ERR_ARG, "negative shape");
    out[0] = out[1] = out[2] = 0.0f;
    if (rows == 0 || cols == 0) return ACOSS_OK;
    if (!S) return fail(ACOSS_ERR_ARG, "S is null");
    const int wp = (cols + 31) / 32;
    std::vector<uint32_t> packed((size_t)rows * wp, 0u);
    for (int i = 0; i < rows; ++i)
        for (int j = 0; j < cols; ++j)
            // SequenceAlignment.c tests S == 1 (qmax/dmax) or S > 0 (Match/Delta); the CRP is {0,1}
            if (S[(size_t)i * cols + j]) packed[(size_t)i * wp + (j >> 5)] |= 1u << (j & 31);
    const int64_t off = 0;
    return acoss_dp_packed(c, packed.data(), &off, &rows, &cols, 1, flags, out);
}

int acoss_qmax(acoss_ctx* c, const uint8_t* S, int32_t M, int32_t N, float* result) {
    if (!result) return fail(ACOSS_ERR_ARG, "result is null");
    float o[3];
    int rc = dp_unpacked(c, S, M, N, kFlagDmaxClean, o);
    *result = o[0];
    return rc;
}

int acoss_dmax(acoss_ctx* c, const uint8_t* S, int32_t M, int32_t N, float* result) {
    if (!result) return fail(ACOSS_ERR_ARG, "result is null");
    float o[3];
    int rc = dp_unpacked(c, S, M, N, kFlagDmaxClean, o);
    *result = o[1];
    return rc;
}

int acoss_qmax_dmax(acoss_ctx* c, const uint8_t* S, int32_t M, int32_t N, float* qmax_result, float* dmax_result) {
    float o[3];
    int rc = dp_unpacked(c, S, M, N, 0u, o);
    if (qmax_result) *qmax_result = o[0];
    if (dmax_result) *dmax_result = o[1];
    return rc;
}

int acoss_swconstrained(acoss_ctx* c, const uint8_t* S, int32_t N, int32_t M, float* result) {
    if (!result) return fail(ACOSS_ERR_ARG, "result is null");
    float o[3];
    int rc = dp_unpacked(c, S, N, M, kFlagDmaxClean | kFlagWantSw, o);
    *result = o[2];
    return rc;
}

// ---------------------------------------------------------------- feature aggregation
int acoss_aggregate_songs(acoss_ctx* c, const void* full, int dtype, int layout, const int64_t* offsets,
                          const int32_t* nframes, int32_t n_songs, int32_t bins, int32_t fac, int32_t aggregate,
                          int32_t zero_nonfinite, void* out, const int64_t* out_offsets, void* gchroma_out) {
    if (!c) return fail(ACOSS_ERR_ARG, "context is null");
    if (!full || !offsets || !nframes || !out || !out_offsets) return fail(ACOSS_ERR_ARG, "null argument");
    if (dtype != ACOSS_F32 && dtype != ACOSS_F64) return fail(ACOSS_ERR_ARG, "dtype must be ACOSS_F32 or ACOSS_F64");
    if (layout != 0 && layout != 1) return fail(ACOSS_ERR_ARG, "layout must be 0 (frames x bins) or 1 (bins x frames)");
    if (aggregate != 0 && aggregate != 1) return fail(ACOSS_ERR_ARG, "aggregate must be 0 (median) or 1 (mean)");
    if (bins < 1 || bins > 64) return fail(ACOSS_ERR_ARG, "bins must be in 1..64");
    if (fac < 1 || fac > kFeatMaxFac) return fail(ACOSS_ERR_ARG, "fac must be in 1..%d", kFeatMaxFac);
    if (gchroma_out && layout != 0) return fail(ACOSS_ERR_ARG, "the global chroma needs the frames x bins layout");
    if (n_songs <= 0) return ACOSS_OK;
    for (int s = 0; s < n_songs; ++s)
        if (nframes[s] < 1) return fail(ACOSS_ERR_ARG, "song %d has no frames", s);
    CU(cudaSetDevice(c->device));
    c->feat_kernel_ms = 0.f;
    c->feat_bytes = 0.0;
    const size_t es = dtype == ACOSS_F32 ? 4 : 8;
    const size_t chunk_budget = (size_t)512 << 20;  // bytes of full-rate input per chunk
    DevBuf d_in, d_out, d_meta, d_g;
    struct Rel {
        DevBuf *a, *b, *c2, *d;
        ~Rel() { a->release(); b->release(); c2->release(); d->release(); }
    } rel{&d_in, &d_out, &d_meta, &d_g};
    int s0 = 0;
    std::vector<int64_t> in_off, out_off;
    while (s0 < n_songs) {
        // ---- cut a chunk of consecutive songs
        int s1 = s0;
        size_t in_el = 0, out_el = 0;
        in_off.clear();
        out_off.clear();
        while (s1 < n_songs && s1 - s0 < 32768) {
            const size_t el = (size_t)nframes[s1] * bins;
            if (s1 > s0 && (in_el + el) * es > chunk_budget) break;
            in_off.push_back((int64_t)in_el);
            out_off.push_back((int64_t)out_el);
            in_el += el;
            out_el += (size_t)((nframes[s1] + fac - 1) / fac) * bins;
            ++s1;
        }
        const int ns = s1 - s0;
        CU(d_in.reserve(in_el * es));
        CU(d_out.reserve(out_el * es));
        CU(d_meta.reserve((size_t)ns * (2 * sizeof(int64_t) + sizeof(int32_t)) + 64));
        if (gchroma_out) CU(d_g.reserve((size_t)ns * bins * es));
        // songs need not be contiguous on the host: one copy per song, unless they are (one copy per chunk)
        int maxL = 0;
        bool packed = true;
        for (int s = s0; s < s1; ++s) {
            packed = packed && (offsets[s] - offsets[s0]) * (int64_t)bins == in_off[s - s0];
            maxL = std::max(maxL, (nframes[s] + fac - 1) / fac);
        }
        if (packed) {
            CU(cudaMemcpyAsync(d_in.p, reinterpret_cast<const char*>(full) + (size_t)offsets[s0] * bins * es, in_el * es,
                               cudaMemcpyHostToDevice, c->stream));
        } else {
            for (int s = s0; s < s1; ++s)
                CU(cudaMemcpyAsync(d_in.as<char>() + (size_t)in_off[s - s0] * es,
                                   reinterpret_cast<const char*>(full) + (size_t)offsets[s] * bins * es,
                                   (size_t)nframes[s] * bins * es, cudaMemcpyHostToDevice, c->stream));
        }
        int64_t* m_in = d_meta.as<int64_t>();
        int64_t* m_out = m_in + ns;
        int32_t* m_n = reinterpret_cast<int32_t*>(m_out + ns);
        CU(cudaMemcpyAsync(m_in, in_off.data(), (size_t)ns * sizeof(int64_t), cudaMemcpyHostToDevice, c->stream));
        CU(cudaMemcpyAsync(m_out, out_off.data(), (size_t)ns * sizeof(int64_t), cudaMemcpyHostToDevice, c->stream));
        CU(cudaMemcpyAsync(m_n, nframes + s0, (size_t)ns * sizeof(int32_t), cudaMemcpyHostToDevice, c->stream));
        FeatArgs A;
        A.in = d_in.p;
        A.in_off = m_in;
        A.nframes = m_n;
        A.out = d_out.p;
        A.out_off = m_out;
        A.gchroma = gchroma_out ? d_g.p : nullptr;
        A.n_songs = ns;
        A.bins = bins;
        A.fac = fac;
        A.layout = layout;
        A.aggregate = aggregate;
        A.zero_nonfinite = zero_nonfinite;
        A.blocks_per_song = (maxL * bins + 127) / 128;
        const int64_t grid64 = (int64_t)A.blocks_per_song * ns;
        if (grid64 > 0x7fffffffLL) return fail(ACOSS_ERR_ARG, "aggregation chunk too large");
        const int grid = (int)grid64;
        CU(cudaEventRecord(c->ev[0], c->stream));
        if (dtype == ACOSS_F32) {
            auto kf = k_feat_aggregate<float>;
            ACOSS_LAUNCH(kf, grid, 128, 0, c->stream, A);
        } else {
            auto kf = k_feat_aggregate<double>;
            ACOSS_LAUNCH(kf, grid, 128, 0, c->stream, A);
        }
        CU(cudaEventRecord(c->ev[1], c->stream));
        c->launches++;
        CU(cudaGetLastError());
        if (gchroma_out) {
            if (dtype == ACOSS_F32) {
                auto kg = k_feat_gchroma<float>;
                ACOSS_LAUNCH(kg, (ns + 3) / 4, 128, 0, c->stream, A);
            } else {
                auto kg = k_feat_gchroma<double>;
                ACOSS_LAUNCH(kg, (ns + 3) / 4, 128, 0, c->stream, A);
            }
            c->launches++;
            CU(cudaGetLastError());
            CU(cudaMemcpyAsync(reinterpret_cast<char*>(gchroma_out) + (size_t)s0 * bins * es, d_g.p, (size_t)ns * bins * es,
                               cudaMemcpyDeviceToHost, c->stream));
        }
        bool packed_out = true;
        for (int s = s0; s < s1; ++s)
            packed_out = packed_out && (out_offsets[s] - out_offsets[s0]) * (int64_t)bins == out_off[s - s0];
        if (packed_out) {
            CU(cudaMemcpyAsync(reinterpret_cast<char*>(out) + (size_t)out_offsets[s0] * bins * es, d_out.p, out_el * es,
                               cudaMemcpyDeviceToHost, c->stream));
        } else {
            for (int s = s0; s < s1; ++s) {
                const size_t L = (size_t)((nframes[s] + fac - 1) / fac);
                CU(cudaMemcpyAsync(reinterpret_cast<char*>(out) + (size_t)out_offsets[s] * bins * es,
                                   d_out.as<char>() + (size_t)out_off[s - s0] * es, L * bins * es, cudaMemcpyDeviceToHost,
                                   c->stream));
            }
        }
        CU(cudaStreamSynchronize(c->stream));
        {
            float ms = 0.f;
            CU(cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]));
            c->feat_kernel_ms += ms;
            c->feat_bytes += (double)(in_el + out_el) * es;
        }
        s0 = s1;
    }
    return ACOSS_OK;
}

/* F2 row: the rank computation of CoverAlgorithm.getEvalStatistics (CoverAlgorithm.py:345-382). */
int acoss_eval_ranks(acoss_ctx* c, const float* D, int64_t n, const int32_t* order, const int32_t* clique_start,
                     const int32_t* clique_size, int32_t rows, const int64_t* out_offsets, int32_t* iranks_out,
                     uint32_t flags) {
    if (!c) return fail(ACOSS_ERR_ARG, "context is null");
    if (n < 0 || rows < 0 || rows > n) return fail(ACOSS_ERR_ARG, "bad sizes (n %lld, rows %d)", (long long)n, rows);
    if (rows == 0) return ACOSS_OK;
    if (!D || !order || !clique_start || !clique_size || !out_offsets || !iranks_out)
        return fail(ACOSS_ERR_ARG, "null argument");
    if (n > 0x7fffffffLL) return fail(ACOSS_ERR_ARG, "n too large");
    // validate the clique description: order is a permutation, every active row lies inside its clique
    std::vector<int32_t> pos((size_t)n, -1);
    for (int64_t p = 0; p < n; ++p) {
        const int32_t s = order[p];
        if (s < 0 || s >= n || pos[(size_t)s] >= 0) return fail(ACOSS_ERR_ARG, "order is not a permutation of 0..n-1");
        pos[(size_t)s] = (int32_t)p;
    }
    int64_t total = 0;
    for (int p = 0; p < rows; ++p) {
        const int32_t st = clique_start[p], K = clique_size[p];
        if (K < 2 || st < 0 || (int64_t)st + K > n || p < st || p >= st + K)
            return fail(ACOSS_ERR_ARG, "row %d: clique [%d, %d) does not hold it or has fewer than 2 songs", p, st, st + K);
        if (out_offsets[p] != total) return fail(ACOSS_ERR_ARG, "out_offsets[%d] must be the running sum of (K - 1)", p);
        total += K - 1;
    }
    CU(cudaSetDevice(c->device));
    const float* d_D = D;
    if (!(flags & ACOSS_FLAG_DEVICE_INPUT)) {
        CU(c->eval_D.reserve((size_t)n * (size_t)n * sizeof(float)));
        CU(cudaMemcpyAsync(c->eval_D.p, D, (size_t)n * (size_t)n * sizeof(float), cudaMemcpyHostToDevice, c->stream));
        d_D = c->eval_D.as<float>();
    }
    DevBuf meta, ranks;
    struct Rel {
        DevBuf *a, *b;
        ~Rel() { a->release(); b->release(); }
    } rel{&meta, &ranks};
    const size_t meta_bytes = (size_t)n * 8 + (size_t)rows * 16 + 64;
    CU(meta.reserve(meta_bytes));
    CU(ranks.reserve((size_t)total * 8 + 64));
    int32_t* d_order = meta.as<int32_t>();
    int32_t* d_pos = d_order + n;
    int32_t* d_cs = d_pos + n;
    int32_t* d_ck = d_cs + rows;
    int64_t* d_off = reinterpret_cast<int64_t*>(meta.as<char>() + (((size_t)n * 8 + (size_t)rows * 8 + 15) & ~(size_t)15));
    CU(cudaMemcpyAsync(d_order, order, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d_pos, pos.data(), (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d_cs, clique_start, (size_t)rows * 4, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d_ck, clique_size, (size_t)rows * 4, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d_off, out_offsets, (size_t)rows * 8, cudaMemcpyHostToDevice, c->stream));
    EvalArgs A;
    A.D = d_D;
    A.order = d_order;
    A.pos = d_pos;
    A.cstart = d_cs;
    A.csize = d_ck;
    A.out_off = d_off;
    A.raw = ranks.as<int32_t>();
    A.out = A.raw + total;
    A.n = n;
    A.rows = rows;
    CU(cudaEventRecord(c->ev[0], c->stream));
    int kmax = 0;
    for (int p = 0; p < rows; ++p) kmax = std::max(kmax, (int)clique_size[p] - 1);
    if (kmax <= 1) { ACOSS_LAUNCH(k_eval_ranks<1>, rows, kEvalThreads, 0, c->stream, A); }
    else if (kmax <= 2) { ACOSS_LAUNCH(k_eval_ranks<2>, rows, kEvalThreads, 0, c->stream, A); }
    else if (kmax <= 4) { ACOSS_LAUNCH(k_eval_ranks<4>, rows, kEvalThreads, 0, c->stream, A); }
    else if (c->opt_eval_direct) {
        if (kmax <= 8) { ACOSS_LAUNCH(k_eval_ranks<8>, rows, kEvalThreads, 0, c->stream, A); }
        else if (kmax <= 12) { ACOSS_LAUNCH(k_eval_ranks<12>, rows, kEvalThreads, 0, c->stream, A); }
        else { ACOSS_LAUNCH(k_eval_ranks<kEvalMembers>, rows, kEvalThreads, 0, c->stream, A); }
    } else { ACOSS_LAUNCH(k_eval_ranks_bs, rows, kEvalThreads, 0, c->stream, A); }
    c->launches++;
    CU(cudaGetLastError());
    CU(cudaEventRecord(c->ev[1], c->stream));
    CU(cudaMemcpyAsync(iranks_out, A.out, (size_t)total * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaEventElapsedTime(&c->eval_ms, c->ev[0], c->ev[1]));
    return ACOSS_OK;
}

int acoss_last_eval_ms(acoss_ctx* c, float* out) {
    if (!c || !out) return fail(ACOSS_ERR_ARG, "null argument");
    if (c->eval_ms < 0.f) return fail(ACOSS_ERR_STATE, "no acoss_eval_ranks call yet");
    *out = c->eval_ms;
    return ACOSS_OK;
}

int acoss_set_option(acoss_ctx* c, const char* name, int64_t value) {
    if (!c || !name) return fail(ACOSS_ERR_ARG, "null argument");
    const std::string n(name);
    if (n == "force_wide_sweep")
        c->opt_force_wide = value != 0;
    else if (n == "force_gscratch")
        c->opt_force_gscratch = value != 0;
    else if (n == "eval_direct")
        c->opt_eval_direct = value != 0;
    else if (n == "tiled_wide")
        c->opt_tiled_wide = value != 0;
    else if (n == "sweep_gen") {
        if (value != 2 && value != 4 && value != 7) return fail(ACOSS_ERR_ARG, "sweep_gen must be 2, 4 or 7");
        c->opt_sweep_gen = (int)value;
    } else if (n == "gap_units") {
        if (value < 2 || value > (1 << 24)) return fail(ACOSS_ERR_ARG, "gap_units must be in [2, 2^24]");
        c->opt_gap = (int)value;
    } else if (n == "single_stream") {
        c->opt_single_stream = value ? 1 : 0;
    } else if (n == "debug_flags")
        c->opt_debug_flags = (uint32_t)value;
    else if (n == "dp_packed")
        c->opt_dp_packed = value ? 1 : 0;
    else if (n == "dp_max_cpl") {
        if (value != 4 && value != 8 && value != 16) return fail(ACOSS_ERR_ARG, "dp_max_cpl must be 4, 8 or 16");
        c->opt_dp_max_cpl = (int)value;
    } else if (n == "chunk_pairs") {
        if (value < 1) return fail(ACOSS_ERR_ARG, "chunk_pairs must be >= 1");
        c->opt_chunk_pairs = (int)std::min<int64_t>(value, 1 << 20);
    } else
        return fail(ACOSS_ERR_ARG, "unknown option %s", name);
    return ACOSS_OK;
}

int64_t acoss_kernel_launches(acoss_ctx* c) { return c ? c->launches : 0; }
void acoss_reset_counters(acoss_ctx* c) {
    if (c) c->launches = 0;
}

int acoss_debug_sweep7_stats(acoss_ctx* c, uint64_t out[16], int reset) {
    if (!c || !out) return fail(ACOSS_ERR_ARG, "null argument");
    unsigned long long h[16];
    const int rc = acoss_sweep7_stats(c, h, reset);
    for (int i = 0; i < 16; ++i) out[i] = h[i];
    return rc;
}

int acoss_last_aggregate_stats(acoss_ctx* c, double out[2]) {
    if (!c || !out) return fail(ACOSS_ERR_ARG, "null argument");
    out[0] = c->feat_kernel_ms;
    out[1] = c->feat_bytes;
    return ACOSS_OK;
}

int acoss_debug_host_ms(acoss_ctx* c, float out[6]) {
    if (!c || !out) return fail(ACOSS_ERR_ARG, "null argument");
    for (int s = 0; s < 6; ++s) out[s] = c->host_ms[s];
    return ACOSS_OK;
}

int acoss_last_stage_ms(acoss_ctx* c, float out[5]) {
    if (!c || !out) return fail(ACOSS_ERR_ARG, "null argument");
    if (!c->have_stage_ms) return fail(ACOSS_ERR_STATE, "no acoss_serra09_pairs call yet");
    for (int s = 0; s < 5; ++s) out[s] = c->stage_ms[s];
    return ACOSS_OK;
}

}  // extern "C"
