// common.cuh -- shared declarations of the Serra09 CUDA path (sm_100a).
//
// Data layout in HBM:
//   frames   : [total_frames][BINS] float32 or float64, frame-major (= chroma.T of the
//              reference's (bins, L) per-song array, Serra09.py:154); song s starts at
//              frame offs[s].
//   norms    : [total_frames] float64, squared L2 norm of each frame (get_csm's
//              np.sum(X**2, 1), CRPUtils.py:82).
//   gchroma  : [n_songs][BINS] float64 (Serra09.py:24-28).
//   PairWork : one record per pair of the chunk being scored.
//   seps     : float64 separators, per pair [M row separators][N column separators]:
//              a cell is a row-neighbour iff its windowed distance <= the row separator
//              (separator = midpoint between the k-th and (k+1)-th smallest of the row).
//   crp      : packed binary cross-recurrence plots, per pair rows x wp uint32.
#pragma once
#include <stdint.h>

#ifdef ACOSS_EMU
#include "cuda_emu.h"
#define ACOSS_DYNSMEM(name) unsigned char* name = emu::st().dyn_smem
#else
#include <cuda_runtime.h>
#define ACOSS_DYNSMEM(name) extern __shared__ __align__(16) unsigned char name[]
#endif

namespace acoss {

constexpr int kMaxBins = 16;
constexpr unsigned kFullMask = 0xffffffffu;

// flags mirrored from include/acoss_b200.h
constexpr uint32_t kFlagNoOti = 1u << 0;
constexpr uint32_t kFlagDmaxClean = 1u << 1;
constexpr uint32_t kFlagNoSwap = 1u << 2;
constexpr uint32_t kFlagNonMutual = 1u << 3;
constexpr uint32_t kFlagWantSw = 1u << 4;

struct PairWork {
    int32_t song_i, song_j;  // the pair as the caller gave it: OTI rolls song_i (Serra09.py:166-167)
    int32_t Li, Lj;          // frames
    int32_t M, N;            // windows: Li-m+1, Lj-m+1 (rows, cols of the reference's CRP)
    int32_t k_row, k_col;    // neighbours kept per row (out of N) / per column (out of M)
    int32_t oti;             // filled by k_oti
    int32_t swap;            // 1: CRP + DP run on the transpose (rows = song_j windows)
    int32_t crp_rows, crp_cols, crp_wp;  // stored CRP shape, words per row
    int32_t expo;            // fixed-point exponent e of the pair's integer keys (k_expo, kernels_sweep4.cuh)
    int64_t sep_off;         // into seps (doubles) / sepk (int32) / sepx (doubles)
    int64_t crp_off;         // into crp (uint32 words)
    int64_t bnd_off;         // into the DP boundary scratch (int4 records), -1 when unused
    int64_t nb_off;          // into the nibble-byte CRP arena of the gen-4 sweep (bytes)
    int32_t nb_stride;       // bytes per row there (= threads of the MODE 2 CTA)
    int32_t pad1;
    int64_t tm_off[2];       // per orientation: the float64 T matrix of the tiled long-song sweep (doubles), -1 unused
};

// Named barriers (PTX bar.sync / bar.arrive with an explicit thread count): the producer /
// consumer hand-over between the compute warps and the selection warps of the sweep kernel.
__device__ __forceinline__ void named_bar_sync(int id, int count) {
#ifdef ACOSS_EMU
    emu::bar_named(id, count, true);
#else
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
#endif
}
__device__ __forceinline__ void named_bar_arrive(int id, int count) {
#ifdef ACOSS_EMU
    emu::bar_named(id, count, false);
#else
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory");
#endif
}

__device__ __forceinline__ int imin(int a, int b) { return a < b ? a : b; }
__device__ __forceinline__ int imax(int a, int b) { return a > b ? a : b; }
__device__ __forceinline__ int imax3(int a, int b, int c) { return imax(imax(a, b), c); }

__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(kFullMask, v, o));
    return v;
}
__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(kFullMask, v, o));
    return v;
}
__device__ __forceinline__ int warp_sum(int v) {
    return __reduce_add_sync(kFullMask, v);  // REDUX.SUM: one instruction instead of five shuffles
}

}  // namespace acoss
