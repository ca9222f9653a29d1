// ctx.h -- the library context and the small host helpers shared by the translation units of
// libacoss_b200.so (api.cu: the C ABI; sweep7.cu: the default sweep kernels and their launcher).
#pragma once
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/acoss_b200.h"
#include "common.cuh"

#ifdef ACOSS_EMU
#define ACOSS_LAUNCH(kfn, grid, block, smem, stream, ...) \
    emu::launch(dim3((unsigned)(grid)), dim3((unsigned)(block)), (size_t)(smem), [&]() { kfn(__VA_ARGS__); })
template <class F>
static cudaError_t cudaOccupancyMaxActiveBlocksPerMultiprocessor(int* n, F, int, size_t) {
    *n = 1;
    return cudaSuccess;
}
#else
#define ACOSS_LAUNCH(kfn, grid, block, smem, stream, ...) kfn<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#endif

std::string& acoss_error_slot();  // thread-local last error (api.cu)

inline int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    acoss_error_slot() = buf;
    return code;
}

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(ACOSS_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <class T>
    T* as() const { return reinterpret_cast<T*>(p); }
};

struct PinBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMallocHost(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() {
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
    }
    template <class T>
    T* as() const { return reinterpret_cast<T*>(p); }
};


struct acoss_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;  // side stream for the few long-song sweeps (general kernels), higher priority
    cudaEvent_t ev_prep = nullptr, ev_side = nullptr;
    int sm_count = 0;
    int max_smem = 0;
    // songs
    DevBuf frames, norms, offs, gchroma;
    std::vector<int64_t> h_offs;
    std::vector<int32_t> h_len;
    int n_songs = 0, bins = 0;
    int64_t total_frames = 0;
    bool has_gchroma = false;
    // second feature arena of the same songs (e.g. mfcc beside chroma) and a host copy of gchroma: the SNF front end
    DevBuf frames2;
    int bins2 = 0;
    std::vector<double> h_gchroma;
    DevBuf snf_desc, snf_mats, snf_md, snf_sidx, snf_sval;
    // per-chunk scratch
    DevBuf pairs, items, counters, seps, crp, bnd, qout, dout, swout, otiout, gscratch;
    DevBuf dense_in, dense_desc, dense_starts;
    DevBuf tmat;                                // tiled long-song sweep: float64 T matrices of the chunk
    DevBuf lens, wnorm, wnmax, sepk, sepx, nb;  // gen-4 sweep: windowed norms, integer separators, nibble CRPs
    int wnorm_m = -1;                           // window the wnorm buffer was computed for
    float* bind_q = nullptr;   // device-resident score matrices (acoss_bind_score_matrices), not owned
    float* bind_d = nullptr;
    int64_t bind_n = 0;
    DevBuf bind_map;           // song id of every uploaded slot (int32), empty = identity
    PinBuf h_stage, h_stage2;  // pinned staging images of two consecutive chunks
    std::vector<uint32_t> h_crp;
    // test/tuning knobs (acoss_set_option)
    int opt_force_wide = 0;      // send every sweep item through the 256-thread kernels
    int opt_force_gscratch = 0;  // ring + strip in global scratch even when shared memory would do
    int opt_dp_max_cpl = 4;      // widest DP kernel used (4, 8, 16).  Measured (profiles/exp/exp_dp_cpl.py):
                                 // 128-column blocks at 80 registers (6 CTAs/SM) beat wider lanes
    int opt_dp_packed = 1;       // Qmax + Dmax on packed 16-bit lanes (k_dp16) where the plots are small enough; 0: k_dp
    int opt_chunk_pairs = 1 << 18;
    int opt_sweep_gen = 7;         // 7 (default): FP64 tensor-core sweep with chained row selection and packed CRP output
                                   // (kernels_sweep7.cuh) for every orientation whose column song has <= 1015 frames
                                   // (m = 9, 12 / 13 bins), the rest as 4; 4: integer-key SIMT sweep
                                   // (kernels_sweep4.cuh, 12 bins, column songs of up to 1320 frames), the rest as 2;
                                   // 2: float64-selection sweep (kernels_sweep2.cuh).  Generations 3, 5 and 6 are
                                   // retired to profiles/exp/.
    int opt_gap = 2;               // gen 4: safety gap G in key units (tests raise it to exercise the exact path)
    int opt_single_stream = 0;     // 1: no side stream (profiling / debugging)
    int opt_eval_direct = 0;       // 1: cliques of more than 5 songs through the one-compare-per-member rank kernel
    int opt_tiled_wide = 1;        // 1: column songs beyond the fast kernels go through the tiled sweep
                                   // (kernels_sweep1t.cuh); 0: through the one-CTA general sweep (kernels_sweep.cuh)
    uint32_t opt_debug_flags = 0;  // OR-ed into the kernel flags (profiling experiments only)
    // accounting
    int64_t launches = 0;
    float stage_ms[5] = {0, 0, 0, 0, 0};
    float feat_kernel_ms = 0.f;   // k_feat_aggregate launches of the last acoss_aggregate_songs call (CUDA events)
    double feat_bytes = 0.0;      // bytes they read + wrote
    float host_ms[6] = {0, 0, 0, 0, 0, 0};  // host wall times of the last acoss_serra09_pairs call (acoss_debug_host_ms)
    bool have_stage_ms = false;
    float eval_ms = -1.f;  // device time of the last k_eval_ranks launch
    DevBuf eval_D;         // score matrix of acoss_eval_ranks when it arrives from the host
    cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
};

// api.cu: the EarlySNF tail (binarise + DPs) over B float64 / float32 matrices ALREADY in ctx->dense_in, matrix b at the
// running sum of rows x cols elements; same outputs as acoss_align_dense.
int acoss_align_dense_resident(acoss_ctx* c, int dtype, const int32_t* rows, const int32_t* cols, int32_t B, double kappa,
                               uint32_t flags, float* out, uint32_t* crp_out, const int64_t* crp_offsets);
