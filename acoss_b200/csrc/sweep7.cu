// sweep7.cu -- translation unit of the default sweep (kernels_sweep7.cuh): its instantiations and the launcher
// api.cu calls.  A separate file so that the library builds in parallel and a change to this kernel does not
// recompile the rest.
#include <algorithm>
#include <cstdlib>

#include "ctx.h"
#include "kernels_sweep7.cuh"
#include "sweep7.h"

using namespace acoss;

namespace {

template <int MODE, int BINS, int NT, int CPT>
int launch_one(acoss_ctx* c, const uint32_t* d_items, int n_items, unsigned int* d_counter, uint32_t flags, cudaStream_t stream) {
    auto kfn = k_sweep7<MODE, BINS, 9, NT, CPT>;
    const size_t smem = (size_t)s7_layout<NT, CPT>(BINS, MODE).total;
    if (smem > (size_t)c->max_smem) return fail(ACOSS_ERR_CUDA, "k_sweep7 needs %zu B of shared memory, the device offers %d", smem, c->max_smem);
    CU(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kfn, NT, smem));
    if (occ < 1) return fail(ACOSS_ERR_CUDA, "k_sweep7 cannot be resident (threads %d, smem %zu B)", NT, smem);
    const int grid = (int)std::min<int64_t>((int64_t)n_items, (int64_t)c->sm_count * occ);
    Sweep7Args A;
    memset(&A, 0, sizeof(A));
    A.frames = c->frames.as<double>();
    A.wnorm = c->wnorm.as<double>();
    A.norms = c->norms.as<double>();
    A.offs = c->offs.as<int64_t>();
    A.pairs = c->pairs.as<PairWork>();
    A.items = d_items;
    A.n_items = n_items;
    A.counter = d_counter;
    A.sepk = c->sepk.as<int32_t>();
    A.sepx = c->sepx.as<double>();
    A.crp = c->crp.as<uint32_t>();
    A.gap = c->opt_gap;
    A.flags = flags;
    ACOSS_LAUNCH(kfn, grid, NT, smem, stream, A);
    c->launches++;
    CU(cudaGetLastError());
    return ACOSS_OK;
}

template <int MODE, int BINS>
int launch_bucket(acoss_ctx* c, const uint32_t* d_items, int n_items, int bucket, unsigned int* d_counter, uint32_t flags, cudaStream_t stream) {
    switch (bucket) {  // capacity 128 * (bucket + 1) columns
        case 0: return launch_one<MODE, BINS, 128, 1>(c, d_items, n_items, d_counter, flags, stream);
        case 1: return launch_one<MODE, BINS, 256, 1>(c, d_items, n_items, d_counter, flags, stream);
        case 2: return launch_one<MODE, BINS, 192, 2>(c, d_items, n_items, d_counter, flags, stream);
        case 3:
            if (c->opt_debug_flags & (1u << 20)) return launch_one<MODE, BINS, 512, 1>(c, d_items, n_items, d_counter, flags, stream);
            return launch_one<MODE, BINS, 256, 2>(c, d_items, n_items, d_counter, flags, stream);
        case 4: return launch_one<MODE, BINS, 320, 2>(c, d_items, n_items, d_counter, flags, stream);
        case 5: return launch_one<MODE, BINS, 384, 2>(c, d_items, n_items, d_counter, flags, stream);
        case 6: return launch_one<MODE, BINS, 448, 2>(c, d_items, n_items, d_counter, flags, stream);
        case 7: return launch_one<MODE, BINS, 512, 2>(c, d_items, n_items, d_counter, flags, stream);
    }
    return fail(ACOSS_ERR_ARG, "k_sweep7: no kernel for bucket %d", bucket);
}

static_assert(kSweep7Buckets == kS7Buckets && kSweep7Pad == kS7Pad && kSweep7MaxCols == kS7MaxCols, "sweep7.h mirrors kernels_sweep7.cuh");

}  // namespace

int acoss_launch_sweep7(acoss_ctx* c, int mode, int bins, const uint32_t* d_items, int n_items, int bucket, unsigned int* d_counter,
                        uint32_t flags) {
    if (n_items <= 0) return ACOSS_OK;
    if (mode == 0 && bins == 12) return launch_bucket<0, 12>(c, d_items, n_items, bucket, d_counter, flags, c->stream);
    if (mode == 2 && bins == 12) return launch_bucket<2, 12>(c, d_items, n_items, bucket, d_counter, flags, c->stream);
    if (mode == 0 && bins == 13) return launch_bucket<0, 13>(c, d_items, n_items, bucket, d_counter, flags, c->stream);
    if (mode == 2 && bins == 13) return launch_bucket<2, 13>(c, d_items, n_items, bucket, d_counter, flags, c->stream);
    return fail(ACOSS_ERR_ARG, "k_sweep7: mode %d / %d bins not built", mode, bins);
}

size_t acoss_sweep7_smem(int bucket, int bins, int mode) {
    switch (bucket) {
        case 0: return (size_t)s7_layout<128, 1>(bins, mode).total;
        case 1: return (size_t)s7_layout<256, 1>(bins, mode).total;
        case 2: return (size_t)s7_layout<192, 2>(bins, mode).total;
        case 3: return std::max((size_t)s7_layout<512, 1>(bins, mode).total, (size_t)s7_layout<256, 2>(bins, mode).total);
        case 4: return (size_t)s7_layout<320, 2>(bins, mode).total;
        case 5: return (size_t)s7_layout<384, 2>(bins, mode).total;
        case 6: return (size_t)s7_layout<448, 2>(bins, mode).total;
        default: return (size_t)s7_layout<512, 2>(bins, mode).total;
    }
}

int acoss_sweep7_stats(acoss_ctx* c, unsigned long long out[16], int reset) {
#if ACOSS_S7_STAT
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    unsigned long long h[16];
    CU(cudaMemcpyFromSymbol(h, g_s7_stats, sizeof(h)));
    for (int i = 0; i < 16; ++i) out[i] = h[i];
    unsigned long long h4[8];
    CU(cudaMemcpyFromSymbol(h4, g_s4_sel_stats, sizeof(h4)));
    if (getenv("ACOSS_S4_SEL_STATS"))  // selection statistics replace the phase cycles [4..11]
        for (int i = 0; i < 8; ++i) out[4 + i] = h4[i];
    if (reset) {
        memset(h, 0, sizeof(h));
        CU(cudaMemcpyToSymbol(g_s7_stats, h, sizeof(h)));
        CU(cudaMemcpyToSymbol(g_s4_sel_stats, h, sizeof(h4)));
    }
#else
    (void)c;
    (void)reset;
    for (int i = 0; i < 16; ++i) out[i] = 0;
#endif
    return ACOSS_OK;
}
