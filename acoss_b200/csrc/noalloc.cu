// noalloc.cu -- acoss_dp_packed_static: the alignment DPs (k_dp, kernels_dp.cuh) for callers that cannot allocate.
//
// No counterpart in the reference.  Code under CUDA graph capture,
// tight serving loops and harnesses that forbid cudaMalloc cannot use acoss_dp_packed, which sizes its PairWork
// records, item list and column-block hand-over scratch with cudaMalloc per call.  Here that scratch lives in ONE
// file-scope __device__ array of fixed size (kStaticBytes, about 4 MiB, held by every GPU that loads the library
// until the process exits); inputs and outputs are the caller's DEVICE buffers; the entry point calls no
// cudaMalloc / cudaFree / allocating library routine and makes no host <-> device copy.
//
// In host code the array's NAME is a host-side stand-in, not a device address: its device address is obtained once
// per device with cudaGetSymbolAddress (with that device current) and that pointer is what the kernels receive.
//
// The array is shared by every call and keeps the last call's contents.  k_dp_static_setup rewrites the PairWork
// records and the item list of the call and zeroes the output; the hand-over records of a plot are all written by
// column block b before column block b + 1 of the same warp reads them, so nothing else has to start at zero.
// Calls that overlap on the GPU (two streams, two host threads, two processes' worth of graphs replayed
// concurrently on one device) RACE on the array: serialise them.
#include <mutex>

#include "ctx.h"
#include "kernels_dp.cuh"

using namespace acoss;

namespace {

constexpr int kStaticMaxPlots = 64;          // plots per call
constexpr int kStaticMaxSide = 16384;        // rows / columns of one plot
constexpr int kStaticMaxWideRows = 65536;    // sum of rows over the plots wider than one column block (128 columns)
constexpr int kStaticCpl = 4;                // k_dp<4, .>: 128-column blocks
constexpr size_t kPairBytes = (size_t)kStaticMaxPlots * sizeof(PairWork);
constexpr size_t kItemOff = (kPairBytes + 255) & ~(size_t)255;
constexpr size_t kBndOff = (kItemOff + kStaticMaxPlots * sizeof(uint32_t) + 255) & ~(size_t)255;
constexpr size_t kBndRecords = (size_t)kStaticMaxWideRows * 2 * 2;  // 2 buffers x 2 int4 per row
constexpr size_t kStaticBytes = kBndOff + kBndRecords * sizeof(int4);

// the scratch itself: device memory of every GPU that loads the library
static __device__ __align__(256) unsigned char g_dp_static_scratch[kStaticBytes];

struct StaticPlots {  // passed BY VALUE as a kernel parameter (1.3 KB): no host -> device copy is needed
    int32_t B;
    int32_t rows[kStaticMaxPlots];
    int32_t cols[kStaticMaxPlots];
    int64_t crp_off[kStaticMaxPlots];
    int64_t bnd_off[kStaticMaxPlots];
};

// One thread per plot: the PairWork record k_dp reads, the item list, and a zeroed output.
__global__ void __launch_bounds__(kStaticMaxPlots) k_dp_static_setup(StaticPlots P, unsigned char* scratch, float* out) {
    const int b = (int)threadIdx.x;
    if (b >= P.B) return;
    PairWork* pw = reinterpret_cast<PairWork*>(scratch) + b;
    PairWork w;
    memset(&w, 0, sizeof(w));
    w.crp_rows = P.rows[b];
    w.crp_cols = P.cols[b];
    w.crp_wp = (P.cols[b] + 31) / 32;
    w.crp_off = P.crp_off[b];
    w.bnd_off = P.bnd_off[b];
    w.tm_off[0] = w.tm_off[1] = -1;
    *pw = w;
    reinterpret_cast<uint32_t*>(scratch + kItemOff)[b] = (uint32_t)b;
    out[b] = 0.0f;
    out[P.B + b] = 0.0f;
    out[2 * P.B + b] = 0.0f;
}

std::mutex g_addr_mutex;
unsigned char* g_scratch_addr[64] = {nullptr};  // device address of g_dp_static_scratch, per device ordinal

// Set-up, once per device, with that device current.
int scratch_address(int device, unsigned char** out) {
    std::lock_guard<std::mutex> lock(g_addr_mutex);
    if (g_scratch_addr[device] == nullptr) {
        void* p = nullptr;
        CU(cudaGetSymbolAddress(&p, g_dp_static_scratch));
        if (p == nullptr) return fail(ACOSS_ERR_CUDA, "cudaGetSymbolAddress returned a null address on device %d", device);
        g_scratch_addr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_scratch_addr[device];
    return ACOSS_OK;
}

}  // namespace

extern "C" int acoss_dp_packed_static(int32_t device, const uint32_t* crp_dev, const int64_t* crp_offsets, const int32_t* rows,
                                      const int32_t* cols, int32_t B, uint32_t flags, float* out_dev, void* stream) {
    if (B < 0 || B > kStaticMaxPlots)
        return fail(ACOSS_ERR_ARG, "acoss_dp_packed_static: B = %d plots, the static scratch holds at most %d per call", B,
                    kStaticMaxPlots);
    if (B == 0) return ACOSS_OK;
    if (!crp_dev || !crp_offsets || !rows || !cols || !out_dev) return fail(ACOSS_ERR_ARG, "acoss_dp_packed_static: null argument");
    if (device < 0 || device >= 64) return fail(ACOSS_ERR_ARG, "acoss_dp_packed_static: device %d out of range (0..63)", device);
    StaticPlots P;
    memset(&P, 0, sizeof(P));
    P.B = B;
    int64_t wide_rows = 0;
    for (int b = 0; b < B; ++b) {
        if (rows[b] < 0 || cols[b] < 0 || crp_offsets[b] < 0) return fail(ACOSS_ERR_ARG, "acoss_dp_packed_static: negative shape / offset for plot %d", b);
        if (rows[b] > kStaticMaxSide || cols[b] > kStaticMaxSide)
            return fail(ACOSS_ERR_ARG, "acoss_dp_packed_static: plot %d is %d x %d, the limit is %d rows / columns", b, rows[b],
                        cols[b], kStaticMaxSide);
        P.rows[b] = rows[b];
        P.cols[b] = cols[b];
        P.crp_off[b] = crp_offsets[b];
        P.bnd_off[b] = -1;
        if (cols[b] > 32 * kStaticCpl) {  // wider than one column block: hand-over records, 2 buffers x 2 per row
            P.bnd_off[b] = wide_rows * 2 * 2;
            wide_rows += rows[b];
        }
    }
    if (wide_rows > kStaticMaxWideRows)
        return fail(ACOSS_ERR_ARG,
                    "acoss_dp_packed_static: %lld rows in plots wider than %d columns, the static scratch (%zu bytes) holds %d",
                    (long long)wide_rows, 32 * kStaticCpl, kStaticBytes, kStaticMaxWideRows);
    int current = -1;
    CU(cudaGetDevice(&current));
    if (current != device) CU(cudaSetDevice(device));
    unsigned char* scratch = nullptr;
    const int rc = scratch_address(device, &scratch);
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    {
        auto ks = k_dp_static_setup;
        ACOSS_LAUNCH(ks, 1, kStaticMaxPlots, 0, st, P, scratch, out_dev);
        CU(cudaGetLastError());
    }
    DpArgs A;
    A.pairs = reinterpret_cast<const PairWork*>(scratch);
    A.items = reinterpret_cast<const uint32_t*>(scratch + kItemOff);
    A.n_items = B;
    A.crp = crp_dev;
    A.bnd = reinterpret_cast<int4*>(scratch + kBndOff);
    A.qmax_out = out_dev;
    A.dmax_out = out_dev + B;
    A.sw_out = out_dev + 2 * B;
    A.flags = flags;
    const int grid = (B + 3) / 4;
    if (flags & kFlagWantSw) {
        auto kfn = k_dp<kStaticCpl, true>;
        ACOSS_LAUNCH(kfn, grid, 128, 0, st, A);
    } else {
        auto kfn = k_dp<kStaticCpl, false>;
        ACOSS_LAUNCH(kfn, grid, 128, 0, st, A);
    }
    CU(cudaGetLastError());
    return ACOSS_OK;
}
