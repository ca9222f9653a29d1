// sweep7.h -- what api.cu needs from sweep7.cu (host-only declarations, no kernel templates).
#pragma once
#include <cstddef>
#include <cstdint>

struct acoss_ctx;

// column capacity classes of k_sweep7 (mirrors kernels_sweep7.cuh; static_asserts in sweep7.cu keep them in step)
constexpr int kSweep7Buckets = 8;
constexpr int kSweep7Pad = 9;
constexpr int kSweep7MaxCols = 1024 - kSweep7Pad;
inline int sweep7_bucket_for(int Lc) {  // -1: the column song does not fit
    const int need = Lc + kSweep7Pad;
    return need <= 1024 ? (need <= 128 ? 0 : (need - 1) / 128) : -1;   // capacity 128 * (bucket + 1)
}

int acoss_launch_sweep7(acoss_ctx* c, int mode, int bins, const uint32_t* d_items, int n_items, int bucket, unsigned int* d_counter,
                        uint32_t flags);
size_t acoss_sweep7_smem(int bucket, int bins, int mode);
int acoss_sweep7_stats(acoss_ctx* c, unsigned long long out[16], int reset);
