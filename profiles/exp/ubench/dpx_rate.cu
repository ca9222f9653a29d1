#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(unsigned* out, unsigned a0, unsigned b0, int iters) {
    unsigned x[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) x[u] = a0 + threadIdx.x * (u + 1);
    unsigned b = b0, c = b0 * 3;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            if (MODE == 0) x[u] = __viaddmax_s16x2(x[u], b, c);
            if (MODE == 1) x[u] = __vimax3_s16x2(x[u], b, c);
            if (MODE == 2) x[u] = __viaddmax_s16x2_relu(x[u], b, c);
            if (MODE == 3) x[u] = (unsigned)__vimax3_s32((int)x[u], (int)b, (int)c);
            if (MODE == 4) x[u] = (unsigned)__viaddmax_s32((int)x[u], (int)b, (int)c);
            if (MODE == 5) x[u] = (unsigned)max((int)x[u] + (int)b, 0);
            if (MODE == 6) x[u] = x[u] * b + c;
        }
        b += 0x00010001u; c ^= b;
    }
    unsigned s = 0;
#pragma unroll
    for (int u = 0; u < 8; ++u) s ^= x[u];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE> void run(const char* name, unsigned* d) {
    const int iters = 4096, grid = 148 * 8, block = 256;
    k<MODE><<<grid, block>>>(d, 1, 2, 16);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<MODE><<<grid, block>>>(d, 1, 2, iters);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double ops = (double)grid * block * iters * 8;
    printf("%-28s %.3f ms  %.1f Gop/s  = %.1f lane-ops/clk/SM at 1.965 GHz\n", name, ms, ops / ms / 1e6, ops / (ms * 1e-3) / 148 / 1.965e9);
}
int main() {
    unsigned* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    run<0>("viaddmax_s16x2", d); run<1>("vimax3_s16x2", d); run<2>("viaddmax_s16x2_relu", d);
    run<3>("vimax3_s32", d); run<4>("viaddmax_s32", d); run<5>("max(a+b,0) s32", d); run<6>("imad", d);
    return 0;
}
