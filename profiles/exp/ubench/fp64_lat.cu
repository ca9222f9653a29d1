// Microbenchmark: DFMA dependent-issue latency and per-SM throughput on B200 (sm_100a).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_lat fp64_lat.cu && ./fp64_lat
#include <cstdio>
#include <cuda_runtime.h>
template <int CHAINS>
__global__ void k_chain(double* out, double a, double b, int iters, long long* cyc) {
    double acc[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) acc[c] = threadIdx.x + c;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int c = 0; c < CHAINS; ++c) acc[c] = fma(acc[c], a, b);
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) s += acc[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int CHAINS>
void run(int blocks, int threads, const char* tag) {
    double* out; long long* cyc; cudaMalloc(&out, sizeof(double) * blocks * threads); cudaMalloc(&cyc, 8);
    const int iters = 2000;
    k_chain<CHAINS><<<blocks, threads>>>(out, 1.0000001, 1e-9, iters, cyc);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k_chain<CHAINS><<<blocks, threads>>>(out, 1.0000001, 1e-9, iters, cyc);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    double n = (double)iters * 8 * CHAINS;
    printf("%-28s chains %d blocks %d threads %d: %.2f cyc per DFMA per warp (block 0), %.2f TFLOP/s\n", tag, CHAINS, blocks, threads,
           (double)c / n, 2.0 * n * blocks * threads / (ms * 1e-3) / 1e12);
    cudaFree(out); cudaFree(cyc);
}
int main() {
    run<1>(1, 32, "latency (1 warp, 1 chain)");
    run<2>(1, 32, "1 warp 2 chains");
    run<4>(1, 32, "1 warp 4 chains");
    run<8>(1, 32, "1 warp 8 chains");
    run<4>(148, 128, "1 warp/sched, 4 chains");
    run<4>(148, 384, "3 warps/sched, 4 chains");
    run<4>(148 * 2, 512, "8 warps/sched, 4 chains");
    run<8>(148 * 2, 512, "8 warps/sched, 8 chains");
    return 0;
}
