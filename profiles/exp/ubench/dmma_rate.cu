// DMMA (mma.sync.m8n8k4.f64) issue rate and latency on one SM: W warps, C independent accumulator chains each.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_rate dmma_rate.cu && ./dmma_rate
#include <cstdio>
#include <cuda_runtime.h>
template <int C>
__global__ void k(double* out, long long* cyc, int iters) {
    double c0[C], c1[C];
    for (int i = 0; i < C; ++i) c0[i] = c1[i] = 0.0;
    double a = threadIdx.x * 1e-3, b = 1.0 + threadIdx.x * 1e-6;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < C; ++i)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c0[i]), "+d"(c1[i]) : "d"(a), "d"(b));
    }
    long long t1 = clock64();
    __syncthreads();
    double s = 0;
    for (int i = 0; i < C; ++i) s += c0[i] + c1[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int C>
void run(int warps, int blocks) {
    double* out; long long* cyc;
    cudaMalloc(&out, sizeof(double) * 1024 * blocks); cudaMalloc(&cyc, 8 * blocks);
    const int iters = 2000;
    k<C><<<blocks, warps * 32>>>(out, cyc, iters);
    k<C><<<blocks, warps * 32>>>(out, cyc, iters);
    long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    cudaDeviceSynchronize();
    const double per = (double)h / ((double)iters * C);
    printf("chains %d warps/SM %2d blocks %d: %.2f cycles per DMMA per warp; %.2f cycles per DMMA per SM sub-partition; %.1f FMA/cycle/SM\n",
           C, warps, blocks, per, per / ((warps + 3) / 4), 256.0 * warps / per);
    cudaFree(out); cudaFree(cyc);
}
int main() {
    run<1>(1, 1); run<2>(1, 1); run<4>(1, 1); run<8>(1, 1);
    run<4>(4, 1); run<4>(8, 1); run<4>(16, 1); run<8>(16, 1); run<4>(16, 148);
    return 0;
}
