// Microbenchmark of the sweep's inner loop structure on B200: per "row" 4 chains x 12 DFMA with
// distinct register operands (x from shared memory, y2 resident), optionally ring traffic.
#include <cstdio>
#include <cuda_runtime.h>
template <int VARIANT>
__global__ void __launch_bounds__(128) k_rows(double* out, const double* yin, int rows, long long* cyc) {
    __shared__ __align__(16) double xs[32 * 12];
    __shared__ double ring[10 * 4 * 132];
    const int tid = threadIdx.x;
    for (int i = tid; i < 32 * 12; i += blockDim.x) xs[i] = 1.0 + 1e-3 * i;
    for (int i = tid; i < 10 * 4 * 132; i += blockDim.x) ring[i] = 0.0;
    double y2[4][12], P[4] = {0, 0, 0, 0};
    for (int r = 0; r < 4; ++r)
        for (int k = 0; k < 12; ++k) y2[r][k] = yin[(tid * 4 + r) * 12 + k];
    __syncthreads();
    long long t0 = clock64();
    int slot = 0;
    double sink = 0.0;
    for (int a = 0; a < rows; ++a) {
        const double* xr = xs + (a & 31) * 12;
        double x[12];
#pragma unroll
        for (int k = 0; k < 12; k += 2) {
            const double2 v = *reinterpret_cast<const double2*>(xr + k);
            x[k] = v.x; x[k + 1] = v.y;
        }
        double acc[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            acc[r] = (r == 0) ? 0.0 : P[r - 1];
#pragma unroll
            for (int k = 0; k < 12; ++k) acc[r] = fma(x[k], y2[r][k], acc[r]);
        }
        if (VARIANT >= 1) {
            double* r0 = ring + slot * 528 + tid + 4;
            const double* rm = ring + ((slot + 1) % 10) * 528 + tid + 4;
            acc[0] += ring[((slot + 9) % 10) * 528 + 3 * 132 + tid + 3];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                sink += (acc[r] - rm[r * 132 - 2]) + 768.0;
                r0[r * 132] = acc[r];
            }
            slot = (slot + 1 == 10) ? 0 : slot + 1;
            if (VARIANT == 1) __syncwarp();
            if (VARIANT == 2) __syncthreads();
        }
#pragma unroll
        for (int r = 0; r < 4; ++r) P[r] = acc[r];
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + tid] = P[0] + P[1] + P[2] + P[3] + sink;
    if (tid == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
// Variant 3: the 12-DFMA chains start from zero (independent of the previous row) and are issued
// one row ahead; the diagonal carry is one DADD per cell.
__global__ void __launch_bounds__(128) k_rows_pipe(double* out, const double* yin, int rows, long long* cyc, int sync_mode) {
    __shared__ __align__(16) double xs[32 * 12];
    __shared__ double ring[10 * 4 * 132];
    const int tid = threadIdx.x;
    for (int i = tid; i < 32 * 12; i += blockDim.x) xs[i] = 1.0 + 1e-3 * i;
    for (int i = tid; i < 10 * 4 * 132; i += blockDim.x) ring[i] = 0.0;
    double y2[4][12], P[4] = {0, 0, 0, 0}, dot[4];
    for (int r = 0; r < 4; ++r)
        for (int k = 0; k < 12; ++k) y2[r][k] = yin[(tid * 4 + r) * 12 + k];
    __syncthreads();
    long long t0 = clock64();
    int slot = 0;
    double sink = 0.0;
    {
        const double* xr = xs;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            dot[r] = 0.0;
#pragma unroll
            for (int k = 0; k < 12; ++k) dot[r] = fma(xr[k], y2[r][k], dot[r]);
        }
    }
    for (int a = 0; a < rows; ++a) {
        double Pn[4];
        Pn[0] = ring[((slot + 9) % 10) * 528 + 3 * 132 + tid + 3] + dot[0];
#pragma unroll
        for (int r = 1; r < 4; ++r) Pn[r] = P[r - 1] + dot[r];
        // next row's dots
        const double* xr = xs + ((a + 1) & 31) * 12;
        double x[12];
#pragma unroll
        for (int k = 0; k < 12; k += 2) {
            const double2 v = *reinterpret_cast<const double2*>(xr + k);
            x[k] = v.x; x[k + 1] = v.y;
        }
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            dot[r] = 0.0;
#pragma unroll
            for (int k = 0; k < 12; ++k) dot[r] = fma(x[k], y2[r][k], dot[r]);
        }
        double* r0 = ring + slot * 528 + tid + 4;
        const double* rm = ring + ((slot + 1) % 10) * 528 + tid + 4;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            sink += (Pn[r] - rm[r * 132 - 2]) + 768.0;
            r0[r * 132] = Pn[r];
            P[r] = Pn[r];
        }
        slot = (slot + 1 == 10) ? 0 : slot + 1;
        if (sync_mode == 1) __syncwarp(); else __syncthreads();
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + tid] = P[0] + P[1] + P[2] + P[3] + sink + dot[0];
    if (tid == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
void run_pipe(int blocks, int sync_mode, const char* tag) {
    double *out, *yin; long long* cyc;
    cudaMalloc(&out, 8 * blocks * 128); cudaMalloc(&yin, 8 * 128 * 48); cudaMalloc(&cyc, 8);
    cudaMemset(yin, 0, 8 * 128 * 48);
    const int rows = 20000;
    k_rows_pipe<<<blocks, 128>>>(out, yin, rows, cyc, sync_mode);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k_rows_pipe<<<blocks, 128>>>(out, yin, rows, cyc, sync_mode);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-44s blocks/SM %d: %.0f cyc per row per CTA, %.1f cyc per row per SM, %.2f TFLOP/s (DFMA only)\n", tag, blocks / 148,
           (double)c / rows, (double)c / rows / (blocks / 148), 2.0 * 48 * 128 * (double)blocks * rows / (ms * 1e-3) / 1e12);
}
template <int V>
void run(int blocks, const char* tag) {
    double *out, *yin; long long* cyc;
    cudaMalloc(&out, 8 * blocks * 128); cudaMalloc(&yin, 8 * 128 * 48); cudaMalloc(&cyc, 8);
    cudaMemset(yin, 0, 8 * 128 * 48);
    const int rows = 20000;
    k_rows<V><<<blocks, 128>>>(out, yin, rows, cyc);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k_rows<V><<<blocks, 128>>>(out, yin, rows, cyc);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-44s blocks/SM %d: %.0f cyc per row per CTA, %.1f cyc per row per SM, %.2f TFLOP/s (DFMA only)\n", tag, blocks / 148,
           (double)c / rows, (double)c / rows / (blocks / 148), 2.0 * 48 * 128 * (double)blocks * rows / (ms * 1e-3) / 1e12);
}
int main() {
    run<0>(148, "dots only, 1 CTA/SM");
    run<0>(148 * 3, "dots only, 3 CTA/SM");
    run<1>(148, "dots + ring + syncwarp, 1 CTA/SM");
    run<1>(148 * 3, "dots + ring + syncwarp, 3 CTA/SM");
    run<2>(148 * 3, "dots + ring + syncthreads, 3 CTA/SM");
    run_pipe(148, 1, "pipelined dots + ring + syncwarp, 1 CTA/SM");
    run_pipe(148 * 3, 1, "pipelined dots + ring + syncwarp, 3 CTA/SM");
    run_pipe(148 * 3, 2, "pipelined dots + ring + syncthreads, 3 CTA/SM");
    run_pipe(148 * 2, 1, "pipelined dots + ring + syncwarp, 2 CTA/SM");
    return 0;
}
