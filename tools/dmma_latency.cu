// DMMA latency microbenchmark for B200 (sm_100a): cycles per mma.sync.m8n8k4.f64 in ONE warp when every DMMA depends on the
// previous one (1 chain) and with 2 / 4 / 8 independent accumulator chains, alone on the SM and with 2 warps per sub-partition.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/dmma_latency tools/dmma_latency.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma(double (&c)[2], double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
template <int CH>
__global__ void lat_kernel(double* out, long long* cyc, int iters) {
    double acc[CH][2];
#pragma unroll
    for (int c = 0; c < CH; ++c) acc[c][0] = acc[c][1] = 0.0;
    const double a = 1e-3 * (threadIdx.x % 7), b = 1e-3 * (threadIdx.x % 5);
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int c = 0; c < CH; ++c) dmma(acc[c], a, b);
    }
    const long long t1 = clock64();
    double s = 0.0;
#pragma unroll
    for (int c = 0; c < CH; ++c) s += acc[c][0] + acc[c][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int CH>
void run(int warps, double* out, long long* cyc) {
    const int iters = 2000;
    lat_kernel<CH><<<1, warps * 32>>>(out, cyc, iters);
    lat_kernel<CH><<<1, warps * 32>>>(out, cyc, iters);
    long long h = 0;
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("chains %d, warps/CTA %2d: %.1f cycles per DMMA per warp, %.1f per dependent step\n", CH, warps, (double)h / (iters * 8.0 * CH),
           (double)h / (iters * 8.0));
}
int main() {
    double* out; long long* cyc;
    cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 8);
    for (int warps : {1, 4, 8, 16}) { run<1>(warps, out, cyc); run<2>(warps, out, cyc); run<4>(warps, out, cyc); run<8>(warps, out, cyc); }
    cudaError_t e = cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
