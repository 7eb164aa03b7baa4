// DMMA operand-delivery microbenchmark for B200 (sm_100a): how close does mma.sync.m8n8k4.f64 get to its register-only
// peak (tools/fp64_peak.cu: 37.1 TFLOP/s) when the operands come from shared memory every k-step, as in the GP and
// Cholesky kernels?  Variants: MI x NF accumulator fragments per warp, warps per CTA, CTAs per SM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/dmma_feed tools/dmma_feed.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1);} } while (0)

__device__ __forceinline__ void dmma(double (&c)[2], double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

// smem: KS k-steps x (MI + NF) fragments x 32 lanes, shared by all warps (broadcast-free: each warp reads the same data)
template <int MI, int NF, int KS>
__global__ void feed_kernel(double* out, int iters) {
    extern __shared__ double sm[];
    for (int i = threadIdx.x; i < KS * (MI + NF) * 32; i += blockDim.x) sm[i] = 1e-3 * (i % 97);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    double acc[MI][NF][2];
#pragma unroll
    for (int m = 0; m < MI; ++m)
#pragma unroll
        for (int n = 0; n < NF; ++n) acc[m][n][0] = acc[m][n][1] = 0.0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll 4
        for (int ks = 0; ks < KS; ++ks) {
            const double* p = sm + (size_t)ks * (MI + NF) * 32 + lane;
            double a[MI], b[NF];
#pragma unroll
            for (int m = 0; m < MI; ++m) a[m] = p[m * 32];
#pragma unroll
            for (int n = 0; n < NF; ++n) b[n] = p[(MI + n) * 32];
#pragma unroll
            for (int m = 0; m < MI; ++m)
#pragma unroll
                for (int n = 0; n < NF; ++n) dmma(acc[m][n], a[m], b[n]);
        }
    }
    double s = 0;
#pragma unroll
    for (int m = 0; m < MI; ++m)
#pragma unroll
        for (int n = 0; n < NF; ++n) s += acc[m][n][0] + acc[m][n][1];
    if (s == 123.456) out[0] = s;
}

template <int MI, int NF>
void run(const char* name, int warps, int ctas_per_sm, int sms) {
    constexpr int KS = 64;
    const int iters = 200;
    double* out; CK(cudaMalloc(&out, 8));
    const size_t smem = (size_t)KS * (MI + NF) * 32 * 8;
    CK(cudaFuncSetAttribute(feed_kernel<MI, NF, KS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        feed_kernel<MI, NF, KS><<<sms * ctas_per_sm, warps * 32, smem>>>(out, iters);
        cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    const double flops = (double)sms * ctas_per_sm * warps * iters * KS * MI * NF * 512.0;
    printf("%-10s MI=%d NF=%d warps/CTA=%2d CTAs/SM=%d  %.3f ms  %.2f TFLOP/s\n", name, MI, NF, warps, ctas_per_sm, best, flops / best * 1e-9);
    cudaFree(out);
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    const int sms = p.multiProcessorCount;
    printf("device %s, %d SMs\n", p.name, sms);
    for (int w : {4, 8, 16}) {
        for (int c : {1, 2}) {
            if (w * c > 32) continue;
            run<2, 4>("lds", w, c, sms);
            run<4, 4>("lds", w, c, sms);
            run<1, 4>("lds", w, c, sms);
            run<2, 2>("lds", w, c, sms);
        }
    }
    return 0;
}
