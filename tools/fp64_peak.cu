// FP64 pipe microbenchmark for B200 (sm_100a): DFMA vs mma.sync f64 shapes.
// Used once to pick the inner loop of the Cholesky / GEMM kernels and to give
// the FP64 roofline denominator a second, library-independent reading.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp64_peak tools/fp64_peak.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { \
  printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1);} } while (0)

__global__ void __launch_bounds__(256) k_dfma(double* out, int iters, double a, double b) {
  double acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x * 1e-9 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = fma(acc[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += acc[i];
  if (s == 123.456) out[0] = s;
}

// NT independent accumulator tiles per warp
template <int NT>
__global__ void __launch_bounds__(256) k_dmma884(double* out, int iters, double a, double b) {
  double c[NT][2];
#pragma unroll
  for (int i = 0; i < NT; ++i) { c[i][0] = i; c[i][1] = threadIdx.x; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NT; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NT; ++i) s += c[i][0] + c[i][1];
  if (s == 123.456) out[0] = s;
}

template <int NT>
__global__ void __launch_bounds__(256) k_dmma1684(double* out, int iters, double a, double b) {
  double c[NT][4];
#pragma unroll
  for (int i = 0; i < NT; ++i) { c[i][0] = i; c[i][1] = threadIdx.x; c[i][2] = 1; c[i][3] = 2; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NT; ++i)
      asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                   : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3]) : "d"(a), "d"(b), "d"(b));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NT; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  if (s == 123.456) out[0] = s;
}

template <int NT>
__global__ void __launch_bounds__(256) k_dmma1688(double* out, int iters, double a, double b) {
  double c[NT][4];
#pragma unroll
  for (int i = 0; i < NT; ++i) { c[i][0] = i; c[i][1] = threadIdx.x; c[i][2] = 1; c[i][3] = 2; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NT; ++i)
      asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                   : "d"(a), "d"(b), "d"(a), "d"(b), "d"(b), "d"(a));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NT; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  if (s == 123.456) out[0] = s;
}

template <int NT>
__global__ void __launch_bounds__(256) k_dmma16816(double* out, int iters, double a, double b) {
  double c[NT][4];
#pragma unroll
  for (int i = 0; i < NT; ++i) { c[i][0] = i; c[i][1] = threadIdx.x; c[i][2] = 1; c[i][3] = 2; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NT; ++i)
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
                   : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                   : "d"(a), "d"(b), "d"(a), "d"(b), "d"(a), "d"(b), "d"(a), "d"(b), "d"(b), "d"(a), "d"(b), "d"(a));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NT; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  if (s == 123.456) out[0] = s;
}

template <typename F>
static void run(const char* name, F launch, double flop_per_launch) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  for (int i = 0; i < 3; ++i) launch();
  CK(cudaDeviceSynchronize());
  float best = 1e30f, tot = 0;
  const int reps = 10;
  for (int i = 0; i < reps; ++i) {
    CK(cudaEventRecord(e0)); launch(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
    tot += ms;
  }
  CK(cudaGetLastError());
  printf("%-28s best %8.3f ms  %7.2f TFLOP/s   mean %8.3f ms  %7.2f TFLOP/s\n", name, best,
         flop_per_launch / best * 1e-9, tot / reps, flop_per_launch / (tot / reps) * 1e-9);
}

int main(int argc, char** argv) {
  int dev = 0; CK(cudaSetDevice(dev));
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, dev));
  printf("device %s, %d SMs, clock %d kHz\n", p.name, p.multiProcessorCount, p.clockRate);
  double* out; CK(cudaMalloc(&out, 8));
  const int sms = p.multiProcessorCount;
  const int iters = 20000;
  for (int bps = 1; bps <= 8; bps *= 2) {
    const int grid = sms * bps, thr = 256;
    const double warps = (double)grid * thr / 32;
    char nm[64];
    printf("--- %d blocks/SM x %d threads\n", bps, thr);
    snprintf(nm, 64, "dfma x16 chains");
    run(nm, [&] { k_dfma<<<grid, thr>>>(out, iters, 1.0000001, 1e-9); }, (double)grid * thr * iters * 16 * 2);
    run("dmma m8n8k4 x4", [&] { k_dmma884<4><<<grid, thr>>>(out, iters, 1.0000001, 1e-9); }, warps * iters * 4 * 8 * 8 * 4 * 2);
    run("dmma m8n8k4 x8", [&] { k_dmma884<8><<<grid, thr>>>(out, iters, 1.0000001, 1e-9); }, warps * iters * 8 * 8 * 8 * 4 * 2);
    run("dmma m8n8k4 x16", [&] { k_dmma884<16><<<grid, thr>>>(out, iters, 1.0000001, 1e-9); }, warps * iters * 16 * 8 * 8 * 4 * 2);
    run("dmma m16n8k4 x4", [&] { k_dmma1684<4><<<grid, thr>>>(out, iters, 1.0000001, 1e-9); }, warps * iters * 4 * 16 * 8 * 4 * 2);
    run("dmma m16n8k8 x4", [&] { k_dmma1688<4><<<grid, thr>>>(out, iters / 2, 1.0000001, 1e-9); }, warps * (iters / 2) * 4 * 16 * 8 * 8 * 2);
    run("dmma m16n8k8 x8", [&] { k_dmma1688<8><<<grid, thr>>>(out, iters / 2, 1.0000001, 1e-9); }, warps * (iters / 2) * 8 * 16 * 8 * 8 * 2);
    run("dmma m16n8k16 x4", [&] { k_dmma16816<4><<<grid, thr>>>(out, iters / 4, 1.0000001, 1e-9); }, warps * (iters / 4) * 4 * 16 * 8 * 16 * 2);
    run("dmma m16n8k16 x8", [&] { k_dmma16816<8><<<grid, thr>>>(out, iters / 4, 1.0000001, 1e-9); }, warps * (iters / 4) * 8 * 16 * 8 * 16 * 2);
  }
  return 0;
}
