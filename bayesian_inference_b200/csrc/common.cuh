// common.cuh -- shared device helpers for the sm_100a log-posterior kernels.
//
// All arithmetic is FP64.  Dense contractions go through the FP64 tensor pipe with mma.sync m8n8k4 (SASS DMMA.8x8x4):
// tcgen05 / TMEM have no f64 kind, so DMMA is the Blackwell FP64 tensor path (measured 37.1 TFLOP/s on B200,
// profiles/r01_fp64_peak_microbench.txt).
//
// Tile layouts (all tiles are 32 x 32 unless a K extent is given; lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3)
//
//   operand-fragment-major ("OF"):  idx = ((mf * nks + ks) * 32 + lane),  value = tile[mf*8 + g][ks*4 + t]
//       -- one coalesced 256-byte load per warp gives the A fragment (rows mf*8.., k-slice ks) of D = A * B^T,
//          and, because B^T is taken, the very same layout of the other operand's rows gives the B fragment.
//   accumulator-fragment-major ("CF"):  idx = ((mf*4 + nf) * 32 + lane) * 2 + e,
//          value = tile[mf*8 + g][nf*8 + 2*t + e]      -- what mma.sync leaves in registers.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define BI_TILE 32

namespace bi {

__device__ __forceinline__ void dmma(double (&c)[2], double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
        : "+d"(c[0]), "+d"(c[1])
        : "d"(a), "d"(b));
}

__device__ __forceinline__ double ld_cg(const double* p) {   // L2 only: streamed per-walker factor tiles
    double v;
    asm volatile("ld.global.cg.f64 %0, [%1];" : "=d"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ double ld_ca(const double* p) {   // L1 + L2: operand shared by the warps of a CTA
    double v;
    asm volatile("ld.global.ca.f64 %0, [%1];" : "=d"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_cg(double* p, double v) {
    asm volatile("st.global.cg.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}

__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }

// Programmatic dependent launch (the kernels of a sampler half-step, bi_api.cu launch_k): a kernel launched with the
// programmatic-serialisation attribute may START before the previous kernel of the stream has finished.  pdl_wait() returns
// when that kernel has completed and its writes are visible: everything a kernel reads that another kernel of the stream may
// have written, and every global write, comes after it.  pdl_trigger() lets the NEXT kernel start; it is always issued after
// pdl_wait(), so a kernel that runs ahead can only ever overlap its direct predecessor, never an older kernel.  Both are
// no-ops for a kernel launched the ordinary way.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// offset of element (r, c) of a 32 x (4*nks) tile stored operand-fragment-major
__host__ __device__ __forceinline__ int of_index(int r, int c, int nks) {
    return (((r >> 3) * nks + (c >> 2)) << 5) + ((r & 7) << 2) + (c & 3);
}
// offset of element (r, c) of a 32 x 32 tile stored accumulator-fragment-major
__host__ __device__ __forceinline__ int cf_index(int r, int c) {
    return (((((r >> 3) << 2) + (c >> 3)) << 5) + ((r & 7) << 2) + ((c & 7) >> 1)) * 2 + (c & 1);
}

// Convert the accumulator fragments of one 8-row strip (4 n-fragments = 32 columns) into the 8 operand fragments
// (k-slices of 4 columns) of the same strip.  Pure quad shuffles, no shared memory.
__device__ __forceinline__ void cf_to_of(const double (&c)[4][2], double (&a)[8], int lane) {
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
        const int src = (lane & 28) | ((ks & 1) << 1) | ((lane & 3) >> 1);
        const double v0 = shfl_d(c[ks >> 1][0], src);
        const double v1 = shfl_d(c[ks >> 1][1], src);
        a[ks] = (lane & 1) ? v1 : v0;
    }
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace bi
