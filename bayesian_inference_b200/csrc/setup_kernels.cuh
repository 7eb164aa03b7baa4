// setup_kernels.cuh -- one-off kernels that turn the uploaded fitted objects into the device image (run in bi_create).
// None of these runs per walker.
#pragma once
#include "common.cuh"

namespace bi {

// X = L^-1 for a batch of lower-triangular matrices: grid (ceil(N/32), P), 32 threads, smem N*32 doubles.
// Forward substitution on the identity, one column per lane (the arithmetic of LAPACK dtrtri / dtrtrs).
__global__ void __launch_bounds__(32) tri_invert_kernel(const double* __restrict__ L, double* __restrict__ X, int N) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* sx = reinterpret_cast<double*>(smem_raw);   // [N][32]
    const int lane = threadIdx.x;
    const int c0 = blockIdx.x * 32;
    const int c = c0 + lane;
    const double* Lp = L + (size_t)blockIdx.y * N * N;
    double* Xp = X + (size_t)blockIdx.y * N * N;
    for (int r = 0; r < N; ++r) {
        double val = 0.0;
        if (r >= c0) {
            const double* Lr = Lp + (size_t)r * N;
            double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
            int k = c0;
            for (; k + 3 < r; k += 4) {
                s0 = fma(__ldg(Lr + k), sx[(k)*32 + lane], s0);
                s1 = fma(__ldg(Lr + k + 1), sx[(k + 1) * 32 + lane], s1);
                s2 = fma(__ldg(Lr + k + 2), sx[(k + 2) * 32 + lane], s2);
                s3 = fma(__ldg(Lr + k + 3), sx[(k + 3) * 32 + lane], s3);
            }
            for (; k < r; ++k) s0 = fma(__ldg(Lr + k), sx[k * 32 + lane], s0);
            const double s = (s0 + s1) + (s2 + s3);
            val = ((r == c ? 1.0 : 0.0) - s) / __ldg(Lr + r);
        }
        sx[r * 32 + lane] = val;
        if (c < N) Xp[(size_t)r * N + c] = val;
    }
}

// Pack L^-1 (N x N row-major, lower) plus alpha as row N into operand-fragment-major n-tiles.
// out: [ntn][4][nks][32];   one thread per output element.
__global__ void pack_linv_kernel(const double* __restrict__ X, const double* __restrict__ alpha, double* __restrict__ out,
                                 int N, int nks, int ntn) {
    const size_t per_pc = (size_t)ntn * 4 * nks * 32;
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= per_pc) return;
    const int p = blockIdx.y;
    const int lane = (int)(idx & 31);
    size_t rest = idx >> 5;
    const int ks = (int)(rest % nks);
    rest /= nks;
    const int nf = (int)(rest & 3);
    const int tile = (int)(rest >> 2);
    const int n = tile * 32 + nf * 8 + (lane >> 2);
    const int k = ks * 4 + (lane & 3);
    double val = 0.0;
    if (k < N) {
        if (n < N) {
            if (k <= n) val = X[((size_t)p * N + n) * N + k];
        } else if (n == N) {
            val = alpha[(size_t)p * N + k];
        }
    }
    out[(size_t)p * per_pc + idx] = val;
}

// Us[r][c] = (sum_q S_rest[q][r] * (ev[q] * S_rest[q][c])) * (scale[r] * scale[c])      emulation.py:227-251, :538
__global__ void __launch_bounds__(256) unexplained_kernel(const double* __restrict__ S_rest, const double* __restrict__ ev,
                                                          const double* __restrict__ scale, double* __restrict__ Us,
                                                          int Fg, int R) {
    const int c = blockIdx.x * 32 + (threadIdx.x & 31);
    const int r = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (r >= Fg || c >= Fg) return;
    double s = 0.0;
    for (int q = 0; q < R; ++q) s = fma(__ldg(S_rest + (size_t)q * Fg + r), __ldg(ev + q) * __ldg(S_rest + (size_t)q * Fg + c), s);
    Us[(size_t)r * Fg + c] = s * (scale[r] * scale[c]);
}

// A = Sigma_exp + Ublk / n_div as accumulator-fragment-major lower tiles (index i(i+1)/2 + j); identity padding
// (with aug != 0 the diagonal entry of row F is 0: the augmented dY row of the pipelined Cholesky kernel).
// grid: (ntiles_lower), 256 threads (4 elements each).
__global__ void __launch_bounds__(256) build_a_tiles_kernel(const double* __restrict__ Sig, const double* __restrict__ yerr,
                                                            const double* __restrict__ Ublk, double n_div, int F,
                                                            int aug, double* __restrict__ out) {
    // decode (i, j) from the linear lower-triangle index
    const int lin = blockIdx.x;
    int i = (int)((sqrt(8.0 * lin + 1.0) - 1.0) * 0.5);
    while ((long long)(i + 1) * (i + 2) / 2 <= lin) ++i;
    while ((long long)i * (i + 1) / 2 > lin) --i;
    const int j = lin - i * (i + 1) / 2;
    for (int el = threadIdx.x; el < 1024; el += 256) {
        const int rr = el >> 5, cc = el & 31;
        const int r = i * 32 + rr, c = j * 32 + cc;
        double val;
        if (r < F && c < F) {
            double sg;
            if (Sig) sg = Sig[(size_t)r * F + c];
            else sg = (r == c) ? yerr[r] * yerr[r] : 0.0;
            val = sg + Ublk[(size_t)r * F + c] / n_div;
        } else {
            val = (r == c && !(aug && r == F)) ? 1.0 : 0.0;   // aug: M[F][F] = 0 (the dY row of chol_pipe.cuh)
        }
        out[(size_t)lin * 1024 + cf_index(rr, cc)] = val;
    }
}

// number of rows strictly inside the prior box (log_posterior.py:63-67)
__global__ void count_inside_kernel(const double* __restrict__ theta, long long W, int d, const double* __restrict__ lo,
                                    const double* __restrict__ hi, unsigned long long* __restrict__ count) {
    const long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    int inside = 0;
    if (w < W) {
        inside = 1;
        for (int k = 0; k < d; ++k) {
            const double x = theta[w * d + k];
            if (!(x > lo[k] && x < hi[k])) inside = 0;
        }
    }
    const unsigned b = __ballot_sync(0xffffffffu, inside);
    if ((threadIdx.x & 31) == 0 && b) atomicAdd(count, (unsigned long long)__popc(b));
}

}  // namespace bi
