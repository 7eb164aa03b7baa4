// cov_assemble.cuh -- K3: inverse-PCA mean and (standalone) covariance assembly for the predict() API.
//
// Replaces (reference src/bayesian_inference/emulation.py):
//   :506-509  mean = (m @ S_P) * scale + scaler.mean_                (no pca.mean_)
//   :522-539  cov  = (S diag(v) S^T + U / n_div) * outer(scale, scale)
//   :346-406  merge over groups: block diagonal per observable (mask), sorted-observable order
//
// The covariance is written to HBM here only because predict() returns it; the log-posterior path (chol_loglik.cuh)
// never materialises it.  Bound: HBM writes, 8 F^2 bytes per walker.
#pragma once
#include "common.cuh"

namespace bi {

struct CovParams {
    int Fo;               // output dimension (F merged, or F_g)
    int K0;               // padded number of principal components (columns of T)
    int Ptot;             // row stride of m / v
    int voff;             // first column of m / v used by T's column 0
    int nq;               // number of valid columns (<= K0)
    const double* Trow;   // (Fo, K0) row-major, T[r][q] = scale[r] * S[q][r]
    const double* Uc;     // (Fo, Fo) unexplained covariance already multiplied by outer(scale, scale) (and masked)
    const double* mu;     // (Fo) scaler mean
    const int* obs;       // (Fo) observable id per row, or nullptr (no mask: one emulation group on its own)
    const double* m;      // (W, Ptot)
    const double* v;      // (W, Ptot)
    double n_div;
    double* mean_out;     // (W, Fo) or nullptr
    double* cov_out;      // (W, Fo, Fo) or nullptr
    long long W;
};

__global__ void __launch_bounds__(256) mean_kernel(const CovParams P) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P.W * P.Fo) return;
    const long long w = idx / P.Fo;
    const int r = (int)(idx - w * P.Fo);
    double s = 0.0;
    for (int q = 0; q < P.nq; ++q) s = fma(P.m[w * P.Ptot + P.voff + q], __ldg(P.Trow + (long long)r * P.K0 + q), s);
    P.mean_out[idx] = s + __ldg(P.mu + r);
}

// grid (ceil(Fo/32), W), 256 threads; dynamic smem = (2 * 32 * (K0 + 1) + K0) doubles
__global__ void __launch_bounds__(256) cov_assemble_kernel(const CovParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int K0 = P.K0, ld = K0 + 1;
    double* sTi = reinterpret_cast<double*>(smem_raw);   // rows of tile i, pre-multiplied by v
    double* sTj = sTi + 32 * ld;
    double* sv = sTj + 32 * ld;
    const int tid = threadIdx.x;
    const long long w = blockIdx.y;
    const int i0 = blockIdx.x * 32;
    for (int q = tid; q < K0; q += 256) sv[q] = (q < P.nq) ? P.v[w * P.Ptot + P.voff + q] : 0.0;
    __syncthreads();
    for (int idx = tid; idx < 32 * K0; idx += 256) {
        const int r = idx / K0, q = idx - r * K0;
        sTi[r * ld + q] = (i0 + r < P.Fo) ? __ldg(P.Trow + (long long)(i0 + r) * K0 + q) * sv[q] : 0.0;
    }
    const int c = tid & 31, r0 = tid >> 5;
    const int ntl = (P.Fo + 31) / 32;
    double* out = P.cov_out + w * (long long)P.Fo * P.Fo;
    for (int jt = 0; jt < ntl; ++jt) {
        const int j0 = jt * 32;
        __syncthreads();
        for (int idx = tid; idx < 32 * K0; idx += 256) {
            const int r = idx / K0, q = idx - r * K0;
            sTj[r * ld + q] = (j0 + r < P.Fo) ? __ldg(P.Trow + (long long)(j0 + r) * K0 + q) : 0.0;
        }
        __syncthreads();
        double acc[4] = {0.0, 0.0, 0.0, 0.0};
        for (int q = 0; q < K0; ++q) {
            const double b = sTj[c * ld + q];
#pragma unroll
            for (int rr = 0; rr < 4; ++rr) acc[rr] = fma(sTi[(r0 + 8 * rr) * ld + q], b, acc[rr]);
        }
        const int col = j0 + c;
        if (col < P.Fo) {
            const int oc = P.obs ? P.obs[col] : 0;
#pragma unroll
            for (int rr = 0; rr < 4; ++rr) {
                const int row = i0 + r0 + 8 * rr;
                if (row < P.Fo) {
                    const bool same = P.obs ? (P.obs[row] == oc) : true;
                    const double u = __ldg(P.Uc + (long long)row * P.Fo + col) / P.n_div;
                    out[(long long)row * P.Fo + col] = same ? acc[rr] + u : 0.0;
                }
            }
        }
    }
}

}  // namespace bi
