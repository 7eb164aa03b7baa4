// woodbury.cuh -- K5: matrix-determinant-lemma / Woodbury evaluation of the log-likelihood (BI_MODE_WOODBURY).
//
// Valid when Sigma_exp does not couple observable blocks.  For each observable block b (f_b rows, emulation group g
// with P = n_pc(g) principal components) the per-walker covariance is a rank-P update of a theta-independent matrix:
//     C_b = A_b + T_b D T_b^T ,   A_b = Sigma_exp[b,b] + (U[b,b] * s s^T) / n_div ,   D = diag(v)
// so with  G_b = T_b^T A_b^-1 T_b (P x P),  h_b = T_b^T A_b^-1 e_b,  q0_b = e_b^T A_b^-1 e_b,  e_b = mu_b - y_b:
//     w      = G_b m + h_b
//     M      = I + D^1/2 G_b D^1/2                      (P x P, SPD also when some v_p = 0)
//     dy^T C_b^-1 dy = (q0_b + m.(h_b + w)) - (D^1/2 w)^T M^-1 (D^1/2 w)
//     log det C_b    = log det A_b + log det M
// Only the P x P matrix M is factorised per walker (by the explicit-matrix Cholesky kernel); everything that involves
// f_b is folded into G_b, h_b, q0_b, log det A_b once per (handle, n_div) by woodbury_setup_kernel.
// This restates the same Gaussian log-likelihood as log_posterior.py:104-146; it must agree with BI_MODE_DENSE to
// rounding (tests/test_gpu_parity.py) and is reported separately by bench.py.
#pragma once
#include "common.cuh"

namespace bi {

struct WbBlock {
    int row0;       // first merged row of the block
    int f;          // rows
    int col0;       // first column of T_merged of the block's group
    int P;          // n_pc of the group
    long long a_off;   // offset (doubles) of this block's f x f scratch in Awork
    long long y_off;   // offset of its f x (P+1) scratch in Ywork
    long long g_off;   // offset of its Pmax x Pmax G in Gall
};

// One CTA per block: factor A_b, Y = L^-1 [T_b | e_b], then G, h, q0, log det A.  One-off per n_div.
__global__ void __launch_bounds__(256) woodbury_setup_kernel(const WbBlock* __restrict__ blocks, int F, int K0, int Pmax,
                                                             const double* __restrict__ Sig, const double* __restrict__ yerr,
                                                             const double* __restrict__ Ublk, double n_div,
                                                             const double* __restrict__ Trow, const double* __restrict__ e,
                                                             double* __restrict__ Awork, double* __restrict__ Ywork,
                                                             double* __restrict__ Gall, double* __restrict__ hall,
                                                             double* __restrict__ q0all, double* __restrict__ ldall) {
    const WbBlock B = blocks[blockIdx.x];
    const int f = B.f, P = B.P, tid = threadIdx.x, nth = blockDim.x;
    double* A = Awork + B.a_off;
    double* Y = Ywork + B.y_off;
    const int ldy = P + 1;
    for (int idx = tid; idx < f * f; idx += nth) {
        const int r = idx / f, c = idx - r * f;
        const int gr = B.row0 + r, gc = B.row0 + c;
        const double sg = Sig ? Sig[(size_t)gr * F + gc] : (r == c ? yerr[gr] * yerr[gr] : 0.0);
        A[idx] = sg + Ublk[(size_t)gr * F + gc] / n_div;
    }
    for (int idx = tid; idx < f * ldy; idx += nth) {
        const int r = idx / ldy, q = idx - r * ldy;
        Y[idx] = (q < P) ? Trow[(size_t)(B.row0 + r) * K0 + B.col0 + q] : e[B.row0 + r];
    }
    __syncthreads();
    // right-looking Cholesky, lower triangle, in place
    for (int c = 0; c < f; ++c) {
        const double d = sqrt(A[(size_t)c * f + c]);
        __syncthreads();
        if (tid == 0) A[(size_t)c * f + c] = d;
        for (int r = c + 1 + tid; r < f; r += nth) A[(size_t)r * f + c] /= d;
        __syncthreads();
        const int rem = f - c - 1;
        for (long long idx = tid; idx < (long long)rem * rem; idx += nth) {
            const int r = c + 1 + (int)(idx / rem), cc = c + 1 + (int)(idx % rem);
            if (cc <= r) A[(size_t)r * f + cc] = fma(-A[(size_t)r * f + c], A[(size_t)cc * f + c], A[(size_t)r * f + cc]);
        }
        __syncthreads();
    }
    // forward substitution, one thread per right-hand side
    for (int q = tid; q < ldy; q += nth) {
        for (int r = 0; r < f; ++r) {
            double s = Y[(size_t)r * ldy + q];
            for (int k = 0; k < r; ++k) s = fma(-A[(size_t)r * f + k], Y[(size_t)k * ldy + q], s);
            Y[(size_t)r * ldy + q] = s / A[(size_t)r * f + r];
        }
    }
    __syncthreads();
    double* G = Gall + B.g_off;
    for (int idx = tid; idx < Pmax * Pmax; idx += nth) {
        const int p = idx / Pmax, q = idx - p * Pmax;
        double s = 0.0;
        if (p < P && q < P)
            for (int r = 0; r < f; ++r) s = fma(Y[(size_t)r * ldy + p], Y[(size_t)r * ldy + q], s);
        G[idx] = s;
    }
    for (int p = tid; p < Pmax; p += nth) {
        double s = 0.0;
        if (p < P)
            for (int r = 0; r < f; ++r) s = fma(Y[(size_t)r * ldy + p], Y[(size_t)r * ldy + P], s);
        hall[(size_t)blockIdx.x * Pmax + p] = s;
    }
    if (tid == 0) {
        double q0 = 0.0, ld = 0.0;
        for (int r = 0; r < f; ++r) {
            q0 = fma(Y[(size_t)r * ldy + P], Y[(size_t)r * ldy + P], q0);
            ld += log(A[(size_t)r * f + r]);
        }
        q0all[blockIdx.x] = q0;
        ldall[blockIdx.x] = 2.0 * ld;
    }
}

// grid (n_blocks, W), 64 threads: M (Pmax x Pmax, identity padded), u (Pmax), quadA for one (walker, block).
__global__ void __launch_bounds__(64) woodbury_build_kernel(const WbBlock* __restrict__ blocks, int nB, int Pmax, int Ptot,
                                                            const double* __restrict__ m, const double* __restrict__ v,
                                                            const double* __restrict__ Gall, const double* __restrict__ hall,
                                                            const double* __restrict__ q0all, double* __restrict__ Mout,
                                                            double* __restrict__ uout, double* __restrict__ quadA) {
    __shared__ double sm[64], ssv[64], sred[64];
    const WbBlock B = blocks[blockIdx.x];
    const long long w = blockIdx.y;
    const int tid = threadIdx.x, P = B.P;
    const double* G = Gall + B.g_off;
    const long long slot = w * nB + blockIdx.x;
    if (tid < Pmax) {
        const bool ok = tid < P;
        sm[tid] = ok ? m[w * Ptot + B.col0 + tid] : 0.0;
        ssv[tid] = ok ? sqrt(v[w * Ptot + B.col0 + tid]) : 0.0;
    }
    __syncthreads();
    double part = 0.0;
    if (tid < Pmax) {
        double s = hall[(size_t)blockIdx.x * Pmax + tid];
        const double hh = s;
        for (int q = 0; q < P; ++q) s = fma(G[(size_t)q * Pmax + tid], sm[q], s);   // G symmetric: coalesced column read
        uout[slot * Pmax + tid] = ssv[tid] * s;
        part = sm[tid] * (hh + s);
    }
    sred[tid] = part;
    __syncthreads();
    if (tid == 0) {
        double q = q0all[blockIdx.x];
        for (int p = 0; p < Pmax; ++p) q += sred[p];
        quadA[slot] = q;
    }
    double* Mo = Mout + slot * (long long)Pmax * Pmax;
    for (int idx = tid; idx < Pmax * Pmax; idx += 64) {
        const int r = idx / Pmax, c = idx - r * Pmax;
        Mo[idx] = (r == c ? 1.0 : 0.0) + ssv[r] * G[idx] * ssv[c];
    }
}

// logp[w] = sum_b ( -1/2 (quadA - quadM) - 1/2 logdetA_b - logdetMhalf ), -inf outside the prior box
__global__ void __launch_bounds__(128) woodbury_combine_kernel(int nB, long long W, int d, const double* __restrict__ theta,
                                                               const double* __restrict__ lo, const double* __restrict__ hi,
                                                               const double* __restrict__ quadA, const double* __restrict__ quadM,
                                                               const double* __restrict__ ldMhalf, const double* __restrict__ ldA,
                                                               const int* __restrict__ info_in, double* __restrict__ logp,
                                                               int* __restrict__ info) {
    const long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= W) return;
    bool inside = true;
    for (int k = 0; k < d; ++k) {
        const double x = theta[w * d + k];
        if (!(x > lo[k] && x < hi[k])) inside = false;
    }
    if (!inside) {
        logp[w] = -INFINITY;
        if (info) info[w] = -1;
        return;
    }
    double s = 0.0;
    int bad = 0;
    for (int b = 0; b < nB; ++b) {
        const long long slot = w * nB + b;
        s += -0.5 * (quadA[slot] - quadM[slot]) - 0.5 * ldA[b] - ldMhalf[slot];
        if (info_in[slot] != 0 && bad == 0) bad = b + 1;
    }
    logp[w] = s;
    if (info) info[w] = bad;
}

}  // namespace bi
