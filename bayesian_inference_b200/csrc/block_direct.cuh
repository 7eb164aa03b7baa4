// block_direct.cuh -- K5 (direct): the log-likelihood observable block by observable block, one warp per (walker, block).
//
// The reference merges the emulation groups into a covariance that is block diagonal per observable
// (emulation.py:362-388) and then factorises the whole F x F matrix with dpotrf (log_posterior.py:123-146).
// When Sigma_exp does not couple observables the same value is the sum over the blocks
//     logp = sum_b [ -1/2 dy_b^T C_b^-1 dy_b - sum log diag chol(C_b) ],   C_b = A_b + T_b D T_b^T,  dy_b = e_b + T_b m
// (structured variant: "if f_b <= P: direct f_b x f_b Cholesky").  With f_b <= 31 a block lives in ONE
// warp's registers, one matrix row per lane:
//   * assembly  a[r][c] = A_b[r][c] + sum_p (T[r][p] v_p) T[c][p]  -- p outer, c inner; T_b is staged transposed in shared
//     memory once per CTA (a CTA works on one block for a batch of walkers), so T[c][p], T[c+1][p] come as one broadcast
//     16-byte load per two FMAs;
//   * lane 31 carries the augmented row dy_b (the same instruction stream with m in place of T[r][.] v and e_b in place of
//     A_b[r][.]), so the right-looking elimination turns it into z = L^-1 dy and the quadratic form is sum z^2: no separate
//     substitution (the device K4 kernels use the same augmentation);
//   * right-looking Cholesky in registers: pivot and multipliers by warp shuffles, rows between f and the padded size are
//     identity rows.
// The per-block terms go to a (W, n_blocks) scratch and block_combine_kernel adds them in block order (deterministic) and
// applies the strict prior box (log_posterior.py:63-64).
//
// Algorithmic work: sum_b f_b^3/3 + f_b^2 P  -- 0.44 MFLOP per walker at C3a (30 blocks of 20 bins, P = 30).
#pragma once
#include "common.cuh"

namespace bi {

struct BdBlock {
    int row0;         // first merged row of the block
    int f;            // rows (<= 31)
    int col0;         // first column of T_merged of the block's group
    int P;            // n_pc of the group (<= BD_PMAX)
    long long a_off;  // offset (doubles) of this block's f x f matrix A_b in the per-divisor array
};

constexpr int BD_WARPS = 8;
constexpr int BD_THREADS = BD_WARPS * 32;
constexpr int BD_PMAX = 64;    // principal components per group the staging buffers hold

// A_b = Sigma_exp[b, b] + U[b, b] / n_div for every block (theta independent; once per divisor)
__global__ void __launch_bounds__(256) block_a_kernel(const BdBlock* __restrict__ blocks, int F, const double* __restrict__ Sig,
                                                      const double* __restrict__ yerr, const double* __restrict__ Ublk, double n_div,
                                                      double* __restrict__ Aout) {
    const BdBlock B = blocks[blockIdx.x];
    double* A = Aout + B.a_off;
    for (int idx = threadIdx.x; idx < B.f * B.f; idx += blockDim.x) {
        const int r = idx / B.f, c = idx - r * B.f;
        const int gr = B.row0 + r, gc = B.row0 + c;
        const double sg = Sig ? Sig[(size_t)gr * F + gc] : (r == c ? yerr[gr] * yerr[gr] : 0.0);
        A[idx] = sg + Ublk[(size_t)gr * F + gc] / n_div;
    }
}

__host__ __device__ inline size_t block_direct_smem_bytes(int FB) {
    // sTt (BD_PMAX x FB) + sA (32 x FB: rows f..30 identity, row 31 = e_b) + per warp v and m (2 x BD_PMAX)
    return sizeof(double) * ((size_t)BD_PMAX * FB + (size_t)32 * FB + (size_t)BD_WARPS * 2 * BD_PMAX);
}

// grid (n_blocks, ceil(W / (BD_WARPS * wpw))): warp u of CTA (b, y) takes walkers y * BD_WARPS * wpw + u + BD_WARPS * i, i < wpw.
// FB = f padded to a multiple of 8 (template: the matrix row lives in FB registers).
// Resident CTAs per SM by row length: the kernel is bound by the latency of its dependent shuffle / FMA chains (ncu: FP64
// pipe 32 %, issue slots 44 % busy at 16 warps per SM), so the shorter rows trade registers for warps.
template <int FB>
__global__ void __launch_bounds__(BD_THREADS, FB <= 16 ? 4 : (FB <= 24 ? 3 : 2)) block_direct_kernel(const BdBlock* __restrict__ blocks, int nB, long long W, int wpw,
                                                                     int Ptot, int K0, const double* __restrict__ m,
                                                                     const double* __restrict__ v, const double* __restrict__ Trow,
                                                                     const double* __restrict__ e, const double* __restrict__ Aall,
                                                                     double* __restrict__ part, int* __restrict__ bad_out,
                                                                     const int* __restrict__ set_id, long long e_stride,
                                                                     long long a_stride) {
    extern __shared__ __align__(16) unsigned char bd_smem[];
    double* sTt = reinterpret_cast<double*>(bd_smem);        // [p][c]  T_b transposed, columns c >= f are zero
    double* sA = sTt + BD_PMAX * FB;                          // [r][c]  A_b, identity rows f..30, row 31 = e_b
    double* sVM = sA + 32 * FB;                               // [warp][2][BD_PMAX]
    const BdBlock B = blocks[blockIdx.x];
    const int f = B.f, P = B.P;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NCOL = (FB == 32) ? 31 : FB;                // columns eliminated (column 31 would be the augmented row's own)

    for (int idx = tid; idx < BD_PMAX * FB; idx += BD_THREADS) {
        const int p = idx / FB, c = idx - p * FB;
        sTt[idx] = (p < P && c < f) ? __ldg(Trow + (size_t)(B.row0 + c) * K0 + B.col0 + p) : 0.0;
    }
    for (int idx = tid; idx < 32 * FB; idx += BD_THREADS) {
        const int r = idx / FB, c = idx - r * FB;
        double x;
        if (r == 31) x = (c < f) ? __ldg(e + B.row0 + c) : 0.0;
        else if (r < f && c < f) x = __ldg(Aall + B.a_off + (size_t)r * f + c);
        else x = (r == c) ? 1.0 : 0.0;
        sA[idx] = x;
    }
    __syncthreads();

    double* sV = sVM + warp * 2 * BD_PMAX;
    double* sM = sV + BD_PMAX;
    const long long w_base = (long long)blockIdx.y * BD_WARPS * wpw + warp;
    for (int it = 0; it < wpw; ++it) {
        const long long w = w_base + (long long)it * BD_WARPS;
        if (w >= W) break;                                    // warp-uniform
        __syncwarp();
        for (int p = lane; p < BD_PMAX; p += 32) {
            const bool ok = p < P;
            sV[p] = ok ? __ldg(v + w * Ptot + B.col0 + p) : 0.0;
            sM[p] = ok ? __ldg(m + w * Ptot + B.col0 + p) : 0.0;
        }
        __syncwarp();
        // ---- assembly: row `lane` of [C_b ; dy_b]
        double a[FB];
        const bool is_aug = lane == 31;
        if (set_id == nullptr) {
            const double2* init = reinterpret_cast<const double2*>(sA + lane * FB);
#pragma unroll
            for (int c = 0; c < FB; c += 2) {
                const double2 x = init[c >> 1];
                a[c] = x.x;
                a[c + 1] = x.y;
            }
        } else {
            // several experimental data sets on one image (closure tests, steer_analysis.py:168-183): A_b and e_b of this
            // walker's set come straight from global memory (L2 resident), everything after that is the same arithmetic
            const long long sid = set_id[w];
            const double* As = Aall + sid * a_stride + B.a_off + (size_t)(lane < f ? lane : 0) * f;
            const double* es = e + sid * e_stride + B.row0;
#pragma unroll
            for (int c = 0; c < FB; ++c) {
                double x = (lane == c) ? 1.0 : 0.0;
                if (c < f) {
                    if (is_aug) x = __ldg(es + c);
                    else if (lane < f) x = __ldg(As + c);
                } else if (is_aug) x = 0.0;
                a[c] = x;
            }
        }
        const int trow = (lane < f) ? lane : 0;               // lanes f..30 (identity rows) use coefficient 0
        for (int p = 0; p < P; ++p) {
            double coef = sTt[p * FB + trow] * sV[p];
            if (lane >= f) coef = 0.0;
            if (is_aug) coef = sM[p];
            const double2* tp = reinterpret_cast<const double2*>(sTt + p * FB);
#pragma unroll
            for (int c = 0; c < FB; c += 2) {
                const double2 t2 = tp[c >> 1];
                a[c] = fma(coef, t2.x, a[c]);
                a[c + 1] = fma(coef, t2.y, a[c + 1]);
            }
        }
        // ---- right-looking Cholesky in registers; lane 31 becomes z = L^-1 dy
        int bad = 0;
        double mypiv = 1.0, quad = 0.0;
#pragma unroll
        for (int c = 0; c < NCOL; ++c) {
            const double piv = shfl_d(a[c], c);
            if (!(piv > 0.0) && bad == 0) bad = c + 1;
            if (lane == c) mypiv = piv;
            const double rinv = rsqrt(piv);
            a[c] *= rinv;
            if (is_aug) quad = fma(a[c], a[c], quad);
#pragma unroll
            for (int c2 = c + 1; c2 < NCOL; ++c2) {
                const double l = shfl_d(a[c], c2);
                a[c2] = fma(-a[c], l, a[c2]);
            }
        }
        const double ld = warp_sum((lane < f) ? 0.5 * log(mypiv) : 0.0);
        quad = shfl_d(quad, 31);
        if (lane == 0) {
            part[w * nB + blockIdx.x] = bad ? __longlong_as_double(0x7ff8000000000000LL) : (-0.5 * quad - ld);
            bad_out[w * nB + blockIdx.x] = bad ? B.row0 + bad : 0;
        }
    }
}

// logp[w] = sum_b part[w][b] in block order; -inf outside the strict prior box; info = first non-positive pivot (1-based row)
__global__ void __launch_bounds__(128) block_combine_kernel(int nB, long long W, int d, const double* __restrict__ theta,
                                                            const double* __restrict__ lo, const double* __restrict__ hi,
                                                            const double* __restrict__ part, const int* __restrict__ bad_in,
                                                            double* __restrict__ logp, int* __restrict__ info) {
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
        s += part[w * nB + b];
        if (bad_in[w * nB + b] != 0 && bad == 0) bad = bad_in[w * nB + b];
    }
    logp[w] = s;
    if (info) info[w] = bad;
}

}  // namespace bi
