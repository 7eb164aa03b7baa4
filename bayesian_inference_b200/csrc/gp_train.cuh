// gp_train.cuh -- batched GP log-marginal-likelihood and its gradient: the objective of the reference's emulator
// training.
//
// The reference fits one GaussianProcessRegressor per principal component with 50 optimiser restarts
// (emulation.py:128-172); every objective evaluation is scikit-learn's log_marginal_likelihood(theta, eval_gradient=True)
// (_gpr.py:541-639): K = k(X, X; theta) + alpha I, L = chol(K), a = K^-1 y,
//     lml  = -1/2 y.a - sum log L_ii - N/2 log 2 pi,        grad_m = 1/2 tr((a a^T - K^-1) dK/dtheta_m).
// All restarts of all principal components are independent, so they are evaluated as one batch: one CTA per
// hyper-parameter vector, persistent over the batch, its N x N working set in an L2-resident scratch slot.
//
// One code path does the factorisation, the inverse factor and K^-1: a left-looking blocked Cholesky of the augmented
// matrix [[K, I], [I, 0]] with 32 x 32 tiles.  Its lower-left block is M = L^-T (upper block triangular, so its tiles
// (i', k) exist for i' <= k and their sweeps start at k = i'), its Schur complement is -M M^T = -K^-1.  Every tile
// update is an "NT" product  acc -= A_tile * B_tile^T  on the FP64 tensor pipe (mma.sync m8n8k4, SASS DMMA), operands
// read straight from the row-major tiles (32-byte sectors, fully used).  Per panel j there are exactly nt independent
// units (nt - j tiles of L, j tiles of M), one per warp; the diagonal tile is factorised and inverted by warp 0
// (lane = row, shuffles), and the panel solve is one more NT product with the explicit 32 x 32 inverse.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "common.cuh"

namespace bi {

constexpr int TRAIN_MAX_D = 16;       // length scales per kernel (parameters of the model)
constexpr int TRAIN_MAX_NT = 31;      // N <= 992
constexpr int TRAIN_LDY = 36;         // row stride of the diagonal-tile inverse in shared memory (bank-conflict free)

struct LmlParams {
    const double* X;        // (N, d) design, row-major
    const double* Y;        // (n_y, N) training targets (PCA scores), row-major
    const int* y_index;     // (B) row of Y per problem, or nullptr: problem b uses row b
    const double* theta;    // (B, nh) log hyper-parameters in scikit-learn order: log ell[0..d-1], [log c], [log eta]
    long long B;
    int N, d, nh, kind, has_const, has_white;
    double alpha;
    double* scratch;        // gridDim.x slots of scratch_stride doubles
    long long scratch_stride;
    double* lml;            // (B)
    double* grad;           // (B, nh); unused when L_out is set
    double* L_out;          // optional (B, N, N): chol(K + alpha I), lower, zeros above  (scikit-learn's L_)
    double* alpha_out;      // optional (B, N): K^-1 y                                      (scikit-learn's alpha_)
};

__host__ __device__ __forceinline__ int tri_idx(int i, int j) { return i * (i + 1) / 2 + j; }

__host__ inline size_t lml_scratch_doubles(int N) {
    const int nt = (N + 31) / 32;
    return (size_t)nt * (nt + 1) * 1024;      // lower tiles of L (later K^-1) + tiles of M = L^-T
}
__host__ inline size_t lml_smem_bytes(int N, int d, int nwarps) {
    const int nt = (N + 31) / 32, Np = nt * 32;
    size_t dbl = (size_t)((N * d + 1) & ~1) + 3 * (size_t)Np + 32 * TRAIN_LDY + (size_t)nwarps * (TRAIN_MAX_D + 4) + 2 * TRAIN_MAX_D + 8;
    return dbl * 8 + (size_t)nt * (nt + 1) / 2 * 2 * sizeof(int) + 16;
}

// acc (+/-)= A * B^T for 32 x 32 row-major tiles (A: leading dimension 32; B: leading dimension LDB).
template <bool NEG, int LDB>
__device__ __forceinline__ void tile_nt(double (&acc)[4][4][2], const double* A, const double* B, int lane) {
    const double* pa = A + (lane >> 2) * 32 + (lane & 3);
    const double* pb = B + (lane >> 2) * LDB + (lane & 3);
    double a[2][4], b[2][4];
#pragma unroll
    for (int f = 0; f < 4; ++f) {
        a[0][f] = pa[f * 256];
        b[0][f] = pb[f * 8 * LDB];
    }
#pragma unroll
    for (int kk = 0; kk < 8; ++kk) {
        const int cur = kk & 1, nxt = cur ^ 1;
        if (kk < 7) {
#pragma unroll
            for (int f = 0; f < 4; ++f) {
                a[nxt][f] = pa[f * 256 + (kk + 1) * 4];
                b[nxt][f] = pb[f * 8 * LDB + (kk + 1) * 4];
            }
        }
#pragma unroll
        for (int mf = 0; mf < 4; ++mf) {
            const double av = NEG ? -a[cur][mf] : a[cur][mf];
#pragma unroll
            for (int nf = 0; nf < 4; ++nf) dmma(acc[mf][nf], av, b[cur][nf]);
        }
    }
}

__device__ __forceinline__ void acc_zero(double (&acc)[4][4][2]) {
#pragma unroll
    for (int mf = 0; mf < 4; ++mf)
#pragma unroll
        for (int nf = 0; nf < 4; ++nf) acc[mf][nf][0] = acc[mf][nf][1] = 0.0;
}
__device__ __forceinline__ void acc_load(double (&acc)[4][4][2], const double* T, int lane) {
    const double* p = T + (lane >> 2) * 32 + 2 * (lane & 3);
#pragma unroll
    for (int mf = 0; mf < 4; ++mf)
#pragma unroll
        for (int nf = 0; nf < 4; ++nf) {
            const double2 v = *reinterpret_cast<const double2*>(p + mf * 256 + nf * 8);
            acc[mf][nf][0] = v.x;
            acc[mf][nf][1] = v.y;
        }
}
__device__ __forceinline__ void acc_store(const double (&acc)[4][4][2], double* T, int lane) {
    double* p = T + (lane >> 2) * 32 + 2 * (lane & 3);
#pragma unroll
    for (int mf = 0; mf < 4; ++mf)
#pragma unroll
        for (int nf = 0; nf < 4; ++nf)
            *reinterpret_cast<double2*>(p + mf * 256 + nf * 8) = make_double2(acc[mf][nf][0], acc[mf][nf][1]);
}

// Warp 0: factorise the 32 x 32 diagonal tile T in place (lower, zeros above), write Y = L_jj^-1 row-major into sY
// (stride TRAIN_LDY) and M_jj = Y^T into Mjj.  Returns sum log L_kk over the tile's rows; ok = 0 on a non-positive pivot.
__device__ __forceinline__ double warp_potrf_inv(double* T, double* sYinv, double* Mjj, int lane, int& ok) {
    double a[32];
    {
        const double4* row = reinterpret_cast<const double4*>(T + lane * 32);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const double4 v = row[q];
            a[4 * q] = v.x; a[4 * q + 1] = v.y; a[4 * q + 2] = v.z; a[4 * q + 3] = v.w;
        }
    }
    double mypiv = 1.0, myrinv = 1.0;
#pragma unroll
    for (int k = 0; k < 32; ++k) {
        const double piv = shfl_d(a[k], k);
        if (!(piv > 0.0)) ok = 0;
        const double lkk = sqrt(piv);
        const double rinv = 1.0 / lkk;
        if (lane == k) { mypiv = piv; myrinv = rinv; }
        a[k] = (lane == k) ? lkk : a[k] * rinv;
#pragma unroll
        for (int c = k + 1; c < 32; ++c) {
            const double lc = shfl_d(a[k], c);
            a[c] = fma(-a[k], lc, a[c]);
        }
    }
#pragma unroll
    for (int c = 0; c < 32; ++c) {
        if (c > lane) a[c] = 0.0;
        T[lane * 32 + c] = a[c];
    }
    // forward substitution on the identity, lane = column of Y
#pragma unroll
    for (int r = 0; r < 32; ++r) {
        double s = (r == lane) ? 1.0 : 0.0;
#pragma unroll
        for (int k = 0; k < r; ++k) s = fma(-shfl_d(a[k], r), sYinv[k * TRAIN_LDY + lane], s);
        const double yv = s * shfl_d(myrinv, r);
        sYinv[r * TRAIN_LDY + lane] = yv;
        Mjj[lane * 32 + r] = yv;                 // M_jj[c][r] = Y[r][c]
    }
    return warp_sum(0.5 * log(mypiv));
}

// d(kernel)/d(log ell_m) = gfun(sq) * D_m with D_m the squared scaled difference in dimension m (scikit-learn
// kernels.py: Matern.__call__ / RBF.__call__ with eval_gradient=True); kval = the kernel value itself.
__device__ __forceinline__ void kernel_and_gfun(int kind, double sq, double& kval, double& gfun) {
    if (kind == 3) {
        kval = exp(-0.5 * sq);
        gfun = kval;
        return;
    }
    const double r = sqrt(sq);
    if (kind == 0) {
        kval = exp(-r);
        gfun = (r > 0.0) ? kval / r : 0.0;
    } else if (kind == 1) {
        const double t = r * 1.7320508075688772;
        const double e = exp(-t);
        kval = (1.0 + t) * e;
        gfun = 3.0 * e;
    } else {
        const double t = r * 2.23606797749979;
        const double e = exp(-t);
        kval = (1.0 + t + t * t / 3.0) * e;
        gfun = (5.0 / 3.0) * (t + 1.0) * e;
    }
}

__global__ void __launch_bounds__(512, 1) gp_lml_kernel(const LmlParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int N = P.N, d = P.d, nt = (N + 31) / 32, Np = nt * 32, ntri = nt * (nt + 1) / 2;
    const int tid = threadIdx.x, nth = blockDim.x, lane = tid & 31, warp = tid >> 5, NW = nth >> 5;
    double* sXs = reinterpret_cast<double*>(smem_raw);            // [N][d] design / length scale
    double* sY = sXs + ((N * d + 1) & ~1);                        // [Np] targets, zero padded
    double* sZ = sY + Np;                                         // [Np] L^-1 y
    double* sAl = sZ + Np;                                        // [Np] K^-1 y
    double* sYinv = sAl + Np;                                     // [32][TRAIN_LDY] inverse of the diagonal tile
    double* sRed = sYinv + 32 * TRAIN_LDY;                        // [NW][TRAIN_MAX_D + 4]
    double* sHyp = sRed + NW * (TRAIN_MAX_D + 4);                 // [2 * TRAIN_MAX_D + 8]: 1/ell, c, eta, logdet
    int* sTi = reinterpret_cast<int*>(sHyp + 2 * TRAIN_MAX_D + 8);
    int* sTj = sTi + ntri;
    int* sOk = sTj + ntri;

    double* Lt = P.scratch + (long long)blockIdx.x * P.scratch_stride;      // lower tiles (i, j) at tri_idx(i, j)
    double* Mt = Lt + (size_t)ntri * 1024;                                  // M tile (i', k) at tri_idx(k, i')

    for (int i = tid; i < nt; i += nth)
        for (int j = 0; j <= i; ++j) {
            sTi[tri_idx(i, j)] = i;
            sTj[tri_idx(i, j)] = j;
        }

    for (long long b = blockIdx.x; b < P.B; b += gridDim.x) {
        __syncthreads();
        const double* th = P.theta + b * P.nh;
        if (tid < d) sHyp[tid] = exp(th[tid]);
        if (tid == 0) {
            sHyp[TRAIN_MAX_D] = P.has_const ? exp(th[d]) : 0.0;
            sHyp[TRAIN_MAX_D + 1] = P.has_white ? exp(th[d + (P.has_const ? 1 : 0)]) : 0.0;
            *sOk = 1;
        }
        __syncthreads();
        const double cst = sHyp[TRAIN_MAX_D], eta = sHyp[TRAIN_MAX_D + 1];
        {
            const long long yrow = P.y_index ? P.y_index[b] : b;
            for (int idx = tid; idx < N * d; idx += nth) sXs[idx] = P.X[idx] / sHyp[idx % d];     // X / length_scale
            for (int idx = tid; idx < Np; idx += nth) sY[idx] = idx < N ? P.Y[yrow * N + idx] : 0.0;
        }
        __syncthreads();

        // ---- K + alpha I as lower tiles (identity padding past N)
        for (int e = tid; e < ntri * 1024; e += nth) {
            const int tI = e >> 10, p = sTi[tI] * 32 + ((e >> 5) & 31), q = sTj[tI] * 32 + (e & 31);
            double val = (p == q) ? 1.0 : 0.0;
            if (p < N && q < N) {
                double sq = 0.0;
                for (int m = 0; m < d; ++m) {
                    const double df = sXs[p * d + m] - sXs[q * d + m];
                    sq = fma(df, df, sq);
                }
                double kv, gf;
                kernel_and_gfun(P.kind, sq, kv, gf);
                val = kv + cst;
                if (p == q) val += eta + P.alpha;
            }
            Lt[e] = val;
        }
        __syncthreads();

        // ---- left-looking blocked Cholesky of [[K, I], [I, 0]]: nt units per panel
        double logdet = 0.0;
        for (int j = 0; j < nt; ++j) {
            const int nreg = nt - j;          // units 0 .. nreg-1: tiles (j + u, j) of L; nreg .. nt-1: M tiles (u - nreg, j)
            for (int u = warp; u < nt; u += NW) {
                double acc[4][4][2];
                double* T;
                if (u < nreg) {
                    const int i = j + u;
                    T = Lt + (size_t)tri_idx(i, j) * 1024;
                    acc_load(acc, T, lane);
                    for (int k = 0; k < j; ++k)
                        tile_nt<true, 32>(acc, Lt + (size_t)tri_idx(i, k) * 1024, Lt + (size_t)tri_idx(j, k) * 1024, lane);
                } else {
                    const int ip = u - nreg;
                    T = Mt + (size_t)tri_idx(j, ip) * 1024;
                    acc_zero(acc);
                    for (int k = ip; k < j; ++k)
                        tile_nt<true, 32>(acc, Mt + (size_t)tri_idx(k, ip) * 1024, Lt + (size_t)tri_idx(j, k) * 1024, lane);
                }
                acc_store(acc, T, lane);
                if (u == 0) {                 // warp 0 owns the diagonal tile: factorise + invert it right away
                    __syncwarp();
                    int ok = 1;
                    logdet += warp_potrf_inv(T, sYinv, Mt + (size_t)tri_idx(j, j) * 1024, lane, ok);
                    if (!ok && lane == 0) *sOk = 0;
                }
            }
            __syncthreads();
            for (int u = warp; u < nt; u += NW) {
                if (u == 0) continue;
                double* T = (u < nreg) ? Lt + (size_t)tri_idx(j + u, j) * 1024 : Mt + (size_t)tri_idx(j, u - nreg) * 1024;
                double acc[4][4][2];
                acc_zero(acc);
                tile_nt<false, TRAIN_LDY>(acc, T, sYinv, lane);       // T * L_jj^-T
                __syncwarp();
                acc_store(acc, T, lane);
            }
            __syncthreads();
        }

        // ---- z = L^-1 y = M^T y   (thread = element of z)
        for (int idx = tid; idx < Np; idx += nth) {
            const int k = idx >> 5, c = idx & 31;
            double s = 0.0;
            for (int ip = 0; ip <= k; ++ip) {
                const double* M = Mt + (size_t)tri_idx(k, ip) * 1024 + c;
                const double* y = sY + ip * 32;
#pragma unroll 8
                for (int r = 0; r < 32; ++r) s = fma(M[r * 32], y[r], s);
            }
            sZ[idx] = s;
        }
        __syncthreads();
        // ---- a = L^-T z = M z      (warp = row)
        for (int row = warp; row < Np; row += NW) {
            const int ip = row >> 5, r = row & 31;
            double s = 0.0;
            for (int k = ip; k < nt; ++k) s = fma(Mt[(size_t)tri_idx(k, ip) * 1024 + r * 32 + lane], sZ[k * 32 + lane], s);
            s = warp_sum(s);
            if (lane == 0) sAl[row] = s;
        }
        if (P.L_out != nullptr) {       // the fitted state scikit-learn keeps (_gpr.py:335-366): L_, alpha_ and the LML
            __syncthreads();
            double* Lo = P.L_out + (size_t)b * N * N;
            for (int e = tid; e < N * N; e += nth) {
                const int p = e / N, q = e - p * N;
                Lo[e] = (q <= p) ? Lt[(size_t)tri_idx(p >> 5, q >> 5) * 1024 + (p & 31) * 32 + (q & 31)] : 0.0;
            }
            double qd = 0.0;
            for (int idx = tid; idx < Np; idx += nth) {
                if (idx < N) P.alpha_out[(size_t)b * N + idx] = sAl[idx];
                qd = fma(sZ[idx], sZ[idx], qd);
            }
            qd = warp_sum(qd);
            if (lane == 0) sRed[warp] = qd;
            __syncthreads();
            if (tid == 0) {
                double s = 0.0;
                for (int w = 0; w < NW; ++w) s += sRed[w];
                P.lml[b] = *sOk ? (-0.5 * s - logdet - 0.5 * N * 1.8378770664093453) : -INFINITY;
            }
            continue;
        }
        // ---- K^-1 = M M^T, lower tiles, written over L
        for (int tI = warp; tI < ntri; tI += NW) {
            const int a = sTi[tI], bb = sTj[tI];
            double acc[4][4][2];
            acc_zero(acc);
            for (int k = a; k < nt; ++k)
                tile_nt<false, 32>(acc, Mt + (size_t)tri_idx(k, a) * 1024, Mt + (size_t)tri_idx(k, bb) * 1024, lane);
            acc_store(acc, Lt + (size_t)tI * 1024, lane);
        }
        __syncthreads();

        // ---- gradient: 1/2 sum_pq (a_p a_q - Kinv_pq) dK_pq/dtheta_m over the lower triangle (off-diagonal counted twice)
        double g[TRAIN_MAX_D];
#pragma unroll
        for (int m = 0; m < TRAIN_MAX_D; ++m) g[m] = 0.0;
        double gc = 0.0, gw = 0.0, qd = 0.0;
        for (int e = tid; e < ntri * 1024; e += nth) {
            const int tI = e >> 10, p = sTi[tI] * 32 + ((e >> 5) & 31), q = sTj[tI] * 32 + (e & 31);
            if (p < N && q <= p) {
                const double wv = fma(sAl[p], sAl[q], -Lt[e]);
                if (p == q) {
                    gw += wv;
                    gc += wv;            // D = 0 on the diagonal: no length-scale term
                } else {
                    double D[TRAIN_MAX_D];
                    double sq = 0.0;
#pragma unroll
                    for (int m = 0; m < TRAIN_MAX_D; ++m) {
                        if (m < d) {
                            const double df = sXs[p * d + m] - sXs[q * d + m];
                            D[m] = df * df;
                            sq += D[m];
                        }
                    }
                    double kv, gf;
                    kernel_and_gfun(P.kind, sq, kv, gf);
                    const double w2 = 2.0 * wv;
                    gc += w2;
                    const double wg = w2 * gf;
#pragma unroll
                    for (int m = 0; m < TRAIN_MAX_D; ++m)
                        if (m < d) g[m] = fma(wg, D[m], g[m]);
                }
            }
        }
        for (int idx = tid; idx < Np; idx += nth) qd = fma(sZ[idx], sZ[idx], qd);
#pragma unroll
        for (int m = 0; m < TRAIN_MAX_D; ++m) {
            const double s = warp_sum(g[m]);
            if (lane == 0) sRed[warp * (TRAIN_MAX_D + 4) + m] = s;
        }
        gc = warp_sum(gc);
        gw = warp_sum(gw);
        qd = warp_sum(qd);
        if (lane == 0) {
            sRed[warp * (TRAIN_MAX_D + 4) + TRAIN_MAX_D] = gc;
            sRed[warp * (TRAIN_MAX_D + 4) + TRAIN_MAX_D + 1] = gw;
            sRed[warp * (TRAIN_MAX_D + 4) + TRAIN_MAX_D + 2] = qd;
        }
        __syncthreads();
        if (warp == 0) {
            const int ok = *sOk;
            if (lane < TRAIN_MAX_D + 3) {
                double s = 0.0;
                for (int w = 0; w < NW; ++w) s += sRed[w * (TRAIN_MAX_D + 4) + lane];
                if (lane < d) P.grad[b * P.nh + lane] = ok ? 0.5 * s : 0.0;
                int pos = d;
                if (P.has_const) {
                    if (lane == TRAIN_MAX_D) P.grad[b * P.nh + pos] = ok ? 0.5 * s * cst : 0.0;
                    ++pos;
                }
                if (P.has_white && lane == TRAIN_MAX_D + 1) P.grad[b * P.nh + pos] = ok ? 0.5 * s * eta : 0.0;
                if (lane == TRAIN_MAX_D + 2)
                    P.lml[b] = ok ? (-0.5 * s - logdet - 0.5 * N * 1.8378770664093453) : -INFINITY;     // log(2 pi)
            }
        }
    }
}

}  // namespace bi
