// gp_train.cuh -- batched GP log-marginal-likelihood and its gradient: the objective of the reference's emulator
// training.
//
// The reference fits one GaussianProcessRegressor per principal component with 50 optimiser restarts
// (emulation.py:128-172); every objective evaluation is scikit-learn's log_marginal_likelihood(theta, eval_gradient=True)
// (_gpr.py:541-639): K = k(X, X; theta) + alpha I, L = chol(K), a = K^-1 y,
//     lml  = -1/2 y.a - sum log L_ii - N/2 log 2 pi,        grad_m = 1/2 tr((a a^T - K^-1) dK/dtheta_m).
// All restarts of all principal components are independent, so they are evaluated as one batch, in three launches:
//   gp_kbuild_kernel  element-wise: K + alpha I of every problem as 32 x 32 tiles (lower triangle)       [FP64 ALU]
//   gp_factor_kernel  one CTA per problem, persistent: Cholesky, L^-T, K^-1, a, the LML                   [FP64 tensor]
//   gp_grad_kernel    element-wise: the trace contraction with dK/dtheta recomputed on the fly            [FP64 ALU]
// The element-wise phases are separate launches on purpose: inside the factorisation kernel their dependent sqrt / exp
// chains queued behind the other warps' DMMAs on the shared FP64 pipe (~100 cycles per dependent instruction, measured)
// and took a third of its time; at full occupancy on their own they are throughput bound.
//
// One code path does the factorisation, the inverse factor and K^-1: a left-looking blocked Cholesky of the augmented
// matrix [[K, I], [I, 0]] with 32 x 32 tiles.  Its lower-left block is M = L^-T (upper block triangular, so its tiles
// (i', k) exist for i' <= k and their sweeps start at k = i'), its Schur complement is -M M^T = -K^-1.  Every tile
// update is an "NT" product  acc -= A_tile * B_tile^T  on the FP64 tensor pipe (mma.sync m8n8k4, SASS DMMA), operands
// read straight from the row-major tiles (32-byte sectors, fully used).  Per panel j there are exactly nt independent
// units (nt - j tiles of L, j tiles of M), one per warp; the diagonal tile is factorised by warp 0 with K4's blocked
// potrf (8-column panels, DMMA trailing updates, 8 x 8 inverses) and the panel solve is K4's blocked triangular solve on
// DMMA; M_jj = L_jj^-T is that same solve applied to the identity tile.
// The targets ride inside the matrix as row N (the augmented-row device of K4): the factor of [[K, y], [y^T, 0]] has
// z = L^-1 y as row N and -z.z as that row's would-be pivot (replaced by 1), so y.K^-1.y = z.z falls out of the last
// potrf, and column N of M = L_aug^-T is -L^-T z = -alpha: no separate forward / backward substitution.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "chol_pipe.cuh"       // warp_potrf_blocked / strip_trsm_blocked: the diagonal-tile factorisation and panel solve of K4
#include "common.cuh"
#include "gp_predict.cuh"      // exp_neg / sqrt_pos: branch-free exp(-x), sqrt(x) (<= 2 ulp)

namespace bi {

constexpr int TRAIN_MAX_D = 16;       // length scales per kernel (parameters of the model)
constexpr int TRAIN_MAX_NT = 31;      // N + 1 <= 992

struct LmlParams {
    const double* X;        // (N, d) design, row-major
    const double* Y;        // (n_y, N) training targets (PCA scores), row-major
    const int* y_index;     // (B) row of Y per problem, or nullptr: problem b uses row b
    const double* theta;    // (B, nh) log hyper-parameters in scikit-learn order: log ell[0..d-1], [log c], [log eta]
    long long B;
    int N, d, nh, has_const, has_white;
    double alpha;
    double* Kt;             // (B, ntri, 1024) lower tiles: K + alpha I in, L during the factorisation, K^-1 out
    double* Gt;             // (B, ntri, 1024) or nullptr: d(kernel)/d(log ell_m) / D_m per entry, written by the build for the
                            // gradient launch (which then needs no sqrt / exp); nullptr when no gradient is asked for
    double* Ab;             // (B, Np) K^-1 y
    double* scratch;        // gridDim.x slots of scratch_stride doubles: the tiles of M = L^-T
    long long scratch_stride;
    double* lml;            // (B)
    double* grad;           // (B, nh); not written when L_out is set
    double* L_out;          // optional (B, N, N): chol(K + alpha I), lower, zeros above  (scikit-learn's L_)
    double* alpha_out;      // optional (B, N): K^-1 y                                      (scikit-learn's alpha_)
};

#ifdef TRAIN_TIMING
__device__ unsigned long long train_dbg[12];   // developer timing: cycles of thread 0 per phase, summed over problems
#define TR_T(i) do { if (threadIdx.x == 0) { long long t1_ = clock64(); atomicAdd(&train_dbg[i], (unsigned long long)(t1_ - t0_)); t0_ = t1_; } } while (0)
#else
#define TR_T(i)
#endif

__host__ __device__ __forceinline__ int tri_idx(int i, int j) { return i * (i + 1) / 2 + j; }
// tiles per side: the N x N matrix plus one augmented row (the targets y, see gp_factor_kernel), padded to 32
__host__ __device__ __forceinline__ int train_nt(int N) { return (N + 1 + 31) / 32; }

__host__ inline size_t lml_tiles_doubles(int N) {        // per problem: lower tiles of K / L / K^-1
    const int nt = train_nt(N);
    return (size_t)nt * (nt + 1) / 2 * 1024;
}
__host__ inline size_t lml_factor_smem_bytes(int N, int nwarps) {
    const int nt = train_nt(N), Np = nt * 32;
    size_t dbl = (size_t)Np + 1024 + 256 + 32 + 32 * 33 + (size_t)nwarps + 8;
    return dbl * 8 + (size_t)nt * (nt + 1) / 2 * 2 * sizeof(int) + 16;
}
__host__ inline size_t lml_grad_smem_bytes(int N, int d) {
    const int nt = train_nt(N), Np = nt * 32;
    return ((size_t)((N * d + 1) & ~1) + Np + 32 * (TRAIN_MAX_D + 2) + TRAIN_MAX_D + 4) * 8 + (size_t)nt * (nt + 1) / 2 * 2 * sizeof(int) + 16;
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

__device__ __forceinline__ void prefetch_tile(const double* T, int lane) {      // 8 KB = 64 lines of 128 bytes
    asm volatile("prefetch.global.L1 [%0];" ::"l"(T + lane * 16));
    asm volatile("prefetch.global.L1 [%0];" ::"l"(T + 512 + lane * 16));
}

// d(kernel)/d(log ell_m) = gfun(sq) * D_m with D_m the squared scaled difference in dimension m (scikit-learn
// kernels.py: Matern.__call__ / RBF.__call__ with eval_gradient=True); kval = the kernel value itself.
template <int KIND>
__device__ __forceinline__ void kernel_and_gfun(double sq, double& kval, double& gfun) {
    if (KIND == 3) {
        kval = exp_neg(0.5 * sq);
        gfun = kval;
        return;
    }
    const double r = sqrt_pos(sq);            // 1e-150 instead of 0 at coincident points: their D_m are exactly 0
    if (KIND == 0) {
        kval = exp_neg(r);
        gfun = kval / r;
    } else if (KIND == 1) {
        const double t = r * 1.7320508075688772;
        const double e = exp_neg(t);
        kval = fma(t, e, e);
        gfun = 3.0 * e;
    } else {
        const double t = r * 2.23606797749979;
        const double e = exp_neg(t);
        kval = fma(t * t, 1.0 / 3.0, 1.0 + t) * e;
        gfun = (5.0 / 3.0) * (t + 1.0) * e;
    }
}

// ---- launch 1: K + alpha I as lower tiles (identity padding past N).  grid (ntri, B), 256 threads, 4 elements each.
template <int KIND>
__global__ void __launch_bounds__(256) gp_kbuild_kernel(const LmlParams P) {
    __shared__ double sInv[TRAIN_MAX_D];
    __shared__ double sXp[32 * TRAIN_MAX_D], sXq[32 * TRAIN_MAX_D];
    const int N = P.N, d = P.d, nt = train_nt(N), ntri = nt * (nt + 1) / 2;
    const int tI = blockIdx.x, tid = threadIdx.x;
    const long long b = blockIdx.y;
    int ti = (int)((sqrt(8.0 * tI + 1.0) - 1.0) * 0.5);
    while ((ti + 1) * (ti + 2) / 2 <= tI) ++ti;
    while (ti * (ti + 1) / 2 > tI) --ti;
    const int tj = tI - ti * (ti + 1) / 2;
    const double* th = P.theta + b * P.nh;
    if (tid < d) sInv[tid] = exp(th[tid]);
    __syncthreads();
    for (int idx = tid; idx < 32 * d; idx += 256) {          // X / length_scale for the tile's 32 rows and 32 columns
        const int r = idx / d, m = idx - r * d;
        sXp[idx] = P.X[min(ti * 32 + r, N - 1) * d + m] / sInv[m];
        sXq[m * 32 + r] = P.X[min(tj * 32 + r, N - 1) * d + m] / sInv[m];      // transposed: lanes read consecutive columns
    }
    const double cst = P.has_const ? exp(th[d]) : 0.0;
    const double eta = P.has_white ? exp(th[d + (P.has_const ? 1 : 0)]) : 0.0;
    __syncthreads();
    double* T = P.Kt + ((size_t)b * ntri + tI) * 1024;
    const double* y = P.Y + (size_t)(P.y_index ? P.y_index[b] : b) * N;
    double* G = P.Gt ? P.Gt + ((size_t)b * ntri + tI) * 1024 : nullptr;
    auto element = [&](int u, bool live) {
        const int e = tid + 256 * u, r = e >> 5, c = e & 31, p = ti * 32 + r, q = tj * 32 + c;
        double kv = 0.0, gf = 0.0;
        if (live) {
            double sq = 0.0;
            for (int m = 0; m < d; ++m) {
                const double df = sXp[r * d + m] - sXq[m * 32 + c];
                sq = fma(df, df, sq);
            }
            kernel_and_gfun<KIND>(sq, kv, gf);
        }
        const double dg = (p == q) ? 1.0 : 0.0;
        double val = (p < N && q < N) ? kv + cst + dg * (eta + P.alpha) : dg;
        if (p == N || q == N) {                       // the augmented row / column: y, with a zero on the diagonal
            const int o = (p == N) ? q : p;
            val = (o < N) ? y[o] : 0.0;
        }
        T[e] = val;
        if (G != nullptr) G[e] = gf;
    };
    if (ti < nt - 1) {                                // four independent sqrt / exp chains per thread
#pragma unroll
        for (int u = 0; u < 4; ++u) element(u, true);
    } else {                                          // last tile row: rows past N are padding (warp-uniform test)
#pragma unroll
        for (int u = 0; u < 4; ++u) element(u, ti * 32 + (tid >> 5) + 8 * u < N);
    }
}

// ---- launch 2: factorisation, inverse factor, K^-1, alpha, LML.  One CTA per problem, persistent.
__global__ void __launch_bounds__(512, 1) gp_factor_kernel(const LmlParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int N = P.N, nt = train_nt(N), Np = nt * 32, ntri = nt * (nt + 1) / 2;
    const int tid = threadIdx.x, nth = blockDim.x, lane = tid & 31, warp = tid >> 5, NW = nth >> 5;
    double* sAl = reinterpret_cast<double*>(smem_raw);            // [Np] K^-1 y
    double* sLjj = sAl + Np;                                      // [1024] L_jj as B-operand fragments (blocked panel solve)
    double* sLinvB = sLjj + 1024;                                 // [256]  inverses of its four 8 x 8 diagonal blocks
    double* sRinv = sLinvB + 256;                                 // [32]   1 / L_kk
    double* sL = sRinv + 32;                                      // [32][33] the diagonal tile, row-major
    double* sRed = sL + 32 * 33;                                  // [NW + 8]
    int* sTi = reinterpret_cast<int*>(sRed + NW + 8);
    int* sTj = sTi + ntri;
    int* sOk = sTj + ntri;

    double* Mt = P.scratch + (long long)blockIdx.x * P.scratch_stride;      // M tile (i', k) at tri_idx(k, i')

    for (int i = tid; i < nt; i += nth)
        for (int j = 0; j <= i; ++j) {
            sTi[tri_idx(i, j)] = i;
            sTj[tri_idx(i, j)] = j;
        }

    for (long long b = blockIdx.x; b < P.B; b += gridDim.x) {
        __syncthreads();
#ifdef TRAIN_TIMING
        long long t0_ = clock64();
#endif
        double* Lt = P.Kt + (size_t)b * ntri * 1024;              // lower tiles (i, j) at tri_idx(i, j)
        if (tid == 0) *sOk = 1;
        __syncthreads();
        TR_T(0);

        // ---- left-looking blocked Cholesky of [[K, I], [I, 0]]: nt units per panel.
        // The diagonal tile of panel j + 1 is swept one panel ahead (all k < j, by the warp whose own unit is the
        // shortest), so that warp 0 only adds k = j - 1 before it factorises: the serial part of a panel is one product
        // + potrf instead of j products + potrf.
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
                    for (int k = (u == 0 && j > 0) ? j - 1 : 0; k < j; ++k) {
                        if (k + 1 < j) prefetch_tile(Lt + (size_t)tri_idx(i, k + 1) * 1024, lane);
                        tile_nt<true, 32>(acc, Lt + (size_t)tri_idx(i, k) * 1024, Lt + (size_t)tri_idx(j, k) * 1024, lane);
                    }
                } else {
                    const int ip = u - nreg;
                    T = Mt + (size_t)tri_idx(j, ip) * 1024;
                    acc_zero(acc);
                    for (int k = ip; k < j; ++k) {
                        if (k + 1 < j) prefetch_tile(Mt + (size_t)tri_idx(k + 1, ip) * 1024, lane);
                        tile_nt<true, 32>(acc, Mt + (size_t)tri_idx(k, ip) * 1024, Lt + (size_t)tri_idx(j, k) * 1024, lane);
                    }
                }
                acc_store(acc, T, lane);
                if (u == 0) {                 // warp 0 owns the diagonal tile: blocked factorisation as in K4
                    __syncwarp();
                    TR_T(1);
#pragma unroll
                    for (int c = 0; c < 32; ++c) sL[lane * 33 + c] = T[lane * 32 + c];
                    __syncwarp();
                    double mypiv, quad = 0.0;
                    const int aug = (j == nt - 1) ? N - j * 32 : -1;        // row N: the targets
                    if (warp_potrf_blocked(sL, sLjj, sLinvB, sRinv, lane, mypiv, aug, quad) != 0 && lane == 0) *sOk = 0;
                    if (aug >= 0 && lane == 0) sRed[0] = quad;              // z.z = y K^-1 y
                    __syncwarp();
#pragma unroll
                    for (int c = 0; c < 32; ++c) T[lane * 32 + c] = (c <= lane) ? sL[lane * 33 + c] : 0.0;
                    logdet += warp_sum(0.5 * log(mypiv));
                    TR_T(2);
                }
            }
            if (warp == NW - 1 && j + 1 < nt && j > 0) {        // look-ahead: tile (j+1, j+1) -= sum_{k<j} L(j+1,k) L(j+1,k)^T
                double acc[4][4][2];
                double* T = Lt + (size_t)tri_idx(j + 1, j + 1) * 1024;
                acc_load(acc, T, lane);
                for (int k = 0; k < j; ++k) {
                    const double* A = Lt + (size_t)tri_idx(j + 1, k) * 1024;
                    if (k + 1 < j) prefetch_tile(A + 1024, lane);
                    tile_nt<true, 32>(acc, A, A, lane);
                }
                acc_store(acc, T, lane);
            }
            __syncthreads();
            TR_T(3);
            // panel solve X = T L_jj^-T, blocked (L_jj + its 8 x 8 inverses), two 16-row strips per tile; warp 0's unit
            // is the identity tile: M_jj = L_jj^-T
            for (int u = warp; u < nt; u += NW) {
                double* T = (u == 0) ? Mt + (size_t)tri_idx(j, j) * 1024
                          : (u < nreg) ? Lt + (size_t)tri_idx(j + u, j) * 1024 : Mt + (size_t)tri_idx(j, u - nreg) * 1024;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    double c[2][4][2], xa[2][8];
                    double* p = T + (16 * h + (lane >> 2)) * 32 + 2 * (lane & 3);
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                        for (int nf = 0; nf < 4; ++nf) {
                            if (u == 0) {
                                const int r = 16 * h + 8 * mi + (lane >> 2), cc = 8 * nf + 2 * (lane & 3);
                                c[mi][nf][0] = (r == cc) ? 1.0 : 0.0;
                                c[mi][nf][1] = (r == cc + 1) ? 1.0 : 0.0;
                            } else {
                                const double2 v = *reinterpret_cast<const double2*>(p + mi * 256 + nf * 8);
                                c[mi][nf][0] = v.x;
                                c[mi][nf][1] = v.y;
                            }
                        }
                    strip_trsm_blocked(c, xa, sLjj + lane, sLinvB + lane, lane);
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                        for (int nf = 0; nf < 4; ++nf)
                            *reinterpret_cast<double2*>(p + mi * 256 + nf * 8) = make_double2(c[mi][nf][0], c[mi][nf][1]);
                }
            }
            __syncthreads();
            TR_T(4);
        }

        // ---- alpha = -column N of M; then that column is cleared so that M M^T is K^-1 on the first N rows
        {
            const int ral = N - (nt - 1) * 32;
            for (int idx = tid; idx < Np; idx += nth) {
                double* m = Mt + (size_t)tri_idx(nt - 1, idx >> 5) * 1024 + (idx & 31) * 32 + ral;
                const double a = (idx < N) ? -*m : 0.0;
                sAl[idx] = a;
                P.Ab[(size_t)b * Np + idx] = a;
                if (idx != N) *m = 0.0;
            }
        }
        __syncthreads();
        TR_T(5);
        if (P.L_out != nullptr) {       // the fitted state scikit-learn keeps (_gpr.py:335-366): L_, alpha_ and the LML
            double* Lo = P.L_out + (size_t)b * N * N;
            for (int e = tid; e < N * N; e += nth) {
                const int p = e / N, q = e - p * N;
                Lo[e] = (q <= p) ? Lt[(size_t)tri_idx(p >> 5, q >> 5) * 1024 + (p & 31) * 32 + (q & 31)] : 0.0;
            }
            for (int idx = tid; idx < N; idx += nth) P.alpha_out[(size_t)b * N + idx] = sAl[idx];
        } else {
            // ---- K^-1 = M M^T, lower tiles, written over L
            for (int tI = warp; tI < ntri; tI += NW) {
                const int a = sTi[tI], bb = sTj[tI];
                double acc[4][4][2];
                acc_zero(acc);
                for (int k = a; k < nt; ++k) {
                    if (k + 1 < nt) {
                        prefetch_tile(Mt + (size_t)tri_idx(k + 1, a) * 1024, lane);
                        prefetch_tile(Mt + (size_t)tri_idx(k + 1, bb) * 1024, lane);
                    }
                    tile_nt<false, 32>(acc, Mt + (size_t)tri_idx(k, a) * 1024, Mt + (size_t)tri_idx(k, bb) * 1024, lane);
                }
                acc_store(acc, Lt + (size_t)tI * 1024, lane);
            }
            __syncthreads();
        }
        TR_T(7);
        if (tid == 0)
            P.lml[b] = *sOk ? (-0.5 * sRed[0] - logdet - 0.5 * N * 1.8378770664093453) : -INFINITY;     // log(2 pi)
    }
}

// ---- launch 3: grad_m = 1/2 sum_pq (a_p a_q - Kinv_pq) dK_pq/dtheta_m over the lower triangle (off-diagonal entries
// counted twice).  One CTA of 512 threads per problem, warp = row p, lanes = columns q <= p; deterministic reduction.
template <int KIND, int DMAX>
__global__ void __launch_bounds__(512, 2) gp_grad_kernel(const LmlParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int N = P.N, d = P.d, nt = train_nt(N), Np = nt * 32, ntri = nt * (nt + 1) / 2;
    const int tid = threadIdx.x, nth = blockDim.x, lane = tid & 31, warp = tid >> 5, NW = nth >> 5;
    const long long b = blockIdx.x;
    double* sXs = reinterpret_cast<double*>(smem_raw);            // [N][d] design / length scale
    double* sAl = sXs + ((N * d + 1) & ~1);                       // [Np]
    double* sRed = sAl + Np;                                      // [NW <= 32][TRAIN_MAX_D + 2]
    double* sHyp = sRed + 32 * (TRAIN_MAX_D + 2);                 // [TRAIN_MAX_D + 4]
    const double* th = P.theta + b * P.nh;
    if (tid < d) sHyp[tid] = exp(th[tid]);
    __syncthreads();
    for (int idx = tid; idx < N * d; idx += nth) sXs[idx] = P.X[idx] / sHyp[idx % d];
    for (int idx = tid; idx < Np; idx += nth) sAl[idx] = P.Ab[(size_t)b * Np + idx];
    __syncthreads();
    const double* Kinv = P.Kt + (size_t)b * ntri * 1024;
    double g[DMAX];          // DMAX = 8 or 16 >= d: accumulators stay in registers
#pragma unroll
    for (int m = 0; m < DMAX; ++m) g[m] = 0.0;
    double gc = 0.0, gw = 0.0;
    // rows are dealt in pairs (j, N-1-j): every pair has N + 1 columns, so the warps are balanced
    for (int it = 0;; ++it) {
        const int jp = warp + (it >> 1) * NW;
        if (jp >= ((N + 1) >> 1)) break;
        const int p = (it & 1) ? N - 1 - jp : jp;
        if ((it & 1) && p == jp) continue;             // the middle row of an odd N is its own partner
        const double ap = sAl[p];
        const double* xp = sXs + p * d;
        const double* Krow = Kinv + (size_t)tri_idx(p >> 5, 0) * 1024 + (p & 31) * 32;
        const double* Grow = P.Gt + ((size_t)b * ntri + tri_idx(p >> 5, 0)) * 1024 + (p & 31) * 32;
#pragma unroll 4
        for (int q0 = 0; q0 <= p; q0 += 32) {
            const int q = q0 + lane, qc = min(q, p);
            const double wv = fma(ap, sAl[qc], -Krow[(size_t)(q0 >> 5) * 1024 + (qc & 31)]);
            const double wd = (q == p) ? wv : 0.0;                      // diagonal: constant and white terms only
            const double w2 = (q < p) ? 2.0 * wv : 0.0;                 // strictly lower: counted twice
            gw += wd;
            gc += wd + w2;
            const double* xq = sXs + qc * d;
            const double wg = w2 * Grow[(size_t)(q0 >> 5) * 1024 + (qc & 31)];      // gfun(sq) from the build launch
#pragma unroll
            for (int m = 0; m < DMAX; ++m)
                if (m < d) {
                    const double df = xp[m] - xq[m];
                    g[m] = fma(wg * df, df, g[m]);
                }
        }
    }
#pragma unroll
    for (int m = 0; m < DMAX; ++m) {
        const double s = warp_sum(g[m]);
        if (lane == 0) sRed[warp * (TRAIN_MAX_D + 2) + m] = s;
    }
    gc = warp_sum(gc);
    gw = warp_sum(gw);
    if (lane == 0) {
        sRed[warp * (TRAIN_MAX_D + 2) + TRAIN_MAX_D] = gc;
        sRed[warp * (TRAIN_MAX_D + 2) + TRAIN_MAX_D + 1] = gw;
    }
    __syncthreads();
    if (warp == 0 && lane < TRAIN_MAX_D + 2) {
        const bool ok = P.lml[b] > -INFINITY;           // not positive definite: (-inf, zero gradient), _gpr.py:586-589
        double s = 0.0;
        for (int w = 0; w < NW; ++w) s += sRed[w * (TRAIN_MAX_D + 2) + lane];
        if (lane < d) P.grad[b * P.nh + lane] = ok ? 0.5 * s : 0.0;
        int pos = d;
        if (P.has_const) {
            if (lane == TRAIN_MAX_D) P.grad[b * P.nh + pos] = ok ? 0.5 * s * exp(th[pos]) : 0.0;
            ++pos;
        }
        if (P.has_white && lane == TRAIN_MAX_D + 1) P.grad[b * P.nh + pos] = ok ? 0.5 * s * exp(th[pos]) : 0.0;
    }
}

}  // namespace bi
