// gp_predict.cuh -- K1 (kernel vectors) + K2 (GP predictive mean and variance), one CTA per (walker tile, PC).
//
// Replaces, per principal component (reference emulation.py:496-499 -> sklearn):
//   _gpr.py:446        K_trans = kernel_(X, X_train_)       kernels.py:1685-1712 (Matern), :1530-1545 (RBF),
//                                                            :1244-1280 (+constant), :1374-1419 (White: 0 off-diagonal)
//   _gpr.py:447        mean = K_trans @ alpha_
//   _gpr.py:460-481    V = solve_triangular(L_, K_trans^T) ;  var = diag - sum V^2 ; clip at 0 (:485-491)
//   emulation.py:499   v = std**2
//
// K1: thread per (walker, design point).  The design, already divided by this PC's length scales and transposed
// (d, N), is staged in shared memory; theta / ell per walker likewise.  Distances are direct differences summed
// over the parameters in order, like scipy cdist.  The kernel vector is written straight into shared memory in
// operand-fragment-major layout (and, for bi_kernel_vectors, to HBM row-major).
// K2: V = k L^-T as an FP64 DMMA contraction against L^-1 (computed once at load time), row n = N of the B operand
// holds alpha so the mean falls out of the same product; sum V^2 is reduced in registers -> shuffles -> shared.
#pragma once
#include "common.cuh"

namespace bi {

struct PcDesc {
    const double* XsT;     // (d, N)  X[j][k] / ell[k], transposed
    const double* ell;     // (d)
    const double* LinvF;   // ntn tiles of (32 rows n) x (4*nks) in operand-fragment-major; row N = alpha
    int N;
    int nks;               // ceil(N / 4) rounded up to a multiple of 4 (k-steps are consumed in groups of 4)
    int ntn;               // ceil((N + 1) / 32)
    int kind;
    double cst;            // ConstantKernel value
    double diag;           // kernel_.diag = 1 + constant + noise
    int group;
    int local;             // index of this PC inside its group
};

constexpr int GP_WT = 32;          // walkers per CTA
constexpr int GP_THREADS = 256;

__host__ __device__ inline size_t gp_smem_bytes(int d, int Nmax) {
    const int nks = ((Nmax + 15) / 16) * 4;   // K steps, padded to whole groups of 4
    size_t b = 0;
    b += sizeof(double) * (size_t)GP_WT * 4 * nks;   // sA  (32 walkers x 4*nks), operand-fragment-major
    b += sizeof(double) * (size_t)d * Nmax;          // sX
    b += sizeof(double) * (size_t)GP_WT * d;         // sTh
    b += sizeof(double) * GP_WT * 4;                 // sSq  [wn][row]
    b += sizeof(double) * GP_WT * 4;                 // sMn  [wn][row]
    return b;
}

__device__ __forceinline__ double stationary_kernel(int kind, double sq) {
    // sq = squared scaled distance
    if (kind == 3) return exp(-0.5 * sq);
    const double r = sqrt(sq);
    if (kind == 0) return exp(-r);
    if (kind == 1) {
        const double tt = r * 1.7320508075688772;     // math.sqrt(3) * dists
        return (1.0 + tt) * exp(-tt);
    }
    const double tt = r * 2.23606797749979;            // math.sqrt(5) * dists
    return (1.0 + tt + tt * tt / 3.0) * exp(-tt);
}

// grid: (ceil(W / 32), n_pc_selected).  pcs[blockIdx.y + pc_first] describes the PC.
// k_out (optional): (W, k_ld_pc, N) row-major slice for this group, PC index = desc.local.
// m_out / v_out (optional): (W, Ptot) at column blockIdx.y + pc_first.
template <bool DO_GP>
__global__ void __launch_bounds__(GP_THREADS, 2) gp_predict_kernel(const PcDesc* __restrict__ pcs, int pc_first, int d,
                                                                const double* __restrict__ theta, long long W,
                                                                double* __restrict__ k_out, int k_npc,
                                                                double* __restrict__ m_out, double* __restrict__ v_out,
                                                                int Ptot) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const PcDesc pc = pcs[pc_first + blockIdx.y];
    const int N = pc.N, nks = pc.nks;
    double* sA = reinterpret_cast<double*>(smem_raw);
    double* sX = sA + (size_t)GP_WT * 4 * nks;
    double* sTh = sX + (size_t)d * N;
    double* sSq = sTh + GP_WT * d;
    double* sMn = sSq + GP_WT * 4;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long w0 = (long long)blockIdx.x * GP_WT;

    for (int idx = tid; idx < d * N; idx += GP_THREADS) sX[idx] = __ldg(pc.XsT + idx);
    for (int idx = tid; idx < GP_WT * d; idx += GP_THREADS) {
        const int wl = idx / d, k = idx - wl * d;
        const long long w = w0 + wl;
        sTh[idx] = (w < W) ? theta[w * d + k] / __ldg(pc.ell + k) : 0.0;
    }
    // zero the K padding of the operand tile (columns N .. 4*nks-1)
    if (DO_GP) {
        const int npad = 4 * nks - N;
        for (int idx = tid; idx < GP_WT * npad; idx += GP_THREADS) {
            const int wl = idx / npad, c = N + (idx - wl * npad);
            sA[of_index(wl, c, nks)] = 0.0;
        }
    }
    __syncthreads();

    // ---- K1 -------------------------------------------------------------------------------------------------
    for (int idx = tid; idx < GP_WT * N; idx += GP_THREADS) {
        const int wl = idx / N, j = idx - wl * N;
        double sq = 0.0;
        for (int k = 0; k < d; ++k) {
            const double df = sTh[wl * d + k] - sX[k * N + j];
            sq = fma(df, df, sq);
        }
        const double kv = stationary_kernel(pc.kind, sq) + pc.cst;
        if (DO_GP) sA[of_index(wl, j, nks)] = kv;
        if (k_out != nullptr && w0 + wl < W) k_out[((w0 + wl) * k_npc + pc.local) * N + j] = kv;
    }
    if (!DO_GP) return;
    __syncthreads();

    // ---- K2 -------------------------------------------------------------------------------------------------
    // warp (wm, wn): walker rows wm*16.., n-tiles {t : snake(t) == wn}.  The B operand (rows of L^-1, L2 resident,
    // shared by every CTA of this PC) is streamed global -> registers in groups of 4 k-steps, one group ahead of the
    // DMMAs that consume it (the loads' L2 latency is what bounded the first version of this kernel).
    const int wm = warp & 1;    // rows wm*16 .. wm*16+15
    const int wn = warp >> 1;   // n-tile owner 0..3
    const int g = lane >> 2, t = lane & 3;
    double sumsq[2] = {0.0, 0.0};
    double meanv[2] = {0.0, 0.0};
    const double* sAw = sA + (size_t)(2 * wm) * nks * 32 + lane;
    const size_t tile_stride = (size_t)4 * nks * 32;

    auto tile_of = [&](int idx) {   // idx-th n-tile owned by this warp (snake order balances the triangle)
        const int blk = idx >> 1;
        return blk * 8 + ((idx & 1) ? 7 - wn : wn);
    };
    auto load_group = [&](double (&bb)[4][4], int tile, int grp) {
        const double* Bt = pc.LinvF + (size_t)tile * tile_stride + (size_t)(grp * 4) * 32 + lane;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
#pragma unroll
            for (int nf = 0; nf < 4; ++nf) bb[kk][nf] = __ldg(Bt + ((size_t)nf * nks + kk) * 32);
    };

    int idx = 0;
    int tile = tile_of(0);
    if (tile < pc.ntn) {
        double bcur[4][4], bnxt[4][4];
        load_group(bcur, tile, 0);
        while (tile < pc.ntn) {
            const int ngrp = min(nks, 8 * (tile + 1)) >> 2;   // nks is a multiple of 4
            double acc[2][4][2];
#pragma unroll
            for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                for (int nf = 0; nf < 4; ++nf) acc[mi][nf][0] = acc[mi][nf][1] = 0.0;
            const int next_tile = tile_of(idx + 1);
            for (int grp = 0; grp < ngrp; ++grp) {
                // prefetch the next group (of this tile, or the first of the next tile)
                if (grp + 1 < ngrp) load_group(bnxt, tile, grp + 1);
                else if (next_tile < pc.ntn) load_group(bnxt, next_tile, 0);
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                    double a[2];
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi) a[mi] = sAw[(mi * nks + grp * 4 + kk) * 32];
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                        for (int nf = 0; nf < 4; ++nf) dmma(acc[mi][nf], a[mi], bcur[kk][nf]);
                }
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)
#pragma unroll
                    for (int nf = 0; nf < 4; ++nf) bcur[kk][nf] = bnxt[kk][nf];
            }
#pragma unroll
            for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                for (int nf = 0; nf < 4; ++nf)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int n = tile * 32 + nf * 8 + 2 * t + e;
                        const double x = acc[mi][nf][e];
                        if (n < N) sumsq[mi] = fma(x, x, sumsq[mi]);
                        else if (n == N) meanv[mi] = x;
                    }
            ++idx;
            tile = next_tile;
        }
    }
#pragma unroll
    for (int mi = 0; mi < 2; ++mi) {
        sumsq[mi] += __shfl_xor_sync(0xffffffffu, sumsq[mi], 1);
        sumsq[mi] += __shfl_xor_sync(0xffffffffu, sumsq[mi], 2);
        meanv[mi] += __shfl_xor_sync(0xffffffffu, meanv[mi], 1);   // only one lane of the quad holds column N
        meanv[mi] += __shfl_xor_sync(0xffffffffu, meanv[mi], 2);
        if (t == 0) {
            const int row = wm * 16 + mi * 8 + g;
            sSq[wn * GP_WT + row] = sumsq[mi];
            sMn[wn * GP_WT + row] = meanv[mi];
        }
    }
    __syncthreads();
    if (tid < GP_WT) {
        const long long w = w0 + tid;
        if (w < W) {
            const double ss = (sSq[tid] + sSq[GP_WT + tid]) + (sSq[2 * GP_WT + tid] + sSq[3 * GP_WT + tid]);
            const double mu = (sMn[tid] + sMn[GP_WT + tid]) + (sMn[2 * GP_WT + tid] + sMn[3 * GP_WT + tid]);
            double var = pc.diag - ss;
            if (var < 0.0) var = 0.0;                       // _gpr.py:485-491
            const double sd = sqrt(var);                    // _gpr.py:500 returns the std ...
            const long long o = w * Ptot + pc_first + blockIdx.y;
            m_out[o] = mu;
            v_out[o] = sd * sd;                             // ... and the reference squares it, emulation.py:499
        }
    }
}

}  // namespace bi
