// cov_assemble.cuh -- K3: inverse-PCA mean and (standalone) covariance assembly for the predict() API.
//
// Replaces (reference src/bayesian_inference/emulation.py):
//   :506-509  mean = (m @ S_P) * scale + scaler.mean_                (no pca.mean_)
//   :522-539  cov  = (S diag(v) S^T + U / n_div) * outer(scale, scale)
//   :346-406  merge over groups: block diagonal per observable (mask), sorted-observable order
//
// The covariance is written to HBM here only because predict() returns it; the log-posterior path (chol_pipe.cuh)
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
    const int* tile_omin; // smallest / largest observable id of each 32-row tile (with obs), else nullptr
    const int* tile_omax;
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

// ---- pieces shared by the two covariance kernels (one warp per 32 x 32 tile (it, jt), jt <= it) ---------------------------

// structurally zero tile and its mirror: plain zero fill, one 256-byte row segment per 16 lanes
__device__ __forceinline__ void k3_zero_tile(const CovParams& P, double* out, int i0, int j0, int lane) {
    const int rr = lane >> 4, cc = (lane & 15) * 2;
    if ((P.Fo & 1) == 0) {             // even row length: every (row, even column) pair is 16-byte aligned
        const double2 z2 = make_double2(0.0, 0.0);
#pragma unroll 4
        for (int r = rr; r < 32; r += 2) {
            const int row = i0 + r, mcol = i0 + cc;          // jt < it here: j0 + cc + 1 < Fo and j0 + r < Fo always
            if (row < P.Fo) *reinterpret_cast<double2*>(out + (long long)row * P.Fo + j0 + cc) = z2;
            if (mcol < P.Fo) *reinterpret_cast<double2*>(out + (long long)(j0 + r) * P.Fo + mcol) = z2;
        }
        return;
    }
#pragma unroll 4
    for (int r = rr; r < 32; r += 2) {
        const int row = i0 + r, col = j0 + cc;           // jt < it here, so col + 1 < Fo always
        if (row < P.Fo) {
            out[(long long)row * P.Fo + col] = 0.0;
            out[(long long)row * P.Fo + col + 1] = 0.0;
        }
        const int mrow = j0 + r, mcol = i0 + cc;
        if (mcol < P.Fo) out[(long long)mrow * P.Fo + mcol] = 0.0;
        if (mcol + 1 < P.Fo) out[(long long)mrow * P.Fo + mcol + 1] = 0.0;
    }
}

// The accumulators start from U / n_div: the 16 loads are in flight together while the first operands arrive (one L2
// round trip per tile; as an epilogue they cost four dependent round trips: 2.40 -> 1.35 ms on one dense 600-bin
// block).  Rows / columns past Fo are clamped (never stored).
__device__ __forceinline__ void k3_init_acc(const CovParams& P, double (&acc)[4][4][2], int i0, int j0, double rinv, int lane) {
    const int g = lane >> 2, t = lane & 3;
    const bool even = (P.Fo & 1) == 0;
#pragma unroll
    for (int mf = 0; mf < 4; ++mf) {
        const double* Ur = P.Uc + (long long)min(i0 + mf * 8 + g, P.Fo - 1) * P.Fo;
#pragma unroll
        for (int nf = 0; nf < 4; ++nf) {
            const int col = j0 + nf * 8 + 2 * t;
            if (even) {
                const double2 uu = __ldcg(reinterpret_cast<const double2*>(Ur + min(col, P.Fo - 2)));
                acc[mf][nf][0] = uu.x * rinv;
                acc[mf][nf][1] = uu.y * rinv;
            } else {
                acc[mf][nf][0] = __ldcg(Ur + min(col, P.Fo - 1)) * rinv;
                acc[mf][nf][1] = __ldcg(Ur + min(col + 1, P.Fo - 1)) * rinv;
            }
        }
    }
}

// observable mask (block-diagonal-per-observable merge), tile (it, jt) and, with the same values, its mirror (jt, it)
__device__ __forceinline__ void k3_store_tile(const CovParams& P, double* out, const double (&acc)[4][4][2], int it, int jt, int lane) {
    const int g = lane >> 2, t = lane & 3;
    const int i0 = it * 32, j0 = jt * 32;
    int ocol[4][2];
#pragma unroll
    for (int nf = 0; nf < 4; ++nf)
#pragma unroll
        for (int e = 0; e < 2; ++e) ocol[nf][e] = P.obs ? P.obs[min(j0 + nf * 8 + 2 * t + e, P.Fo - 1)] : 0;
    const bool even = (P.Fo & 1) == 0;
#pragma unroll
    for (int mf = 0; mf < 4; ++mf) {
        const int row = i0 + mf * 8 + g;
        if (row >= P.Fo) continue;
        const int orow = P.obs ? P.obs[row] : 0;
        double* po = out + (long long)row * P.Fo;
#pragma unroll
        for (int nf = 0; nf < 4; ++nf) {
            const int col = j0 + nf * 8 + 2 * t;        // columns col, col + 1
            if (col >= P.Fo) continue;
            const bool two = col + 1 < P.Fo;
            const bool s0 = ocol[nf][0] == orow, s1 = two && ocol[nf][1] == orow;
            const double o0 = s0 ? acc[mf][nf][0] : 0.0;
            const double o1 = s1 ? acc[mf][nf][1] : 0.0;
            if (two && even) {
                *reinterpret_cast<double2*>(po + col) = make_double2(o0, o1);
            } else {
                po[col] = o0;
                if (two) po[col + 1] = o1;
            }
            if (jt < it) {                                                   // mirrored tile (j, i): the same values
                out[(long long)col * P.Fo + row] = o0;
                if (two) out[(long long)(col + 1) * P.Fo + row] = o1;
            }
        }
    }
}

// grid (ceil(Fo/32), W), 256 threads = 8 warps.  CTA (i, w) produces row tile i of walker w's covariance and, by
// symmetry of T diag(v) T^T, the mirrored column tile: warp q takes the tiles j = q, q + 8, ... <= i, contracts
// (T_i diag(v)) T_j^T on the FP64 tensor pipe (DMMA m8n8k4, operand fragments read straight from the row-major,
// L1/L2-resident T) on top of U_ij / n_div and writes tile (i, j) and, with the same values, its transpose (j, i) -- the
// reference's U = S_u diag(ev) S_u^T is symmetric up to the rounding of its own matrix product, ~1e-16.  Tiles whose
// observables do not overlap (block-diagonal-per-observable merge, emulation.py:346-406) skip the contraction and
// write zeros.  Half the flops of the full product, so the kernel is bound by its 8 F^2 bytes of HBM writes.
// Used when T does not fit in shared memory (cov_persist_kernel below otherwise).
__global__ void __launch_bounds__(256, 2) cov_assemble_kernel(const CovParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int K0 = P.K0, nks = K0 >> 2;
    double* sv = reinterpret_cast<double*>(smem_raw);
    double* sA = sv + K0;                      // T_i diag(v), operand-fragment-major: [(mf * nks + ks) * 32 + lane]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const long long w = blockIdx.y;
    const int it = blockIdx.x, i0 = it * 32;
    for (int q = tid; q < K0; q += 256) sv[q] = (q < P.nq) ? P.v[w * P.Ptot + P.voff + q] : 0.0;
    __syncthreads();
    for (int idx = tid; idx < 32 * K0; idx += 256) {
        const int r = idx / K0, q = idx - r * K0;
        sA[of_index(r, q, nks)] = (i0 + r < P.Fo) ? __ldg(P.Trow + (long long)(i0 + r) * K0 + q) * sv[q] : 0.0;
    }
    __syncthreads();
    double* out = P.cov_out + w * (long long)P.Fo * P.Fo;
    const int omin_i = P.tile_omin ? P.tile_omin[it] : 0;
    const double rinv = 1.0 / P.n_div;         // U / n_div as a product: within 1 ulp of the reference's quotient
    for (int jt = warp; jt <= it; jt += 8) {
        const int j0 = jt * 32;
        const bool overlap = P.tile_omax ? (P.tile_omax[jt] >= omin_i) : true;
        if (!overlap) {
            k3_zero_tile(P, out, i0, j0, lane);
            continue;
        }
        double acc[4][4][2];
        k3_init_acc(P, acc, i0, j0, rinv, lane);
        // operands: A = rows of T_i scaled by v (shared by the 8 warps: staged once in shared memory, fragment-major),
        // B = rows of T_j read straight from the row-major, L1/L2-resident T (row index clamped, never predicated)
        {
            int rb[4];
#pragma unroll
            for (int nf = 0; nf < 4; ++nf) rb[nf] = min(j0 + nf * 8 + g, P.Fo - 1) * K0 + t;
            for (int ks = 0; ks < nks; ++ks) {
                double a[4], b[4];
#pragma unroll
                for (int mf = 0; mf < 4; ++mf) a[mf] = sA[(mf * nks + ks) * 32 + lane];
#pragma unroll
                for (int nf = 0; nf < 4; ++nf) b[nf] = __ldg(P.Trow + rb[nf] + ks * 4);
#pragma unroll
                for (int mf = 0; mf < 4; ++mf)
#pragma unroll
                    for (int nf = 0; nf < 4; ++nf) dmma(acc[mf][nf], a[mf], b[nf]);
            }
        }
        k3_store_tile(P, out, acc, it, jt, lane);
    }
}

// Persistent variant, used whenever T (padded to 32-row tiles) fits in shared memory: one 512-thread CTA per SM keeps all of
// T in shared memory as operand fragments, and every WARP walks the global list of (walker, tile) items with stride
// (warps in the grid) -- no CTA-level synchronisation after the prologue, every warp busy until the list ends (the
// (row tile, walker) grid above leaves 28 % of its warp slots without a tile), both operands are conflict-free LDS.64
// (A is scaled by v in registers: 4 DMUL per 16 DMMA), and consecutive items share a walker so the writes of a moment
// fall into a dozen covariance matrices.
constexpr int K3P_THREADS = 512;
constexpr size_t K3P_SMEM_MAX = 200 * 1024;
__host__ __device__ inline size_t k3p_smem_bytes(int Fo, int K0) { return (size_t)((Fo + 31) / 32) * 32 * K0 * sizeof(double); }

__global__ void __launch_bounds__(K3P_THREADS, 1) cov_persist_kernel(const CovParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* sT = reinterpret_cast<double*>(smem_raw);   // [tile][mf][ks][lane]
    const int K0 = P.K0, nks = K0 >> 2, nt = (P.Fo + 31) / 32;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int t = lane & 3;
    for (int idx = tid; idx < nt * 32 * K0; idx += K3P_THREADS) {
        const int r = idx / K0, q = idx - r * K0;
        sT[(size_t)(r >> 5) * 32 * K0 + of_index(r & 31, q, nks)] = (r < P.Fo) ? __ldg(P.Trow + (long long)r * K0 + q) : 0.0;
    }
    __syncthreads();
    const double rinv = 1.0 / P.n_div;
    const long long ntile = (long long)nt * (nt + 1) / 2;
    const long long nitem = P.W * ntile;
    const long long stride = (long long)gridDim.x * (K3P_THREADS / 32);
    for (long long item = (long long)blockIdx.x * (K3P_THREADS / 32) + warp; item < nitem; item += stride) {
        const long long w = item / ntile;
        const int q = (int)(item - w * ntile);
        int it = (int)((sqrtf(8.0f * (float)q + 1.0f) - 1.0f) * 0.5f);      // q = it (it + 1) / 2 + jt, 0 <= jt <= it
        while (it * (it + 1) / 2 > q) --it;
        while ((it + 1) * (it + 2) / 2 <= q) ++it;
        const int jt = q - it * (it + 1) / 2;
        const int i0 = it * 32, j0 = jt * 32;
        double* out = P.cov_out + w * (long long)P.Fo * P.Fo;
        const bool overlap = P.tile_omax ? (P.tile_omax[jt] >= P.tile_omin[it]) : true;
        if (!overlap) {
            k3_zero_tile(P, out, i0, j0, lane);
            continue;
        }
        double acc[4][4][2];
        k3_init_acc(P, acc, i0, j0, rinv, lane);
        const double* vw = P.v + w * P.Ptot + P.voff;
        const double* sAi = sT + (size_t)it * 32 * K0 + lane;
        const double* sBj = sT + (size_t)jt * 32 * K0 + lane;
        for (int ks = 0; ks < nks; ++ks) {
            const int qk = ks * 4 + t;
            const double vk = (qk < P.nq) ? __ldg(vw + qk) : 0.0;
            double a[4], b[4];
#pragma unroll
            for (int mf = 0; mf < 4; ++mf) a[mf] = sAi[(mf * nks + ks) * 32] * vk;
#pragma unroll
            for (int nf = 0; nf < 4; ++nf) b[nf] = sBj[(nf * nks + ks) * 32];
#pragma unroll
            for (int mf = 0; mf < 4; ++mf)
#pragma unroll
                for (int nf = 0; nf < 4; ++nf) dmma(acc[mf][nf], a[mf], b[nf]);
        }
        k3_store_tile(P, out, acc, it, jt, lane);
    }
}

}  // namespace bi
