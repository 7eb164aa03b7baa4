// chol_loglik.cuh -- K4: fused covariance assembly + blocked Cholesky + triangular solve + log-det, one CTA per walker.
//
// Replaces, per walker (reference src/bayesian_inference/):
//   emulation.py:522-539   cov = (S diag(v) S^T + U / n_div) * outer(scale, scale)
//   emulation.py:362-388   block-diagonal-per-observable merge
//   log_posterior.py:86-94 dY = mean - y ;  C = cov + Sigma_exp
//   log_posterior.py:123-146  dpotrf, dpotrs, -1/2 dY.alpha - sum log diag
//
// Algorithm: left-looking blocked Cholesky with 32 x 32 tiles.  For panel (block column) j and every row tile i >= j
//     C_ij = A_ij + mask(T_i D T_j^T) - sum_{k<j} L_ik L_jk^T ,   A = Sigma_exp + blockdiag(U)/n_div  (theta independent)
// so the covariance is never materialised: its theta-dependent part enters as one extra K-chunk ("chunk 0") of the
// same DMMA contraction that applies the already-factorised panels.  dY rides along as one extra row of the matrix
// (row tile nt): after panel j its entries are z_j = (L^-1 dY)_j, so the triangular solve costs no extra pass and
//     logp = -1/2 sum z^2 - sum log L_ii .
// The diagonal tile is factorised by one warp in registers (shuffle-broadcast right-looking potrf), inverted, and the
// panel's triangular solve is a DMMA product with that 32 x 32 inverse.
//
// Memory: the per-walker factor lives in a per-CTA slab of HBM/L2 scratch in operand-fragment-major tiles (every
// operand load is one coalesced 256-byte warp access); the slab is re-used for each walker the persistent CTA takes.
// Streamed operand (row tile i, read once) bypasses L1 (.cg); the panel's own row tile j, shared by all warps, goes
// through L1.  A, T are theta independent, shared by all CTAs and stay L2 resident.
#pragma once
#include "common.cuh"

namespace bi {

struct CholParams {
    int F, Fp, nt, K0, nks0, Ptot, d;
    long long W;
    const double* theta;      // (W, d)
    const double* m;          // (W, Ptot)
    const double* v;          // (W, Ptot)
    const double* prior_min;  // (d)
    const double* prior_max;  // (d)
    const double* Tt;         // (K0, Fp)  T_merged transposed (for dY)
    const double* Tm;         // nt tiles, operand-fragment-major, 4*nks0*32 doubles each
    const double* e;          // (Fp) scaler mean - y
    const double* Atiles;     // lower tiles incl. diagonal, accumulator-fragment-major, index i(i+1)/2 + j
    const int* obs_id;        // (Fp)
    const int* tile_omin;     // (nt)
    const int* tile_omax;     // (nt)
    const int* kmin;          // (nt) first structurally non-zero tile column of row tile i (0 when dense)
    const int* imax;          // (nt) last structurally non-zero row tile of panel j (nt-1 when dense)
    double* ws;               // scratch, ws_stride doubles per CTA
    long long ws_stride;
    double* logp;             // (W)
    int* info;                // (W) or nullptr
    int check_prior;          // 1: strict prior box -> -inf ; 0: rows are known to be inside
    // EXPLICIT instantiation only (log_posterior._loglikelihood(y, cov), log_posterior.py:104):
    const double* cov_x;      // (W, F, F) row-major, lower triangle read
    const double* dy_x;       // (W, F)
    // several experimental data sets on one emulator image (closure tests, steer_analysis.py:168-183): walker w uses
    // e + set_id[w] * e_stride and Atiles + set_id[w] * a_stride; set_id == nullptr -> the handle's own data
    const int* set_id;
    long long e_stride, a_stride;
    double* quad_out;         // optional (W): dy^T C^-1 dy on its own (Woodbury combine), else nullptr
    double* logdet_out;       // optional (W): sum log diag chol(C)
};

#ifndef CHOL_MIN_BLOCKS
#define CHOL_MIN_BLOCKS 2
#endif
constexpr int CHOL_WARPS = 8;
constexpr int CHOL_THREADS = CHOL_WARPS * 32;
constexpr int TILE_ELEMS = BI_TILE * BI_TILE;  // 1024 doubles = 8 KB

__host__ __device__ inline long long chol_ws_doubles(int nt) {
    // strictly-lower tiles + one z tile per panel
    return ((long long)nt * (nt - 1) / 2 + nt) * TILE_ELEMS;
}
__host__ __device__ inline size_t chol_smem_bytes(int Fp, int K0) {
    size_t b = 0;
    b += sizeof(double) * (size_t)K0;        // sNegV
    b += sizeof(double) * (size_t)K0;        // sM
    b += sizeof(double) * (size_t)Fp;        // sDy
    b += sizeof(double) * 32 * 33;           // sD
    b += sizeof(double) * TILE_ELEMS;        // sLinv
    b += sizeof(double) * 32;                // sRinv
    b += sizeof(double) * CHOL_WARPS;        // sRed
    b += sizeof(int) * (size_t)Fp;           // sObs
    b += sizeof(int) * 4;                    // flags
    return b;
}

// One warp factorises the 32 x 32 tile held row-major (stride 33) in sD, leaves L in sD and L^-1 in sLinv
// (operand-fragment-major).  Returns 0 or the 1-based index of the first non-positive pivot; adds sum(log L_ii).
__device__ __forceinline__ int warp_potrf_inv(double* sD, double* sLinv, double* sRinv, int lane, double& logdet) {
    double a[32];
#pragma unroll
    for (int c = 0; c < 32; ++c) a[c] = sD[lane * 33 + c];
    int bad = 0;
    double mypiv = 1.0;
#pragma unroll
    for (int c = 0; c < 32; ++c) {
        const double p = shfl_d(a[c], c);
        if (!(p > 0.0) && bad == 0) bad = c + 1;
        if (lane == c) mypiv = p;
        const double rinv = rsqrt(p);
        a[c] *= rinv;
        if (lane == c) sRinv[c] = rinv;
#pragma unroll
        for (int c2 = c + 1; c2 < 32; ++c2) {
            const double l = shfl_d(a[c], c2);
            a[c2] = fma(-a[c], l, a[c2]);
        }
    }
    if (bad) return bad;
    logdet += warp_sum(0.5 * log(mypiv));
#pragma unroll
    for (int c = 0; c < 32; ++c)
        if (c <= lane) sD[lane * 33 + c] = a[c];
    __syncwarp();
    // lane = column of X = L^-1 (forward substitution on the identity)
#pragma unroll
    for (int r = 0; r < 32; ++r) {
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
#pragma unroll
        for (int k = 0; k < r; ++k) {
            const double l = sD[r * 33 + k];
            if ((k & 3) == 0) s0 = fma(l, a[k], s0);
            else if ((k & 3) == 1) s1 = fma(l, a[k], s1);
            else if ((k & 3) == 2) s2 = fma(l, a[k], s2);
            else s3 = fma(l, a[k], s3);
        }
        const double s = (s0 + s1) + (s2 + s3);
        a[r] = ((r == lane ? 1.0 : 0.0) - s) * sRinv[r];   // a[] now holds column `lane` of L^-1
    }
#pragma unroll
    for (int n = 0; n < 32; ++n) sLinv[of_index(n, lane, 8)] = a[n];
    return 0;
}

// EXPLICIT = false: covariance assembled on the fly from the emulator image (the log-posterior path).
// EXPLICIT = true : covariance and residual given per row (the _loglikelihood(y, cov) entry point); no chunk 0.
template <bool EXPLICIT>
__global__ void __launch_bounds__(CHOL_THREADS, CHOL_MIN_BLOCKS) chol_loglik_kernel(const CholParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* sNegV = reinterpret_cast<double*>(smem_raw);
    double* sM = sNegV + P.K0;
    double* sDy = sM + P.K0;
    double* sD = sDy + P.Fp;
    double* sLinv = sD + 32 * 33;
    double* sRinv = sLinv + TILE_ELEMS;
    double* sRed = sRinv + 32;
    int* sObs = reinterpret_cast<int*>(sRed + CHOL_WARPS);
    int* sFlag = sObs + P.Fp;

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    const int g = lane >> 2;
    const int t = lane & 3;
    const int nt = P.nt;
    const int nks0 = P.nks0;
    const int tstride = 4 * nks0 * 32;   // doubles per T tile

    double* wsL = P.ws + (long long)blockIdx.x * P.ws_stride;
    double* wsZ = wsL + (long long)nt * (nt - 1) / 2 * TILE_ELEMS;

    if (!EXPLICIT)
        for (int r = tid; r < P.Fp; r += CHOL_THREADS) sObs[r] = P.obs_id[r];

    for (long long w = blockIdx.x; w < P.W; w += gridDim.x) {
        __syncthreads();   // previous walker fully done with smem
        if (tid == 0) {
            int inside = 1;
            if (!EXPLICIT && P.check_prior) {
                for (int k = 0; k < P.d; ++k) {
                    const double x = P.theta[w * P.d + k];
                    if (!(x > P.prior_min[k] && x < P.prior_max[k])) inside = 0;
                }
            }
            sFlag[0] = inside;
            sFlag[1] = 0;   // failure index
        }
        if (!EXPLICIT)
            for (int q = tid; q < P.K0; q += CHOL_THREADS) {
                const bool ok = q < P.Ptot;
                sM[q] = ok ? P.m[w * P.Ptot + q] : 0.0;
                sNegV[q] = ok ? -P.v[w * P.Ptot + q] : 0.0;
            }
        __syncthreads();
        if (!sFlag[0]) {
            if (tid == 0) {
                P.logp[w] = -INFINITY;
                if (P.info) P.info[w] = -1;
            }
            continue;
        }
        // dY = T m + (mu - y)      (emulation.py:506-509 + log_posterior.py:87)
        for (int r = tid; r < P.Fp; r += CHOL_THREADS) {
            double s;
            if (EXPLICIT) {
                s = (r < P.F) ? P.dy_x[w * P.F + r] : 0.0;
            } else {
                s = __ldg(P.e + (P.set_id ? P.set_id[w] * P.e_stride : 0) + r);
                for (int q = 0; q < P.K0; ++q) s = fma(__ldg(P.Tt + (long long)q * P.Fp + r), sM[q], s);
            }
            sDy[r] = s;
        }
        __syncthreads();

        double logdet = 0.0;   // meaningful in warp 0
        double zpart = 0.0;    // per-lane partial of sum z^2
        int failed = 0;

        for (int j = 0; j < nt; ++j) {
            const int i_hi = P.imax[j];
            const int nreg = 2 * (i_hi - j + 1);   // regular 16-row units of this panel
            const int nunits = nreg + 1;           // + the dY row
            const int omin_j = EXPLICIT ? 0 : P.tile_omin[j];

            // ---------------- phase 1: update every unit of the panel -----------------------------------------
            for (int q = warp; q < nunits; q += CHOL_WARPS) {
                const bool is_z = (q == nreg);
                const int i = is_z ? nt : j + (q >> 1);
                const int h = is_z ? 0 : (q & 1);
                double acc[2][4][2];
#pragma unroll
                for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                    for (int nf = 0; nf < 4; ++nf) acc[mi][nf][0] = acc[mi][nf][1] = 0.0;

                const int k_lo = P.kmin[is_z ? j : i];   // the dY row is dense; L_jk is not
                if (!EXPLICIT && !is_z) {
                    if (P.tile_omax[j] >= P.tile_omin[i]) {
                        // chunk 0: -T_i D T_j^T, masked to pairs of rows of the same observable (emulation.py:362-388)
                        const double* At = P.Tm + (long long)i * tstride + (long long)(2 * h) * nks0 * 32 + lane;
                        const double* Bt = P.Tm + (long long)j * tstride + lane;
                        for (int ks = 0; ks < nks0; ++ks) {
                            const double nv = sNegV[4 * ks + t];
                            double a[2], b[4];
#pragma unroll
                            for (int mi = 0; mi < 2; ++mi) a[mi] = __ldg(At + (mi * nks0 + ks) * 32);
#pragma unroll
                            for (int nf = 0; nf < 4; ++nf) b[nf] = __ldg(Bt + (nf * nks0 + ks) * 32) * nv;
#pragma unroll
                            for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                                for (int nf = 0; nf < 4; ++nf) dmma(acc[mi][nf], a[mi], b[nf]);
                        }
                        if (omin_j != P.tile_omax[i]) {
#pragma unroll
                            for (int mi = 0; mi < 2; ++mi) {
                                const int orow = sObs[i * 32 + (2 * h + mi) * 8 + g];
#pragma unroll
                                for (int nf = 0; nf < 4; ++nf)
#pragma unroll
                                    for (int e = 0; e < 2; ++e)
                                        if (sObs[j * 32 + nf * 8 + 2 * t + e] != orow) acc[mi][nf][e] = 0.0;
                            }
                        }
                    }
                }
                // chunks k < j: + L_ik L_jk^T
                {
                    const double* Abase = is_z ? wsZ : wsL + ((long long)i * (i - 1) / 2) * TILE_ELEMS;
                    const double* Bbase = wsL + ((long long)j * (j - 1) / 2) * TILE_ELEMS;
                    const int aoff = (2 * h) * 8 * 32 + lane;
                    for (int k = k_lo; k < j; ++k) {
                        const double* At = Abase + (long long)k * TILE_ELEMS + aoff;
                        const double* Bt = Bbase + (long long)k * TILE_ELEMS + lane;
                        double a[2][8];
#pragma unroll
                        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                            for (int ks = 0; ks < 8; ++ks) a[mi][ks] = __ldcg(At + (mi * 8 + ks) * 32);
#pragma unroll
                        for (int ks = 0; ks < 8; ++ks) {
                            double b[4];
#pragma unroll
                            for (int nf = 0; nf < 4; ++nf) b[nf] = Bt[(nf * 8 + ks) * 32];
#pragma unroll
                            for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                                for (int nf = 0; nf < 4; ++nf) dmma(acc[mi][nf], a[mi][ks], b[nf]);
                        }
                    }
                }
                // C = A_ij - acc
                if (is_z) {
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                        for (int nf = 0; nf < 4; ++nf)
#pragma unroll
                            for (int e = 0; e < 2; ++e) {
                                const double init = (mi == 0 && g == 0) ? sDy[j * 32 + nf * 8 + 2 * t + e] : 0.0;
                                acc[mi][nf][e] = init - acc[mi][nf][e];
                            }
                } else if (EXPLICIT) {
                    const double* Cw = P.cov_x + w * (long long)P.F * P.F;
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi) {
                        const int r = i * 32 + (2 * h + mi) * 8 + g;
#pragma unroll
                        for (int nf = 0; nf < 4; ++nf)
#pragma unroll
                            for (int e = 0; e < 2; ++e) {
                                const int c = j * 32 + nf * 8 + 2 * t + e;
                                const double init = (r < P.F && c < P.F) ? Cw[(long long)r * P.F + c] : (r == c ? 1.0 : 0.0);
                                acc[mi][nf][e] = init - acc[mi][nf][e];
                            }
                    }
                } else {
                    const double2* At = reinterpret_cast<const double2*>(
                        P.Atiles + (P.set_id ? P.set_id[w] * P.a_stride : 0) + ((long long)i * (i + 1) / 2 + j) * TILE_ELEMS);
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                        for (int nf = 0; nf < 4; ++nf) {
                            const double2 a0 = __ldg(At + ((2 * h + mi) * 4 + nf) * 32 + lane);
                            acc[mi][nf][0] = a0.x - acc[mi][nf][0];
                            acc[mi][nf][1] = a0.y - acc[mi][nf][1];
                        }
                }
                if (!is_z && i == j) {
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                        for (int nf = 0; nf < 4; ++nf)
#pragma unroll
                            for (int e = 0; e < 2; ++e)
                                sD[(h * 16 + mi * 8 + g) * 33 + nf * 8 + 2 * t + e] = acc[mi][nf][e];
                } else {
                    // park the un-solved strip in its own slot of the factor (same lanes read it back in phase 3)
                    double* Ct = (is_z ? wsZ + (long long)j * TILE_ELEMS
                                       : wsL + ((long long)i * (i - 1) / 2 + j) * TILE_ELEMS) + (2 * h) * 8 * 32 + lane;
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi) {
                        double a8[8];
                        cf_to_of(acc[mi], a8, lane);
#pragma unroll
                        for (int ks = 0; ks < 8; ++ks) Ct[(mi * 8 + ks) * 32] = a8[ks];
                    }
                }
            }
            __syncthreads();

            // ---------------- phase 2: factorise + invert the diagonal tile ------------------------------------
            if (warp == 0) {
                const int bad = warp_potrf_inv(sD, sLinv, sRinv, lane, logdet);
                if (bad && lane == 0) sFlag[1] = j * 32 + bad;
            }
            __syncthreads();
            if (sFlag[1]) {
                failed = 1;
                break;
            }

            // ---------------- phase 3: triangular solve of the panel: L_ij = C_ij L_jj^-T ------------------------
            for (int q = warp; q < nunits; q += CHOL_WARPS) {
                if (q < 2) continue;   // the diagonal tile itself
                const bool is_z = (q == nreg);
                const int i = is_z ? nt : j + (q >> 1);
                const int h = is_z ? 0 : (q & 1);
                double* Ct = (is_z ? wsZ + (long long)j * TILE_ELEMS
                                   : wsL + ((long long)i * (i - 1) / 2 + j) * TILE_ELEMS) + (2 * h) * 8 * 32 + lane;
                double a[2][8];
#pragma unroll
                for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                    for (int ks = 0; ks < 8; ++ks) a[mi][ks] = Ct[(mi * 8 + ks) * 32];
                double x[2][4][2];
#pragma unroll
                for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                    for (int nf = 0; nf < 4; ++nf) x[mi][nf][0] = x[mi][nf][1] = 0.0;
#pragma unroll
                for (int ks = 0; ks < 8; ++ks) {
#pragma unroll
                    for (int nf = 0; nf < 4; ++nf) {
                        if (ks <= 2 * nf + 1) {   // L^-1 is lower triangular: B[k][n] = Linv[n][k] = 0 for k > n
                            const double b = sLinv[(nf * 8 + ks) * 32 + lane];
#pragma unroll
                            for (int mi = 0; mi < 2; ++mi) dmma(x[mi][nf], a[mi][ks], b);
                        }
                    }
                }
                if (is_z) {
                    if (g == 0) {
#pragma unroll
                        for (int nf = 0; nf < 4; ++nf) zpart += x[0][nf][0] * x[0][nf][0] + x[0][nf][1] * x[0][nf][1];
                    }
                }
#pragma unroll
                for (int mi = 0; mi < 2; ++mi) {
                    double a8[8];
                    cf_to_of(x[mi], a8, lane);
#pragma unroll
                    for (int ks = 0; ks < 8; ++ks) Ct[(mi * 8 + ks) * 32] = a8[ks];
                }
            }
            __syncthreads();   // the panel's tiles are visible to every warp before the next panel reads them
        }

        // ---------------- epilogue ---------------------------------------------------------------------------
        const double zs = warp_sum(zpart);
        if (lane == 0) sRed[warp] = zs;
        __syncthreads();
        if (tid == 0) {
            if (failed) {
                P.logp[w] = __longlong_as_double(0x7ff8000000000000LL);
                if (EXPLICIT && P.quad_out) P.quad_out[w] = P.logdet_out[w] = __longlong_as_double(0x7ff8000000000000LL);
                if (P.info) P.info[w] = sFlag[1] > P.F ? P.F : sFlag[1];
            } else {
                double tot = 0.0;
                for (int k = 0; k < CHOL_WARPS; ++k) tot += sRed[k];
                P.logp[w] = -0.5 * tot - logdet;
                if (EXPLICIT && P.quad_out) {
                    P.quad_out[w] = tot;
                    P.logdet_out[w] = logdet;
                }
                if (P.info) P.info[w] = 0;
            }
        }
    }
}

}  // namespace bi
