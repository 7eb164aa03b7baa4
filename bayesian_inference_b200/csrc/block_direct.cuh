// block_direct.cuh -- K5 (direct): the log-likelihood observable block by observable block, one warp per (walker, block).
//
// The reference merges the emulation groups into a covariance that is block diagonal per observable
// (emulation.py:362-388) and then factorises the whole F x F matrix with dpotrf (log_posterior.py:123-146).
// When Sigma_exp does not couple observables the same value is the sum over the blocks
//     logp = sum_b [ -1/2 dy_b^T C_b^-1 dy_b - sum log diag chol(C_b) ],   C_b = A_b + T_b D T_b^T,  dy_b = e_b + T_b m
// (structured variant: "if f_b <= P: direct f_b x f_b Cholesky").  With f_b <= 31 a block lives in ONE
// warp's registers, one matrix row per lane:
//   * assembly  a[r][c] = A_b[r][c] + sum_p (T[r][p] v_p) T[c][p] on the FP64 tensor pipe (mma.sync m8n8k4): T_b is staged
//     in shared memory once per CTA as operand fragments (a CTA works on one block for a batch of walkers); only the
//     lower 8 x 8 tiles are computed; the accumulators go through a per-warp row-major tile to one row per lane;
//   * one lane carries the augmented row dy_b (m in place of T[r][.] v, e_b in place of A_b[r][.]): lane FB - 1 inside the
//     last 8-row fragment when every block leaves that row free (f < FB), else lane 31 with a fragment row of its own, so the right-looking elimination turns it into z = L^-1 dy and the quadratic form is sum z^2: no
//     separate substitution (the device K4 kernels use the same augmentation);
//   * right-looking Cholesky in registers: the pivot column is broadcast through a double-buffered 32-entry line of shared
//     memory (the first version used warp shuffles and was bound by the shuffle / load-store unit); rows between f and the
//     padded size are identity rows.
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

// the combine step fused into block_direct_kernel (counter == nullptr: separate block_combine_kernel launch)
struct BdFuse {
    int* counter;             // (max_batch) zero between launches: block terms delivered per walker
    const double* theta;      // (W, d)
    const double* lo;         // (d) strict prior box
    const double* hi;
    double* logp;             // (W)
    int* info;                // (W) or nullptr
    int d;
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

__host__ __device__ inline size_t block_direct_smem_bytes(int FB, int nks) {
    // CTA: sTf (T_b as DMMA operand fragments, FB/8 x nks x 32) + sA (32 x (FB + 2): A_b rows, identity rows f.., one row = e_b;
    //      the row stride FB + 2 is an odd number of 16-byte words: a lane per row reads it with LDS.128 free of bank conflicts)
    // per warp: sC ((FB+1) x (FB+2): the assembled block, row-major) + sCol (2 x 32: column broadcast) + v and m (2 x 4 nks)
    return sizeof(double) * ((size_t)(FB / 8) * nks * 32 + (size_t)32 * (FB + 2) +
                             (size_t)BD_WARPS * ((size_t)(FB + 1) * (FB + 2) + 64 + 8 * nks));
}

// grid (n_blocks, ceil(W / (BD_WARPS * wpw))): warp u of CTA (b, y) takes walkers y * BD_WARPS * wpw + u + BD_WARPS * i, i < wpw.
// FB = f padded to a multiple of 8 (template: the matrix row lives in FB registers); nksmax = ceil(max P / 4).
//
// Per (walker, block), one warp:
//  1. assembly on the FP64 tensor pipe: [T_b diag(v) ; m^T] T_b^T by mma.sync m8n8k4 -- lower 8 x 8 tiles of the f x f block plus
//     one row of tiles for the augmented row (T_b m); T_b sits in shared memory as operand fragments, staged once per CTA
//     (a CTA works on one block for a batch of walkers), and the same fragment is the B operand and, scaled by v, the A
//     operand: 5 shared-memory loads per 9 DMMAs;
//  2. accumulator fragments -> the warp's row-major tile in shared memory -> one matrix row per lane (+ A_b, theta
//     independent; lane 31 = the augmented row dy_b = e_b + T_b m);
//  3. right-looking Cholesky in registers; column c of the current matrix is broadcast through shared memory (one store,
//     one 16-byte broadcast load per two multipliers) -- the first version used two SHFL per multiplier and was bound by the
//     SM's shuffle / load-store unit (ncu: 1073 of 3200 instructions per task, short-scoreboard + MIO-throttle stalls 50 %).
//     The update uses u = a[r][c] / pivot and the unscaled column, L[r][c] L[c2][c] = u * a[c2][c].
// Resident CTAs per SM by row length (registers for warps).
// AUGIN: every block has f <= FB - 1, so the augmented row rides in row FB - 1 of the last 8-row fragment (lane FB - 1)
// instead of a fragment row of its own -- a third fewer DMMAs in the assembly at FB = 24.  FB = 32 is always AUGIN (f <= 31).
template <int FB, bool AUGIN_>
__global__ void __launch_bounds__(BD_THREADS, FB <= 16 ? 4 : (FB <= 24 ? 3 : 2)) block_direct_kernel(
    const BdBlock* __restrict__ blocks, int nB, long long W, int wpw, int Ptot, int K0, int nksmax, const double* __restrict__ m,
    const double* __restrict__ v, const double* __restrict__ Trow, const double* __restrict__ e, const double* __restrict__ Aall,
    double* __restrict__ part, int* __restrict__ bad_out, const int* __restrict__ set_id, long long e_stride, long long a_stride,
    const BdFuse fuse) {
    constexpr int NFR = FB / 8;                               // 8-row fragments holding matrix rows
    constexpr bool AUGIN = AUGIN_ || FB == 32;
    constexpr bool AUGF = !AUGIN;                             // the augmented row has a fragment row of its own
    constexpr int LDC = FB + 2;
    constexpr int NCOL = AUGIN ? FB - 1 : FB;                 // columns eliminated (column FB - 1 would be the augmented row's own)
    constexpr int AUGROW = AUGIN ? FB - 1 : FB;               // row of sC holding the augmented row
    constexpr int AL = AUGIN ? FB - 1 : 31;                   // lane (= row of sA) that carries the augmented row
    extern __shared__ __align__(16) unsigned char bd_smem[];
    double* sTf = reinterpret_cast<double*>(bd_smem);        // [mf][ks][lane] = T_b[8 mf + g][4 ks + t]
    double* sA = sTf + (size_t)NFR * nksmax * 32;             // [r][c]  A_b, identity rows f.., row AL = e_b
    double* sW = sA + 32 * LDC;                               // per-warp scratch
    const BdBlock B = blocks[blockIdx.x];
    const int f = B.f, P = B.P;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const int wstride = (FB + 1) * LDC + 64 + 8 * nksmax;
    double* sC = sW + (size_t)warp * wstride;                 // [(FB + 1)][LDC]
    double* sCol = sC + (FB + 1) * LDC;                       // [2][32]
    double* sV = sCol + 64;                                   // [4 nksmax]
    double* sM = sV + 4 * nksmax;

    for (int idx = tid; idx < NFR * nksmax * 32; idx += BD_THREADS) {
        const int l = idx & 31, ks = (idx >> 5) % nksmax, mf = (idx >> 5) / nksmax;
        const int r = 8 * mf + (l >> 2), p = 4 * ks + (l & 3);
        sTf[idx] = (r < f && p < P) ? __ldg(Trow + (size_t)(B.row0 + r) * K0 + B.col0 + p) : 0.0;
    }
    for (int idx = tid; idx < 32 * FB; idx += BD_THREADS) {
        const int r = idx / FB, c = idx - r * FB;
        double x;
        if (r == AL) x = (c < f) ? __ldg(e + B.row0 + c) : 0.0;
        else if (r < f && c < f) x = __ldg(Aall + B.a_off + (size_t)r * f + c);
        else x = (r == c) ? 1.0 : 0.0;
        sA[r * LDC + c] = x;
    }
    for (int idx = lane; idx < (FB + 1) * LDC; idx += 32) sC[idx] = 0.0;   // tiles above the diagonal are never written
    __syncthreads();
    pdl_wait();        // T_b, A_b, e_b above are theta independent and older than the previous kernel (K1+K2: m, v)
    pdl_trigger();

    const bool is_aug = lane == AL;
    const int nks = (P + 3) >> 2;
    const long long w_base = (long long)blockIdx.y * BD_WARPS * wpw + warp;
    for (int it = 0; it < wpw; ++it) {
        const long long w = w_base + (long long)it * BD_WARPS;
        if (w >= W) break;                                    // warp-uniform
        __syncwarp();
        for (int p = lane; p < 4 * nks; p += 32) {
            const bool ok = p < P;
            sV[p] = ok ? __ldg(v + w * Ptot + B.col0 + p) : 0.0;
            sM[p] = ok ? __ldg(m + w * Ptot + B.col0 + p) : 0.0;
        }
        __syncwarp();
        // ---- 1. assembly on the tensor pipe
        {
            double acc[NFR][NFR][2], aug[NFR][2];
#pragma unroll
            for (int mi = 0; mi < NFR; ++mi) {
                aug[mi][0] = aug[mi][1] = 0.0;
#pragma unroll
                for (int nf = 0; nf < NFR; ++nf) acc[mi][nf][0] = acc[mi][nf][1] = 0.0;
            }
            for (int ks = 0; ks < nks; ++ks) {
                double tf[NFR], ta[NFR];
                const double vk = sV[4 * ks + t], mk = sM[4 * ks + t];
#pragma unroll
                for (int mf = 0; mf < NFR; ++mf) {
                    tf[mf] = sTf[((size_t)mf * nksmax + ks) * 32 + lane];
                    ta[mf] = tf[mf] * vk;
                }
                if (!AUGF && g == 7) ta[NFR - 1] = mk;        // row FB - 1 of the last fragment is the augmented row
#pragma unroll
                for (int mi = 0; mi < NFR; ++mi)
#pragma unroll
                    for (int nf = 0; nf <= mi; ++nf) dmma(acc[mi][nf], ta[mi], tf[nf]);
                if (AUGF) {
                    const double aa = (g == 7) ? mk : 0.0;
#pragma unroll
                    for (int nf = 0; nf < NFR; ++nf) dmma(aug[nf], aa, tf[nf]);
                }
            }
            // ---- 2. fragments -> row-major tile of this warp
#pragma unroll
            for (int mi = 0; mi < NFR; ++mi)
#pragma unroll
                for (int nf = 0; nf <= mi; ++nf)
                    *reinterpret_cast<double2*>(sC + (8 * mi + g) * LDC + 8 * nf + 2 * t) = make_double2(acc[mi][nf][0], acc[mi][nf][1]);
            if (AUGF && g == 7) {
#pragma unroll
                for (int nf = 0; nf < NFR; ++nf)
                    *reinterpret_cast<double2*>(sC + AUGROW * LDC + 8 * nf + 2 * t) = make_double2(aug[nf][0], aug[nf][1]);
            }
        }
        __syncwarp();
        double a[FB];
        {
            const double2* crow = reinterpret_cast<const double2*>(sC + (is_aug ? AUGROW : (lane < FB ? lane : 0)) * LDC);
            if (set_id == nullptr) {
                const double2* init = reinterpret_cast<const double2*>(sA + lane * LDC);
#pragma unroll
                for (int c = 0; c < FB; c += 2) {
                    const double2 x = init[c >> 1], y = crow[c >> 1];
                    a[c] = x.x + y.x;                          // lanes that hold no row compute on row 0 / their (unused) sA row:
                    a[c + 1] = x.y + y.y;                      // nothing of theirs is ever read by another lane
                }
            } else {
                // several experimental data sets on one image (closure tests, steer_analysis.py:168-183): A_b and e_b of this
                // walker's set come straight from global memory (L2 resident), everything else is the same arithmetic
                const long long sid = set_id[w];
                const double* As = Aall + sid * a_stride + B.a_off + (size_t)(lane < f ? lane : 0) * f;
                const double* es = e + sid * e_stride + B.row0;
#pragma unroll
                for (int c = 0; c < FB; c += 2) {
                    const double2 y = crow[c >> 1];
                    double x0 = (lane == c) ? 1.0 : 0.0, x1 = (lane == c + 1) ? 1.0 : 0.0;
                    if (c < f) {
                        if (is_aug) x0 = __ldg(es + c);
                        else if (lane < f) x0 = __ldg(As + c);
                    } else if (is_aug) x0 = 0.0;
                    if (c + 1 < f) {
                        if (is_aug) x1 = __ldg(es + c + 1);
                        else if (lane < f) x1 = __ldg(As + c + 1);
                    } else if (is_aug) x1 = 0.0;
                    a[c] = x0 + y.x;
                    a[c + 1] = x1 + y.y;
                }
            }
        }
        // ---- 3. right-looking Cholesky in registers; lane 31 becomes z = L^-1 dy
        double mypiv = 1.0, quad = 0.0;
#pragma unroll
        for (int c = 0; c < NCOL; ++c) {
            double* cb = sCol + (c & 1) * 32;
            cb[lane] = a[c];                                  // column c of the current matrix, all rows
            __syncwarp();
            const double piv = cb[c];
            if (lane == c) mypiv = piv;
            const double rinv = rsqrt(piv);
            const double u = a[c] * (rinv * rinv);            // a[r][c] / pivot
            a[c] *= rinv;                                     // L[r][c]; on lane 31: z_c
            if (is_aug) quad = fma(a[c], a[c], quad);
#pragma unroll
            for (int e2 = 0; e2 < FB; e2 += 2) {             // columns in aligned pairs: one 16-byte broadcast load per two
                if (e2 + 1 <= c || e2 >= NCOL) continue;      // (all of this resolves at compile time: c, e2 are unrolled)
                const double2 l2 = *reinterpret_cast<const double2*>(cb + e2);
                if (e2 > c) a[e2] = fma(-u, l2.x, a[e2]);
                if (e2 + 1 < NCOL) a[e2 + 1] = fma(-u, l2.y, a[e2 + 1]);
            }
        }
        // first non-positive pivot (1-based), found after the fact: lane c kept pivot c, and everything behind a bad pivot is NaN
        const unsigned badmask = __ballot_sync(0xffffffffu, lane < f && !(mypiv > 0.0));
        const int bad = badmask ? __ffs(badmask) : 0;
        const double ld = warp_sum((lane < f) ? 0.5 * log(mypiv) : 0.0);
        quad = shfl_d(quad, AL);
        int last = 0;
        if (lane == 0) {
            part[w * nB + blockIdx.x] = bad ? __longlong_as_double(0x7ff8000000000000LL) : (-0.5 * quad - ld);
            bad_out[w * nB + blockIdx.x] = bad ? B.row0 + bad : 0;
            if (fuse.counter != nullptr) {
                __threadfence();                                            // this block's term is visible before the count
                last = atomicAdd(fuse.counter + w, 1) == nB - 1;
            }
        }
        // the warp that delivers a walker's last block term adds the terms up, in block order (deterministic whichever warp
        // that is), applies the strict prior box and re-arms the counter for the next launch: no combine kernel
        if (__shfl_sync(0xffffffffu, last, 0)) {
            if (lane == 0) {
                __threadfence();
                bool inside = true;
                for (int k = 0; k < fuse.d; ++k) {
                    const double x = fuse.theta[w * fuse.d + k];
                    if (!(x > fuse.lo[k] && x < fuse.hi[k])) inside = false;
                }
                double sum = 0.0;
                int first_bad = 0;
                for (int b = 0; b < nB; ++b) {
                    sum += __ldcg(part + w * nB + b);                       // L2: the terms come from other SMs
                    const int bb = __ldcg(bad_out + w * nB + b);
                    if (bb != 0 && first_bad == 0) first_bad = bb;
                }
                fuse.logp[w] = inside ? sum : -INFINITY;
                if (fuse.info) fuse.info[w] = inside ? first_bad : -1;
                fuse.counter[w] = 0;
            }
        }
    }
}

// logp[w] = sum_b part[w][b] in block order; -inf outside the strict prior box; info = first non-positive pivot (1-based row)
__global__ void __launch_bounds__(128) block_combine_kernel(int nB, long long W, int d, const double* __restrict__ theta,
                                                            const double* __restrict__ lo, const double* __restrict__ hi,
                                                            const double* __restrict__ part, const int* __restrict__ bad_in,
                                                            double* __restrict__ logp, int* __restrict__ info) {
    pdl_wait();
    pdl_trigger();
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
