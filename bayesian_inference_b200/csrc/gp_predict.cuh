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
// Only the lower triangle of L^-1 is multiplied, at the granularity of one 8-row fragment x one 4-column k-step.
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

#ifdef GP_TIMING
__device__ unsigned long long gp_dbg[8];   // developer timing: cycles of thread 0 per phase, summed over CTAs
#define GP_T(i) do { if (threadIdx.x == 0) { long long t1_ = clock64(); atomicAdd(&gp_dbg[i], (unsigned long long)(t1_ - t0_)); t0_ = t1_; } } while (0)
#else
#define GP_T(i)
#endif

// Walkers per CTA is a template parameter: 32 (256 threads, two CTAs per SM: one CTA's K1 overlaps the other's K2 on the
// shared FP64 pipe) or 64 (512 threads, one CTA per SM).  In-kernel clock64 timing on C3: K1 28.6 k + K2 23.6 k cycles
// per 32-walker CTA with a co-resident CTA, K1 18.8 k + K2 39.0 k per 64-walker CTA alone; 32 x 2 wins (0.124 vs 0.138 ms).
constexpr int GP_WT_MAX = 64;
// A third instantiation, 8 walkers per CTA (256 threads, 32 per walker in K1, one 8-row fragment per warp in K2), is the
// LATENCY configuration for small batches (the W/2 proposals of a sampler half-step): it spreads 50 walkers x 30 PCs over
// 210 CTAs instead of 60, and a CTA's K1 + K2 take a quarter of the time.  Same arithmetic per element and the same
// summation order, hence bit-identical (m, v).  A fourth width, 16 walkers (256 threads, 16 per walker in K1, two m-fragments per
// warp in K2), exists only inside gp_predict_mixed_kernel below: the tiles of a partly empty last wave.
__host__ __device__ constexpr int gp_threads(int wt) { return wt >= 32 ? wt * 8 : 256; }

__host__ __device__ inline size_t gp_smem_bytes(int d, int Nmax, int GP_WT) {
    const int nks = ((Nmax + 15) / 16) * 4;   // K steps, padded to whole groups of 4
    size_t b = 0;
    b += sizeof(double) * (size_t)GP_WT * 4 * nks;   // sA  (32 walkers x 4*nks), operand-fragment-major
    b += sizeof(double) * (size_t)d * Nmax;          // sX
    b += sizeof(double) * (size_t)GP_WT * d;         // sTh
    b += sizeof(double) * GP_WT * 8;                 // sSq  [wn][row] (up to 8 owners)
    b += sizeof(double) * GP_WT * 8;                 // sMn  [wn][row]
    return b;
}

// ---- K1 arithmetic ------------------------------------------------------------------------------------------------
// The CUDA math library's sqrt() / exp() cost ~150 issued instructions per kernel value here (slow-path branches,
// BSSY/BSYNC pairs and two UMOVs per polynomial coefficient), and every FP64 instruction K1 issues competes with K2's
// DMMAs for the one FP64 pipe of the SM sub-partition.  Both functions are restated branch-free for the only argument
// range that occurs (sq >= 0 finite, exp of a non-positive number) with as few FP64 instructions as the accuracy allows:
//   exp(-x) = 2^q * 2^(j/64) * e^r,  n = rint(-x 64 / ln2) = 64 q + j,  |r| <= ln2 / 128: a 64-entry table (one L1 load,
//             off the FP64 pipe) and a degree-5 Taylor polynomial (remainder 3.4e-17) -- 12 FP64 instructions, <= 2 ulp;
//   sqrt(s)  = MUFU.RSQ64H seed (2^-20) + two coupled Goldschmidt steps (2^-80: rounding limited, <= 2 ulp) -- 9.
// Both agree with libm on the test fixtures to <= 4e-16 relative, far inside the 1e-13 tolerance of the kernel-vector
// parity test.
__constant__ double GP_EXP_C[12] = {
    92.33248261689366,                   // 0  64 / ln2
    6.93147180369123816490e-01 / 64.0,   // 1  ln2_hi / 64 (32 significant bits: n * hi is exact)
    1.90821492927058770002e-10 / 64.0,   // 2  ln2_lo / 64
    6755399441055744.0,                  // 3  1.5 * 2^52: adding it rounds to nearest integer and leaves it in the low word
    1.0 / 120.0,                         // 4
    1.0 / 24.0,                          // 5
    1.0 / 6.0,                           // 6
    700.0,                               // 7  clamp: exp(-700) = 1e-304 is still a normal number
    0.0, 0.0, 0.0, 0.0};
__device__ const double GP_EXP_T[64] = {   // 2^(j/64), correctly rounded
    1.0, 1.0108892860517005, 1.0218971486541166, 1.0330248790212284,
    1.0442737824274138, 1.0556451783605572, 1.0671404006768237, 1.0787607977571199,
    1.0905077326652577, 1.102382583307841, 1.1143867425958924, 1.1265216186082418,
    1.1387886347566916, 1.1511892299529827, 1.1637248587775775, 1.1763969916502812,
    1.189207115002721, 1.202156731452703, 1.215247359980469, 1.22848053610687,
    1.241857812073484, 1.255380757024691, 1.2690509571917332, 1.2828700160787783,
    1.2968395546510096, 1.3109612115247644, 1.3252366431597413, 1.339667524053303,
    1.3542555469368927, 1.3690024229745905, 1.383909881963832, 1.3989796725383112,
    1.4142135623730951, 1.42961333839197, 1.4451808069770467, 1.460917794180647,
    1.4768261459394993, 1.4929077282912648, 1.5091644275934228, 1.5255981507445384,
    1.5422108254079407, 1.559004400237837, 1.5759808451078865, 1.593142151342267,
    1.6104903319492543, 1.6280274218573478, 1.645755478153965, 1.6636765803267364,
    1.681792830507429, 1.7001063537185235, 1.718619298122478, 1.7373338352737062,
    1.7562521603732995, 1.7753764925265212, 1.7947090750031072, 1.8142521755003989,
    1.8340080864093424, 1.8539791250833855, 1.8741676341103, 1.8945759815869656,
    1.9152065613971474, 1.9360617934922943, 1.9571441241754002, 1.978456026387951};

__device__ __forceinline__ double exp_neg(double x) {   // exp(-x), x >= 0
    x = fmin(x, GP_EXP_C[7]);
    const double tm = fma(x, -GP_EXP_C[0], GP_EXP_C[3]);
    const double nf = tm - GP_EXP_C[3];                 // n = rint(-x 64 / ln2) <= 0
    double r = fma(nf, -GP_EXP_C[1], -x);
    r = fma(nf, -GP_EXP_C[2], r);                       // r = -x - n ln2 / 64, |r| <= ln2 / 128
    const int n = __double2loint(tm);
    const double tj = __ldg(GP_EXP_T + (n & 63));       // n = 64 q + j with j = n & 63, q = n >> 6 (floor)
    double p = fma(GP_EXP_C[4], r, GP_EXP_C[5]);
    p = fma(p, r, GP_EXP_C[6]);
    p = fma(p, r, 0.5);
    p = fma(p, r, 1.0);
    p = fma(p, r, 1.0);
    return (tj * p) * __hiloint2double(((n >> 6) + 1023) << 20, 0);   // q >= -1010: the scale is a normal power of two
}

__device__ __forceinline__ double sqrt_pos(double s) {  // sqrt(s), s >= 0 finite; returns 1e-150 for s < 1e-300
    s = fmax(s, 1e-300);
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(s));   // MUFU.RSQ64H, ~2^-20 relative
    double g = s * y, h = 0.5 * y;                      // coupled (Goldschmidt) iteration: g -> sqrt(s), h -> 1/(2 sqrt(s))
#pragma unroll
    for (int it = 0; it < 2; ++it) {
        const double r = fma(-g, h, 0.5);
        g = fma(g, r, g);
        if (it == 0) h = fma(h, r, h);
    }
    return g;
}

template <int kind>
__device__ __forceinline__ double stationary_kernel(double sq) {
    // sq = squared scaled distance
    if (kind == 3) return exp_neg(0.5 * sq);
    const double r = sqrt_pos(sq);
    if (kind == 0) return exp_neg(r);
    if (kind == 1) {
        const double tt = r * 1.7320508075688772;     // math.sqrt(3) * dists
        const double e = exp_neg(tt);
        return fma(tt, e, e);                          // (1 + tt) exp(-tt)
    }
    const double tt = r * 2.23606797749979;            // math.sqrt(5) * dists
    return fma(tt * tt, 1.0 / 3.0, 1.0 + tt) * exp_neg(tt);
}

// K1 for one (walker tile, PC): thread = (walker tid/8, design points tid%8 + 8u + 32i), four design points at a time
// so that four independent sqrt/exp dependency chains interleave.
template <bool DO_GP, int KIND, int TPW>
__device__ __forceinline__ void gp_k1_loop(double* sA, const double* sX, const double* sTh, const PcDesc& pc, int d, int nks,
                                           long long w0, long long W, double* __restrict__ k_out, int k_npc, int tid) {
    const int N = pc.N;
    const int wl = tid / TPW, jl = tid % TPW;      // TPW threads per walker (8, or 32 in the 8-walker instantiation)
    const double* th = sTh + wl * d;
    double* krow = (k_out != nullptr && w0 + wl < W) ? k_out + ((w0 + wl) * k_npc + pc.local) * N : nullptr;
    // of_index(wl, jl + TPW u + 4 TPW i, nks) = base + 8 TPW u + 32 TPW i
    double* sArow = sA + ((((wl >> 3) * nks + (jl >> 2)) << 5) + ((wl & 7) << 2) + (jl & 3));
    const double cst = pc.cst;
    for (int jb = jl; jb < N; jb += 4 * TPW, sArow += 32 * TPW) {
        double sq[4] = {0.0, 0.0, 0.0, 0.0};
        const double* xj[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) xj[u] = sX + min(jb + TPW * u, N - 1);
        for (int k = 0; k < d; ++k) {
            const double tk = th[k];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const double df = tk - xj[u][k * N];
                sq[u] = fma(df, df, sq[u]);
            }
        }
        double kv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) kv[u] = stationary_kernel<KIND>(sq[u]) + cst;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (jb + TPW * u < N) {
                if (DO_GP) sArow[8 * TPW * u] = kv[u];
                if (krow != nullptr) krow[jb + TPW * u] = kv[u];
            }
        }
    }
}

// K2 for one (walker tile, PC): the calling warps (8 per 32 walkers) contract the kernel vectors in sA with L^-1 and leave
// their partial sums of squares and means in sSq / sMn ([owner][walker]).
//
// Warp tile = (8 MI walker rows) x (NF fragments of 8 rows of L^-1), MI = 8 / NF: 16 accumulator doubles either way.
// NF = 2 (MI = 4, eight fragment owners per 32 walkers) is the default: every B fragment is loaded by exactly one warp of
// the CTA and the B double buffer is 32 registers instead of 64 -- with NF = 4 the loop ran out of its 128 registers
// (spills and re-materialised shared-memory addresses inside the k loop: ncu source page, profiles/r01_notes.md).
#ifndef GP_K2_NF
#define GP_K2_NF 2
#endif
constexpr int GP_K2_OWNERS = 16 / GP_K2_NF;   // fragment-group owners per 32 walkers = rows of sSq / sMn
#ifndef GP_K2_DEEP
#define GP_K2_DEEP 3                          // narrow tiles: groups of k-steps in flight ahead of the DMMAs (0 = shallow loop)
#endif

template <int GP_WT>
__device__ __forceinline__ void gp_k2_tile(const PcDesc& pc, const double* sA, int warp, int lane, double* sSq, double* sMn) {
    constexpr int NF = GP_K2_NF, NO = GP_K2_OWNERS;
    constexpr int MI = (GP_WT >= 32) ? 8 / NF : GP_WT / 8;   // m-fragments (8 walkers each) per warp: 4, or 1 in the 8-walker tile
    constexpr int WM = GP_WT / (8 * MI);    // warps along the walker rows
    const int N = pc.N, nks = pc.nks;
    // warp (wm, wn): walker rows wm*8*MI.., fragment groups {q : snake(q) == wn}.  The B operand is cut into n-fragments
    // of 8 rows of L^-1 (fragment f = rows 8f .. 8f+7; row N holds alpha).  Fragment f only needs the k-steps
    // 0 .. min(ceil(N/4), 2f+2) - 1 (L^-1 is lower triangular) and fragments past row N do not exist, so the fragments
    // are grouped NF at a time from the TOP (group 0 = the longest) and the groups dealt to the owners in snake order:
    // every owner gets the same number of DMMAs to within a few per cent.  B (L2 resident, shared by every CTA of this
    // PC) is streamed global -> registers in groups of 4 k-steps, one group ahead of the DMMAs that consume it.
    const int wm = warp % WM;
    const int wn = warp / WM;   // fragment-group owner 0 .. NO-1
    const int g = lane >> 2, t = lane & 3;
    double sumsq[MI], meanv[MI];
#pragma unroll
    for (int mi = 0; mi < MI; ++mi) sumsq[mi] = meanv[mi] = 0.0;
    const double* sAw = sA + (size_t)(MI * wm) * nks * 32 + lane;
    const double* Bl = pc.LinvF + lane;
    const int kfull = (N + 3) >> 2;
    const int ftop = N >> 3;                        // fragment holding row N
    const int ngroups = (ftop + NF) / NF;

    auto group_of = [&](int idx) {   // idx-th fragment group owned by this warp
        const int blk = idx >> 1;
        return blk * 2 * NO + ((idx & 1) ? 2 * NO - 1 - wn : wn);
    };
    auto klim_of = [&](int f) { return f < 0 ? 0 : min(kfull, 2 * f + 2); };
    // loads are never predicated: past a fragment's triangle the packed operand holds zeros (4*grp+3 < nks always),
    // and the fragments a bottom group lacks are clamped to fragment 0 and never multiplied
    auto load_group = [&](double (&bb)[4][NF], int f0, int grp) {
#pragma unroll
        for (int nf = 0; nf < NF; ++nf) {
            const double* bp = Bl + ((size_t)max(f0 - nf, 0) * nks + 4 * grp) * 32;
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) bb[kk][nf] = __ldg(bp + kk * 32);
        }
    };

    int idx = 0;
    int q = group_of(0);
    if (q < ngroups) {
        double bcur[4][NF], bnxt[4][NF];
        load_group(bcur, ftop - NF * q, 0);
        while (q < ngroups) {
            const int f0 = ftop - NF * q;
            const int ngrp = (klim_of(f0) + 3) >> 2;
            const int kl_last = klim_of(f0 - (NF - 1));       // the shortest fragment of the group
            double acc[MI][NF][2];
#pragma unroll
            for (int mi = 0; mi < MI; ++mi)
#pragma unroll
                for (int nf = 0; nf < NF; ++nf) acc[mi][nf][0] = acc[mi][nf][1] = 0.0;
            const int next_q = group_of(idx + 1);
            for (int grp = 0; grp < ngrp; ++grp) {
                // prefetch the next group of k-steps (of this fragment group, or the first of the next one)
                if (grp + 1 < ngrp) load_group(bnxt, f0, grp + 1);
                else if (next_q < ngroups) load_group(bnxt, ftop - NF * next_q, 0);
                if (4 * grp + 4 <= kl_last) {
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) {
                        double a[MI];
#pragma unroll
                        for (int mi = 0; mi < MI; ++mi) a[mi] = sAw[(mi * nks + grp * 4 + kk) * 32];
#pragma unroll
                        for (int mi = 0; mi < MI; ++mi)
#pragma unroll
                            for (int nf = 0; nf < NF; ++nf) dmma(acc[mi][nf], a[mi], bcur[kk][nf]);
                    }
                } else {
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) {
                        const int ks = grp * 4 + kk;
                        double a[MI];
#pragma unroll
                        for (int mi = 0; mi < MI; ++mi) a[mi] = sAw[(mi * nks + ks) * 32];
#pragma unroll
                        for (int nf = 0; nf < NF; ++nf)
                            if (ks < klim_of(f0 - nf)) {
#pragma unroll
                                for (int mi = 0; mi < MI; ++mi) dmma(acc[mi][nf], a[mi], bcur[kk][nf]);
                            }
                    }
                }
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)
#pragma unroll
                    for (int nf = 0; nf < NF; ++nf) bcur[kk][nf] = bnxt[kk][nf];
            }
            if (f0 == ftop) {
                // the top fragment holds row N (= alpha: the mean) and nothing past it counts
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int n = f0 * 8 + 2 * t + e;
#pragma unroll
                    for (int mi = 0; mi < MI; ++mi) {
                        const double x = acc[mi][0][e];
                        if (n < N) sumsq[mi] = fma(x, x, sumsq[mi]);
                        else if (n == N) meanv[mi] = x;
                    }
                }
            }
#pragma unroll
            for (int nf = 0; nf < NF; ++nf) {
                if (f0 - nf < 0 || (nf == 0 && f0 == ftop)) continue;
#pragma unroll
                for (int mi = 0; mi < MI; ++mi)
                    sumsq[mi] = fma(acc[mi][nf][0], acc[mi][nf][0], fma(acc[mi][nf][1], acc[mi][nf][1], sumsq[mi]));
            }
            ++idx;
            q = next_q;
        }
    }
#pragma unroll
    for (int mi = 0; mi < MI; ++mi) {
        sumsq[mi] += __shfl_xor_sync(0xffffffffu, sumsq[mi], 1);
        sumsq[mi] += __shfl_xor_sync(0xffffffffu, sumsq[mi], 2);
        meanv[mi] += __shfl_xor_sync(0xffffffffu, meanv[mi], 1);   // only one lane of the quad holds column N
        meanv[mi] += __shfl_xor_sync(0xffffffffu, meanv[mi], 2);
        if (t == 0) {
            const int row = wm * 8 * MI + mi * 8 + g;
            sSq[wn * GP_WT + row] = sumsq[mi];
            sMn[wn * GP_WT + row] = meanv[mi];
        }
    }
}

// K2 for the narrow tiles (8 and 16 walkers per CTA): the same contraction, the same fragment groups per warp and the same
// order of DMMAs per accumulator as gp_k2_tile (hence the same bits), but the stream of L^-1 runs PF groups of 4 k-steps
// ahead of the DMMAs through a ring of PF + 1 register buffers instead of one group ahead.  A narrow tile issues 8 or 16
// DMMAs per group and warp -- 128 to 256 cycles of the FP64 pipe against an L2 round trip of ~340 (in-kernel clock64, C1,
// 50 walkers: K2 10.1 k cycles of a CTA's 16 k, i.e. one exposed round trip per group) -- so one group ahead leaves the
// loop latency bound; the wide tiles (32 DMMAs per group, 16 warps per SM) are pipe bound and keep the shallow version,
// whose 127 registers have no room for a deeper ring.
template <int GP_WT, int PF>
__device__ __forceinline__ void gp_k2_tile_deep(const PcDesc& pc, const double* sA, int warp, int lane, double* sSq, double* sMn) {
    constexpr int NF = GP_K2_NF, NO = GP_K2_OWNERS, R = PF + 1;
    constexpr int MI = GP_WT / 8;           // m-fragments (8 walkers each) per warp; every warp is a fragment-group owner
    static_assert(GP_WT == 8 || GP_WT == 16, "narrow tiles only");
    const int N = pc.N, nks = pc.nks;
    const int wn = warp;
    const int g = lane >> 2, t = lane & 3;
    double sumsq[MI], meanv[MI];
#pragma unroll
    for (int mi = 0; mi < MI; ++mi) sumsq[mi] = meanv[mi] = 0.0;
    const double* sAw = sA + lane;
    const double* Bl = pc.LinvF + lane;
    const int kfull = (N + 3) >> 2;
    const int ftop = N >> 3;
    const int ngroups = (ftop + NF) / NF;
    auto klim_of = [&](int f) { return f < 0 ? 0 : min(kfull, 2 * f + 2); };

    struct Cursor { int idx, f0, ngrp, grp; bool valid; };
    auto seek = [&](Cursor& c) {            // position c on the first k-step group of its idx-th fragment group
        const int blk = c.idx >> 1;
        const int q = blk * 2 * NO + ((c.idx & 1) ? 2 * NO - 1 - wn : wn);
        c.valid = q < ngroups;
        c.f0 = ftop - NF * q;
        c.ngrp = c.valid ? (klim_of(c.f0) + 3) >> 2 : 0;
        c.grp = 0;
    };
    auto advance = [&](Cursor& c) {
        if (++c.grp == c.ngrp) { ++c.idx; seek(c); }
    };
    auto load_group = [&](double (&bb)[4][NF], const Cursor& c) {
#pragma unroll
        for (int nf = 0; nf < NF; ++nf) {
            const double* bp = Bl + ((size_t)max(c.f0 - nf, 0) * nks + 4 * c.grp) * 32;
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) bb[kk][nf] = __ldg(bp + kk * 32);
        }
    };

    Cursor cc{0, 0, 0, 0, false}, cp{0, 0, 0, 0, false};
    seek(cc);
    cp = cc;
    double buf[R][4][NF];
#pragma unroll
    for (int s = 0; s < PF; ++s)
        if (cp.valid) { load_group(buf[s], cp); advance(cp); }
    double acc[MI][NF][2];
    while (cc.valid) {
#pragma unroll
        for (int s = 0; s < R; ++s) {
            if (!cc.valid) break;
            if (cp.valid) { load_group(buf[(s + PF) % R], cp); advance(cp); }
            const int f0 = cc.f0, grp = cc.grp;
            if (grp == 0) {
#pragma unroll
                for (int mi = 0; mi < MI; ++mi)
#pragma unroll
                    for (int nf = 0; nf < NF; ++nf) acc[mi][nf][0] = acc[mi][nf][1] = 0.0;
            }
            if (4 * grp + 4 <= klim_of(f0 - (NF - 1))) {
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                    double a[MI];
#pragma unroll
                    for (int mi = 0; mi < MI; ++mi) a[mi] = sAw[(mi * nks + grp * 4 + kk) * 32];
#pragma unroll
                    for (int mi = 0; mi < MI; ++mi)
#pragma unroll
                        for (int nf = 0; nf < NF; ++nf) dmma(acc[mi][nf], a[mi], buf[s][kk][nf]);
                }
            } else {
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                    const int ks = grp * 4 + kk;
                    double a[MI];
#pragma unroll
                    for (int mi = 0; mi < MI; ++mi) a[mi] = sAw[(mi * nks + ks) * 32];
#pragma unroll
                    for (int nf = 0; nf < NF; ++nf)
                        if (ks < klim_of(f0 - nf)) {
#pragma unroll
                            for (int mi = 0; mi < MI; ++mi) dmma(acc[mi][nf], a[mi], buf[s][kk][nf]);
                        }
                }
            }
            if (grp == cc.ngrp - 1) {          // the fragment group is complete: fold it into the sums
                if (f0 == ftop) {
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int n = f0 * 8 + 2 * t + e;
#pragma unroll
                        for (int mi = 0; mi < MI; ++mi) {
                            const double x = acc[mi][0][e];
                            if (n < N) sumsq[mi] = fma(x, x, sumsq[mi]);
                            else if (n == N) meanv[mi] = x;
                        }
                    }
                }
#pragma unroll
                for (int nf = 0; nf < NF; ++nf) {
                    if (f0 - nf < 0 || (nf == 0 && f0 == ftop)) continue;
#pragma unroll
                    for (int mi = 0; mi < MI; ++mi)
                        sumsq[mi] = fma(acc[mi][nf][0], acc[mi][nf][0], fma(acc[mi][nf][1], acc[mi][nf][1], sumsq[mi]));
                }
            }
            advance(cc);
        }
    }
#pragma unroll
    for (int mi = 0; mi < MI; ++mi) {
        sumsq[mi] += __shfl_xor_sync(0xffffffffu, sumsq[mi], 1);
        sumsq[mi] += __shfl_xor_sync(0xffffffffu, sumsq[mi], 2);
        meanv[mi] += __shfl_xor_sync(0xffffffffu, meanv[mi], 1);
        meanv[mi] += __shfl_xor_sync(0xffffffffu, meanv[mi], 2);
        if (t == 0) {
            const int row = mi * 8 + g;
            sSq[wn * GP_WT + row] = sumsq[mi];
            sMn[wn * GP_WT + row] = meanv[mi];
        }
    }
}

// var = diag - sum V^2 clipped at 0, std -> squared, mean: thread tid < GP_WT finishes walker w0 + tid
template <int GP_WT>
__device__ __forceinline__ void gp_finish_tile(const PcDesc& pc, const double* sSq, const double* sMn, int tid, long long w0, long long W,
                                               double* __restrict__ m_out, double* __restrict__ v_out, int Ptot, int col) {
    if (tid < GP_WT) {
        const long long w = w0 + tid;
        if (w < W) {
            double ss = 0.0, mu = 0.0;
#pragma unroll
            for (int o = 0; o < GP_K2_OWNERS; o += 2) {
                ss += sSq[o * GP_WT + tid] + sSq[(o + 1) * GP_WT + tid];
                mu += sMn[o * GP_WT + tid] + sMn[(o + 1) * GP_WT + tid];
            }
            double var = pc.diag - ss;
            if (var < 0.0) var = 0.0;                       // _gpr.py:485-491
            const double sd = sqrt(var);                    // _gpr.py:500 returns the std ...
            const long long o = w * Ptot + col;
            m_out[o] = mu;
            v_out[o] = sd * sd;                             // ... and the reference squares it, emulation.py:499
        }
    }
}

// One (walker tile, PC) unit: walkers w0 .. w0 + GP_WT - 1 against principal component `col` (an index into pcs).
// k_out (optional): (W, k_ld_pc, N) row-major slice for this group, PC index = desc.local.
// m_out / v_out (optional): (W, Ptot) at column col.
template <bool DO_GP, int GP_WT>
__device__ __forceinline__ void gp_predict_unit(unsigned char* smem_raw, const PcDesc* __restrict__ pcs, int col, int d,
                                                const double* __restrict__ theta, long long W, long long w0,
                                                double* __restrict__ k_out, int k_npc, double* __restrict__ m_out,
                                                double* __restrict__ v_out, int Ptot) {
    const PcDesc pc = pcs[col];
    const int N = pc.N, nks = pc.nks;
    double* sA = reinterpret_cast<double*>(smem_raw);
    double* sX = sA + (size_t)GP_WT * 4 * nks;
    double* sTh = sX + (size_t)d * N;
    double* sSq = sTh + GP_WT * d;
    double* sMn = sSq + GP_WT * 8;

    constexpr int GP_THREADS = gp_threads(GP_WT), TPW = GP_THREADS / GP_WT;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#ifdef GP_TIMING
    long long t0_ = clock64();
#endif

    for (int idx = tid; idx < d * N; idx += GP_THREADS) sX[idx] = __ldg(pc.XsT + idx);   // the image: never written after bi_create
    pdl_wait();        // theta may come from the previous kernel of the stream (the proposals of a sampler half-step)
    pdl_trigger();
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
    GP_T(0);

    // ---- K1 -------------------------------------------------------------------------------------------------
    switch (pc.kind) {   // the kernel kind is a template parameter of the loop so that its four chains interleave
        case 0: gp_k1_loop<DO_GP, 0, TPW>(sA, sX, sTh, pc, d, nks, w0, W, k_out, k_npc, tid); break;
        case 1: gp_k1_loop<DO_GP, 1, TPW>(sA, sX, sTh, pc, d, nks, w0, W, k_out, k_npc, tid); break;
        case 2: gp_k1_loop<DO_GP, 2, TPW>(sA, sX, sTh, pc, d, nks, w0, W, k_out, k_npc, tid); break;
        default: gp_k1_loop<DO_GP, 3, TPW>(sA, sX, sTh, pc, d, nks, w0, W, k_out, k_npc, tid); break;
    }
    if (!DO_GP) return;
    GP_T(1);
    __syncthreads();
    GP_T(2);

    // ---- K2 -------------------------------------------------------------------------------------------------
#if GP_K2_DEEP > 0
    if constexpr (GP_WT <= 16) gp_k2_tile_deep<GP_WT, GP_K2_DEEP>(pc, sA, warp, lane, sSq, sMn);
    else
#endif
    gp_k2_tile<GP_WT>(pc, sA, warp, lane, sSq, sMn);
    GP_T(3);
    __syncthreads();
    GP_T(4);
    gp_finish_tile<GP_WT>(pc, sSq, sMn, tid, w0, W, m_out, v_out, Ptot, col);
}

// grid: (ceil(W / GP_WT), n_pc_selected).  pcs[blockIdx.y + pc_first] describes the PC.
template <bool DO_GP, int GP_WT>
__global__ void __launch_bounds__(gp_threads(GP_WT), GP_WT <= 32 ? 2 : 1) gp_predict_kernel(const PcDesc* __restrict__ pcs, int pc_first, int d,
                                                                const double* __restrict__ theta, long long W,
                                                                double* __restrict__ k_out, int k_npc,
                                                                double* __restrict__ m_out, double* __restrict__ v_out,
                                                                int Ptot) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    gp_predict_unit<DO_GP, GP_WT>(smem_raw, pcs, pc_first + blockIdx.y, d, theta, W, (long long)blockIdx.x * GP_WT, k_out, k_npc,
                                  m_out, v_out, Ptot);
}

// Mixed tiles in ONE launch (the wave tail): 1024 walkers x 30 PCs in 32-walker tiles are 960 CTAs on 296 resident slots =
// 3.24 waves, i.e. a fourth wave that is a quarter full.  Here the first n_big walker tiles are 32 walkers wide and the
// remaining walkers go in tiles of GP_WT_TAIL = 16; CTAs are dealt in block order, big ones first (for each PC its tiles,
// so that the CTAs resident together share the PC's L^-1 in L2), then the small ones, which fill the slots the last big
// wave leaves free and halve the length of what runs after it.  Same arithmetic per (walker, PC) and the same summation
// order in every tile width, hence bit-identical (m, v) to gp_predict_kernel.
constexpr int GP_WT_TAIL = 16;
__global__ void __launch_bounds__(256, 2) gp_predict_mixed_kernel(const PcDesc* __restrict__ pcs, int n_pc, int d,
                                                                  const double* __restrict__ theta, long long W, int n_big,
                                                                  int n_small, double* __restrict__ m_out,
                                                                  double* __restrict__ v_out, int Ptot) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int big_units = n_big * n_pc;
    int u = (int)blockIdx.x;
    if (u < big_units) {
        const int col = u / n_big, tile = u - col * n_big;
        gp_predict_unit<true, 32>(smem_raw, pcs, col, d, theta, W, (long long)tile * 32, nullptr, 0, m_out, v_out, Ptot);
    } else {
        u -= big_units;
        const int col = u / n_small, tile = u - col * n_small;
        gp_predict_unit<true, GP_WT_TAIL>(smem_raw, pcs, col, d, theta, W, (long long)n_big * 32 + (long long)tile * GP_WT_TAIL, nullptr,
                                          0, m_out, v_out, Ptot);
    }
}

}  // namespace bi
