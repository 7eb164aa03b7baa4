// chol_pipe.cuh -- K4 (v2): fused covariance assembly + blocked Cholesky + solve + log-det, warp-specialised and fed by
// TMA bulk copies through an mbarrier ring.  Same mathematics as chol_loglik.cuh (see there); what changes is how
// operands reach the FP64 tensor pipe:
//
//   warps 0..7  consumers: one 16-row x 32-column strip ("unit") of the panel each per round, DMMA from shared memory
//   warp  8     producer : per 16-wide K half-chunk one cp.async.bulk per active consumer for its 16x16 A piece and
//                          two for the panel's 32x16 B piece (shared by all consumers), all completing on the stage's
//                          `full` mbarrier; 4 stages of 20 KB
//   warp  9     factor   : potrf + inverse of the 32x32 diagonal tile in registers while the consumers keep updating
//
// All consumers sweep K in lock step within a round (they share the B piece), the producer runs ahead across rounds
// and panels, and the only cross-panel dependency -- tiles of block column j-1 written by panel j-1's triangular
// solves -- is the LAST chunk of every sweep, so the ring rarely drains at a panel boundary.  A strip is solved against
// L_jj^-1 straight out of registers when the inverse is already there, otherwise parked in its own slot of the factor
// and solved at the end of the panel.
//
// Tile layout in the factor scratch and in the padded T tiles ("OF2"): [half h][ks][mi][lane],
//   idx = (((h*8 + ks)*2 + mi)*32 + lane),  value = tile[h*16 + mi*8 + g][ks*4 + t]
// so a 16-row strip is 4 KB contiguous and each of its K halves 2 KB contiguous (one bulk copy each).
//
// Generic-proxy stores of factor tiles are made visible to the async proxy (the bulk copies of later panels) by
// fence.proxy.async before the arrive on `panel_done`.
#pragma once
#include "chol_loglik.cuh"

namespace bi {

#ifndef PIPE_PREFETCH
#define PIPE_PREFETCH 0   // L2 prefetch distance of the producer in 32-wide K chunks; measured: no gain (the ring is not DRAM-latency bound)
#endif
#ifndef PIPE_XBYTES
#define PIPE_XBYTES (SUB_ELEMS * 8)   // experiment knob: bytes actually copied per A piece
#endif
// Consumer warps per CTA is a template parameter of the kernel: 8 (two CTAs per SM, the throughput configuration) or 16 (one
// CTA per SM: a panel needs about two thirds of the lock-step rounds, used for the walkers of a last, less than half filled
// wave and for launches of at most one walker per SM, where latency is what counts).
constexpr int PIPE_CONSUMERS_MAIN = 8;
constexpr int PIPE_CONSUMERS_WIDE = 16;
// 4 consumers, three CTAs per SM: for few-tile problems (F <= ~320), where a panel has at most a dozen strips and the
// walker's time is the chain potrf -> solve -> next diagonal tile, a third walker in flight per SM is worth more than wider
// rounds (measured on the C1 shape, profiles/r02_notes.md).
#ifndef PIPE_NC_SMALL
#define PIPE_NC_SMALL 4   // experiment knob: 5 consumers x 3 CTAs needs -DPIPE_NSTAGES=3 to fit the shared memory
#endif
constexpr int PIPE_CONSUMERS_SMALL = PIPE_NC_SMALL;
__host__ __device__ constexpr int pipe_threads(int nc) { return (nc + 2) * 32; }   // + producer + factor warp
#ifndef PIPE_NSTAGES
#define PIPE_NSTAGES 4   // measured: 3, 4, 6 and 8 stages all leave the consumers' wait at 25-30 % -- the ring is not latency bound
#endif
constexpr int PIPE_STAGES = PIPE_NSTAGES;
constexpr int STRIP_ELEMS = TILE_ELEMS / 2;                // 16 x 32 doubles = 4 KB
constexpr int SUB_ELEMS = STRIP_ELEMS / 2;                 // 16 x 16 doubles = 2 KB: one K half of a strip
__host__ __device__ constexpr int pipe_stage_elems(int nc) { return (2 + nc) * SUB_ELEMS; }   // B piece (2 strips) + one A piece per consumer: 20 KB at 8

__host__ __device__ __forceinline__ int of2_index(int r, int c) {
    return ((((r >> 4) * 8 + (c >> 2)) * 2 + ((r >> 3) & 1)) << 5) + ((r & 7) << 2) + (c & 3);
}

__host__ __device__ inline size_t pipe_smem_bytes(int Fp, int K0p, int nc = PIPE_CONSUMERS_MAIN) {
    size_t b = 0;
    b += sizeof(double) * (size_t)PIPE_STAGES * pipe_stage_elems(nc);
    b += sizeof(double) * TILE_ELEMS;          // sLinv
    b += sizeof(double) * 32 * 33;             // sD
    b += sizeof(double) * 32;                  // sRinv
    b += sizeof(double) * 256;                 // sLinvB
    b += sizeof(double) * (size_t)Fp;          // sDy
    b += sizeof(double) * (size_t)K0p * 2;     // sNegV, sM
    b += sizeof(double) * 16;                  // sRed (8) + logdet + spare
    b += sizeof(unsigned long long) * 24;      // mbarriers
    b += sizeof(int) * (size_t)Fp;             // sObs
    b += sizeof(int) * 8;                      // flags
    b += sizeof(int) * 4 * (size_t)(Fp / 32);  // kmin, imax, tile_omin, tile_omax
    return b + 128;                            // alignment slack
}

// ---- mbarrier / bulk-copy primitives ------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try(unsigned long long* bar, uint32_t parity) {   // may suspend for a while
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ bool mbar_test(unsigned long long* bar, uint32_t parity) {  // never blocks
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
    while (!mbar_try(bar, parity)) {
    }
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bulk_prefetch_l2(const void* src, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

struct PipeParams {
    CholParams c;
    const double* TmP;   // T tiles padded to whole 32-wide K chunks: [nt][nTc][1024], OF2 layout
    int nTc;             // number of 32-wide chunks of T (ceil(K0 / 32)); K0p = 32 * nTc
    long long* dbg;      // developer timing buffer (PIPE_TIMING builds only), else nullptr
};

#ifdef PIPE_TIMING
#define TSTART() long long t0_ = clock64()
#define TADD(cat) do { long long t1_ = clock64(); tacc[cat] += t1_ - t0_; t0_ = t1_; } while (0)
#else
#define TSTART()
#define TADD(cat)
#endif

// mbarrier slots
enum { BAR_FULL = 0, BAR_EMPTY = PIPE_STAGES, BAR_DIAG = 2 * PIPE_STAGES, BAR_LINV, BAR_PANEL, BAR_COUNT };

// 32 DMMAs of one K half-chunk: A piece [ks4][mi][lane], B piece [h][ks4][mi][lane]
template <bool SCALE>
__device__ __forceinline__ void half_chunk_mma(double (&acc)[2][4][2], const double* As, const double* Bs, const double* nv4) {
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
        double a[2], b[4];
#pragma unroll
        for (int mi = 0; mi < 2; ++mi) a[mi] = As[(ks * 2 + mi) * 32];
#pragma unroll
        for (int nf = 0; nf < 4; ++nf) {
            b[nf] = Bs[(((nf >> 1) * 4 + ks) * 2 + (nf & 1)) * 32];
            if (SCALE) b[nf] *= nv4[4 * ks];
        }
#pragma unroll
        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
            for (int nf = 0; nf < 4; ++nf) dmma(acc[mi][nf], a[mi], b[nf]);
    }
}
// X = C L_jj^-T for one 16 x 32 strip held as operand fragments a[mi][ks]; sLinv is [nf][ks][lane]
__device__ __forceinline__ void strip_trsm(const double (&a)[2][8], double (&x)[2][4][2], const double* sLinvLane) {
#pragma unroll
    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
        for (int nf = 0; nf < 4; ++nf) x[mi][nf][0] = x[mi][nf][1] = 0.0;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks)
#pragma unroll
        for (int nf = 0; nf < 4; ++nf)
            if (ks <= 2 * nf + 1) {   // L^-1 is lower triangular
                const double b = sLinvLane[(nf * 8 + ks) * 32];
#pragma unroll
                for (int mi = 0; mi < 2; ++mi) dmma(x[mi][nf], a[mi][ks], b);
            }
}

// ---- diagonal tile: blocked factorisation by one warp ------------------------------------------------------------------
// The FP64 tensor pipe is shared with the consumers' DMMAs, so what counts here is the NUMBER of FP64 instructions the
// factor warp issues, not only their dependency chain.  The 32 x 32 tile is factorised in 4 block columns of 8:
// lane r holds row r of the 32 x 8 panel, 8 right-looking scalar steps factor the diagonal block and solve the panel
// below it at once, and the trailing update L_panel L_panel^T goes through DMMA.  No 32 x 32 inverse is formed: the
// consumers' triangular solve is blocked too and only needs L_jj and the four 8 x 8 inverses of its diagonal blocks.
//   sD     32 x 33 row-major: in C_jj (lower), out L_jj (lower)
//   sLjj   [nf][ks][lane] operand-fragment-major L_jj (B operand of the blocked solve), upper part zero
//   sLinvB [b][kk][lane]: (L_bb)^-1 of the four 8 x 8 diagonal blocks as B fragments
// returns 0 or the 1-based index of the first non-positive pivot; mypiv = this lane's pivot L_ii^2 (the caller sums
// 0.5 log of it AFTER it has released the consumers: the logarithm is not on anybody's critical path).
// aug >= 0: row / column `aug` of this tile is the augmented dY row (see the kernel): its "pivot" is -|L^-1 dY|^2, which is
// returned in quad and replaced by 1 so that the (identity) padding rows behind it factorise trivially.
#ifdef PIPE_TIMING
__device__ unsigned long long potrf_dbg[8];
#define PT(i) do { long long t1p = clock64(); if (lane == 0) atomicAdd(&potrf_dbg[i], (unsigned long long)(t1p - t0p)); t0p = t1p; } while (0)
#else
#define PT(i)
#endif
__device__ __forceinline__ int warp_potrf_blocked(double* sD, double* sLjj, double* sLinvB, double* sRinv, int lane, double& mypiv,
                                                  int aug, double& quad) {
    const int g = lane >> 2, t = lane & 3;
    int bad = 0;
    mypiv = 1.0;
#ifdef PIPE_TIMING
    long long t0p = clock64();
#endif
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        double a[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) a[c] = sD[lane * 33 + 8 * b + c];
        PT(0);
        // Two columns at a time.  For the 2 x 2 pivot block [[A, B], [B, C]] the second pivot is det / A with
        // det = A C - B^2, so 1 / sqrt(A) and 1 / sqrt(det) are independent and the chain carries ONE reciprocal square
        // root per pair of columns instead of two (same cancellation as C - B^2 / A; the chain is what the consumers wait for).
#pragma unroll
        for (int c = 0; c < 8; c += 2) {
            const int k0 = 8 * b + c, k1 = k0 + 1;
            double A = shfl_d(a[c], k0);
            const double B = shfl_d(a[c], k1);
            const double Cc = shfl_d(a[c + 1], k1);
            if (k0 == aug) {                    // the augmented dY row: its would-be pivot is -|L^-1 dY|^2
                quad = -A;
                A = 1.0;
                if (lane == aug) a[c] = 1.0;
            }
            const double det = fma(A, Cc, -(B * B));
            double r1 = rsqrt(A);
            double rd = rsqrt(det);
            if (!(A > 0.0) && bad == 0) bad = k0 + 1;
            double p2 = det * (r1 * r1);        // second pivot
            double r2 = rd * (A * r1);          // 1 / sqrt(det / A)
            if (k1 == aug) {
                quad = -p2;
                p2 = 1.0;
                r2 = 1.0;
            }
            if (!(p2 > 0.0) && bad == 0) bad = k1 + 1;
            if (lane == k0) mypiv = A;
            if (lane == k1) mypiv = p2;
            const double l21 = B * r1;
            a[c] *= r1;                         // L[r][k0]; the pivot row: sqrt(A)
            a[c + 1] = fma(-a[c], l21, a[c + 1]) * r2;
            if (k1 == aug && lane == aug) a[c + 1] = 1.0;
            if (lane == k0) sRinv[k0] = r1;
            if (lane == k1) sRinv[k1] = r2;
#pragma unroll
            for (int c2 = c + 2; c2 < 8; ++c2) {
                const double la = shfl_d(a[c], 8 * b + c2), lb = shfl_d(a[c + 1], 8 * b + c2);
                a[c2] = fma(-a[c + 1], lb, fma(-a[c], la, a[c2]));
            }
        }
        PT(1);
#pragma unroll
        for (int c = 0; c < 8; ++c)
            if (lane >= 8 * b + c) sD[lane * 33 + 8 * b + c] = a[c];
        __syncwarp();
        // trailing update: T[8mf.., 8nf..] -= L[8mf.., 8b..8b+7] L[8nf.., 8b..8b+7]^T   for b < nf <= mf <= 3
#pragma unroll
        for (int mf = b + 1; mf < 4; ++mf) {
            double am[2];
#pragma unroll
            for (int kk = 0; kk < 2; ++kk) am[kk] = -sD[(8 * mf + g) * 33 + 8 * b + 4 * kk + t];
#pragma unroll
            for (int nf = b + 1; nf <= mf; ++nf) {
                double cacc[2];
                cacc[0] = sD[(8 * mf + g) * 33 + 8 * nf + 2 * t];
                cacc[1] = sD[(8 * mf + g) * 33 + 8 * nf + 2 * t + 1];
#pragma unroll
                for (int kk = 0; kk < 2; ++kk) dmma(cacc, am[kk], sD[(8 * nf + g) * 33 + 8 * b + 4 * kk + t]);
                sD[(8 * mf + g) * 33 + 8 * nf + 2 * t] = cacc[0];
                sD[(8 * mf + g) * 33 + 8 * nf + 2 * t + 1] = cacc[1];
            }
        }
        __syncwarp();
        PT(2);
    }
    if (bad) return bad;
    // (L_bb)^-1: lane 8b + c computes column c of block b by forward substitution on the identity
    {
        const int b = lane >> 3, c = lane & 7;
        double x[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            double s = 0.0;
#pragma unroll
            for (int k = 0; k < r; ++k) s = fma(sD[(8 * b + r) * 33 + 8 * b + k], x[k], s);
            x[r] = ((r == c ? 1.0 : 0.0) - s) * sRinv[8 * b + r];
        }
#pragma unroll
        for (int n = 0; n < 8; ++n) sLinvB[(b * 2 + (c >> 2)) * 32 + n * 4 + (c & 3)] = x[n];
    }
    PT(3);
    // L_jj as B operand
#pragma unroll 4
    for (int idx = lane; idx < TILE_ELEMS; idx += 32) {
        const int nf = idx >> 8, ks = (idx >> 5) & 7;
        const int n = 8 * nf + g, k = 4 * ks + t;
        sLjj[idx] = (k <= n) ? sD[n * 33 + k] : 0.0;
    }
    PT(4);
    return 0;
}

// one accumulator fragment (8 rows x 8 columns) -> its two operand fragments (k-slices of 4 columns)
__device__ __forceinline__ void cfrag_to_of2(const double (&c)[2], double& a0, double& a1, int lane) {
    const int base = lane & 28, hi = (lane & 3) >> 1;
    const double v00 = shfl_d(c[0], base | hi), v01 = shfl_d(c[1], base | hi);
    const double v10 = shfl_d(c[0], base | 2 | hi), v11 = shfl_d(c[1], base | 2 | hi);
    a0 = (lane & 1) ? v01 : v00;
    a1 = (lane & 1) ? v11 : v10;
}

// Blocked triangular solve X = C L_jj^-T of one 16 x 32 strip, in place: c[mi][nf] (accumulator fragments) -> X as
// accumulator fragments, xa[mi][ks] the same X as operand fragments.  sLjjLane / sLinvBLane already offset by lane.
__device__ __forceinline__ void strip_trsm_blocked(double (&c)[2][4][2], double (&xa)[2][8], const double* sLjjLane,
                                                   const double* sLinvBLane, int lane) {
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        // t = sum_{a<b} X_a L_{b,a}^T
        double tacc[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
#pragma unroll
        for (int ks = 0; ks < 2 * b; ++ks) {
            const double bl = sLjjLane[(b * 8 + ks) * 32];
#pragma unroll
            for (int mi = 0; mi < 2; ++mi) dmma(tacc[mi], xa[mi][ks], bl);
        }
        const double i0 = sLinvBLane[(b * 2) * 32], i1 = sLinvBLane[(b * 2 + 1) * 32];
#pragma unroll
        for (int mi = 0; mi < 2; ++mi) {
            double sfrag[2] = {c[mi][b][0] - tacc[mi][0], c[mi][b][1] - tacc[mi][1]};
            double s0, s1;
            cfrag_to_of2(sfrag, s0, s1, lane);
            double x[2] = {0.0, 0.0};
            dmma(x, s0, i0);
            dmma(x, s1, i1);
            c[mi][b][0] = x[0];
            c[mi][b][1] = x[1];
            cfrag_to_of2(x, xa[mi][2 * b], xa[mi][2 * b + 1], lane);
        }
    }
}

// The dY row rides inside the matrix: the factorised matrix is the (F+1) x (F+1) augmented
//     M = [ C   dY ]        chol(M) = [ L                     0 ]
//         [ dY' 0  ]                  [ (L^-1 dY)'   sqrt(-|L^-1 dY|^2) ]
// so row F of the factor is z = L^-1 dY and the would-be pivot of that row is -z'z: the quadratic form falls out of the
// last diagonal tile's factorisation (warp_potrf_blocked, aug) and costs no unit of its own -- the separate dY unit
// per panel of the first version cost 9 % of the kernel (one more lock-step round on every fourth panel).
// Row F lives in strip hA of the last row tile.  With structured skipping that strip is the only dense one: in panels
// that do not couple to the last tile it is swept as a single extra unit, and it always starts at the panel's own kmin.
template <int NC>
__global__ void __launch_bounds__(pipe_threads(NC), NC == PIPE_CONSUMERS_MAIN ? 2 : (NC == PIPE_CONSUMERS_SMALL ? 3 : 1))
chol_pipe_kernel(const PipeParams PP) {
    constexpr int PIPE_CONSUMERS = NC, PIPE_THREADS = pipe_threads(NC), STAGE_ELEMS = pipe_stage_elems(NC);
    const CholParams& P = PP.c;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    // 128-byte alignment as an offset into the shared array (rounding the generic address through uintptr_t made every
    // shared access of this kernel a generic LD.E / ST.E instead of LDS / STS)
    const unsigned smem_base = (unsigned)__cvta_generic_to_shared(smem_raw);
    double* sStage = reinterpret_cast<double*>(smem_raw + ((128u - (smem_base & 127u)) & 127u));
    double* sLinv = sStage + (size_t)PIPE_STAGES * STAGE_ELEMS;
    double* sD = sLinv + TILE_ELEMS;
    double* sRinv = sD + 32 * 33;
    double* sLinvB = sRinv + 32;
    double* sDy = sLinvB + 256;
    const int K0p = PP.nTc * 32;
    double* sNegV = sDy + P.Fp;
    double* sM = sNegV + K0p;
    double* sRed = sM + K0p;
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(sRed + 16);
    int* sObs = reinterpret_cast<int*>(bars + 24);
    int* sFlag = sObs + P.Fp;
    int* sKmin = sFlag + 8;
    int* sImax = sKmin + P.nt;
    int* sOmin = sImax + P.nt;
    int* sOmax = sOmin + P.nt;

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    const int g = lane >> 2;
    const int t = lane & 3;
    const int nt = P.nt;
    const int nTc = PP.nTc;
    const int ra = P.F - (nt - 1) * 32;   // augmented row inside the last tile
    const int hA = ra >> 4;               // ... its strip

    double* wsL = P.ws + (long long)blockIdx.x * P.ws_stride;

    if (tid == 0) {
        for (int s = 0; s < PIPE_STAGES; ++s) {
            mbar_init(&bars[BAR_FULL + s], 1);
            mbar_init(&bars[BAR_EMPTY + s], PIPE_CONSUMERS);
        }
        mbar_init(&bars[BAR_DIAG], 2);
        mbar_init(&bars[BAR_LINV], 1);
        mbar_init(&bars[BAR_PANEL], PIPE_CONSUMERS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int r = tid; r < P.Fp; r += PIPE_THREADS) sObs[r] = P.obs_id[r];
    for (int r = tid; r < P.nt; r += PIPE_THREADS) {
        sKmin[r] = P.kmin[r];
        sImax[r] = P.imax[r];
        sOmin[r] = P.tile_omin[r];
        sOmax[r] = P.tile_omax[r];
    }
    __syncthreads();

    uint32_t it = 0;       // K half-chunks issued / consumed so far (identical sequence in producer and consumers)
    uint32_t gpanel = 0;   // panels processed so far by this CTA (parity of the per-panel barriers)

    for (long long w = blockIdx.x; w < P.W; w += gridDim.x) {
        __syncthreads();   // previous walker fully retired
        if (tid == 0) {
            int inside = 1;
            if (P.check_prior) {
                for (int k = 0; k < P.d; ++k) {
                    const double x = P.theta[w * P.d + k];
                    if (!(x > P.prior_min[k] && x < P.prior_max[k])) inside = 0;
                }
            }
            sFlag[0] = inside;
            sFlag[1] = 0;
        }
        for (int q = tid; q < K0p; q += PIPE_THREADS) {
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
        const long long sid = P.set_id ? P.set_id[w] : 0;     // experimental data set of this walker
        const double* e_w = P.e + sid * P.e_stride;
        const double* At_w = P.Atiles + sid * P.a_stride;
        for (int r = tid; r < P.Fp; r += PIPE_THREADS) {
            double s = __ldg(e_w + r);
            for (int q = 0; q < P.K0; ++q) s = fma(__ldg(P.Tt + (long long)q * P.Fp + r), sM[q], s);
            sDy[r] = s;
        }
        __syncthreads();

        // unit q of panel j -> (row tile i, strip h): q < nreg regular, q == nreg the lone augmented strip
#define PIPE_UNIT(q, nreg, j, i, h)                      \
    const bool uaug_ = ((q) >= (nreg));                  \
    const int i = uaug_ ? nt - 1 : (j) + ((q) >> 1);     \
    const int h = uaug_ ? hA : ((q) & 1)

        if (warp == PIPE_CONSUMERS) {
            // =============================== producer ==========================================================
            // lane u < 8 owns the A piece of the round's u-th unit, lanes 8 and 9 the two halves of the shared B piece,
            // lane 0 arms the barrier
            uint32_t gp = gpanel;
            for (int j = 0; j < nt; ++j, ++gp) {
                const int i_hi = sImax[j];
                const int nreg = 2 * (i_hi - j + 1);
                const int nunits = nreg + (i_hi < nt - 1 ? 1 : 0);
                const int qA = (i_hi < nt - 1) ? nreg : 2 * (nt - 1 - j) + hA;   // the unit holding the augmented row
                const int kA = (j == nt - 1) ? 0 : sKmin[j];   // ... is dense: it starts where the panel does, and in the last
                                                               // panel (whose column F is the row itself) at 0
                const double* Bsrc = wsL + (size_t)j * (j - 1) / 2 * TILE_ELEMS;
                const int bh = lane - PIPE_CONSUMERS;   // 0 / 1 on the B lanes
                bool waited = false;
                for (int r0 = 0; r0 < nunits; r0 += PIPE_CONSUMERS) {
                    const int nact = min(PIPE_CONSUMERS, nunits - r0);
                    const bool first_aug = (r0 == nreg);          // the round is the lone augmented strip: no T chunks
                    const int i_first = first_aug ? nt - 1 : j + (r0 >> 1);
                    const int q = r0 + lane;
                    const bool has = lane < nact;
                    PIPE_UNIT(q, nreg, j, i, h);
                    const bool uz = has && uaug_;
                    const int my_kmin = has ? ((q == qA) ? kA : sKmin[i]) : 0x7fffffff;
                    const bool my_need0 = has && !uz && sOmax[j] >= sOmin[i];
                    // ---- T chunks (theta-dependent part of the covariance)
                    if (!first_aug && sOmax[j] >= sOmin[i_first]) {
                        const uint32_t bytes = (uint32_t)((2 + __popc(__ballot_sync(0xffffffffu, my_need0))) * SUB_ELEMS * 8);
                        const double* Asrc = PP.TmP + (size_t)i * nTc * TILE_ELEMS + h * STRIP_ELEMS;
                        const double* Bt = PP.TmP + (size_t)j * nTc * TILE_ELEMS;
                        for (int c2 = 0; c2 < 2 * nTc; ++c2, ++it) {
                            const int s = it % PIPE_STAGES;
                            mbar_wait(&bars[BAR_EMPTY + s], ((it / PIPE_STAGES) & 1) ^ 1);
                            double* st = sStage + (size_t)s * STAGE_ELEMS;
                            const size_t off = (size_t)(c2 >> 1) * TILE_ELEMS + (c2 & 1) * SUB_ELEMS;
                            if (lane == 0) mbar_expect_tx(&bars[BAR_FULL + s], bytes);
                            __syncwarp();
                            if (my_need0) bulk_g2s(st + (2 + lane) * SUB_ELEMS, Asrc + off, SUB_ELEMS * 8, &bars[BAR_FULL + s]);
                            if (bh == 0 || bh == 1)
                                bulk_g2s(st + bh * SUB_ELEMS, Bt + off + bh * STRIP_ELEMS, SUB_ELEMS * 8, &bars[BAR_FULL + s]);
                        }
                    }
                    // ---- factor chunks k < j
                    const bool a_in_round = (qA >= r0) && (qA < r0 + PIPE_CONSUMERS);
                    const int klo = a_in_round ? kA : sKmin[i_first];
                    const double* Asrc = wsL + (size_t)i * (i - 1) / 2 * TILE_ELEMS + h * STRIP_ELEMS;
                    for (int k = klo; k < j; ++k) {
                        if (k == j - 1 && !waited) {
                            // block column j-1 was written by panel j-1's solves: wait for all of them
                            mbar_wait(&bars[BAR_PANEL], (gp - 1) & 1);
                            waited = true;
                        }
                        const bool part = k >= my_kmin;
                        const uint32_t bytes = (uint32_t)(2 * SUB_ELEMS * 8 + __popc(__ballot_sync(0xffffffffu, part)) * PIPE_XBYTES);
                        for (int kh = 0; kh < 2; ++kh, ++it) {
                            const int s = it % PIPE_STAGES;
                            mbar_wait(&bars[BAR_EMPTY + s], ((it / PIPE_STAGES) & 1) ^ 1);
                            double* st = sStage + (size_t)s * STAGE_ELEMS;
                            const size_t off = (size_t)k * TILE_ELEMS + kh * SUB_ELEMS;
                            if (lane == 0) mbar_expect_tx(&bars[BAR_FULL + s], bytes);
                            __syncwarp();
                            if (part) bulk_g2s(st + (2 + lane) * SUB_ELEMS, Asrc + off, PIPE_XBYTES, &bars[BAR_FULL + s]);
                            if (bh == 0 || bh == 1)
                                bulk_g2s(st + bh * SUB_ELEMS, Bsrc + off + bh * STRIP_ELEMS, SUB_ELEMS * 8, &bars[BAR_FULL + s]);
                        }
                    }
                }
                // every phase of BAR_PANEL is observed exactly once, in order (parity waits must not skip a phase)
                if (!waited && gp > 0) mbar_wait(&bars[BAR_PANEL], (gp - 1) & 1);
            }
        } else if (warp == PIPE_CONSUMERS + 1) {
            // =============================== factor warp ========================================================
            double lsum = 0.0, quad = 0.0;     // lsum: this lane's share of sum log L_ii
            int first_bad = 0;
            uint32_t gp = gpanel;
#ifdef PIPE_TIMING
            long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#endif
            TSTART();
            for (int j = 0; j < nt; ++j, ++gp) {
                mbar_wait(&bars[BAR_DIAG], gp & 1);
                TADD(7);
                double mypiv;
                const int bad = warp_potrf_blocked(sD, sLinv, sLinvB, sRinv, lane, mypiv, j == nt - 1 ? ra : -1, quad);
                TADD(6);
                if (bad) {
                    if (first_bad == 0) first_bad = j * 32 + bad;
                    // keep the data flow alive (control flow must not depend on the data): poison the inverse
                    for (int n = 0; n < 32; ++n) sLinv[n * 32 + lane] = __longlong_as_double(0x7ff8000000000000LL);
                    for (int n = 0; n < 8; ++n) sLinvB[n * 32 + lane] = __longlong_as_double(0x7ff8000000000000LL);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&bars[BAR_LINV]);
                if (!bad) lsum += 0.5 * log(mypiv);
            }
            const double logdet = warp_sum(lsum);
            if (lane == 0) {
                sRed[8] = logdet;
                sRed[9] = quad;
                sFlag[1] = first_bad;
            }
#ifdef PIPE_TIMING
            if (PP.dbg && lane == 0) {
                atomicAdd((unsigned long long*)&PP.dbg[6], (unsigned long long)tacc[6]);
                atomicAdd((unsigned long long*)&PP.dbg[7], (unsigned long long)tacc[7]);
            }
#endif
        } else {
            // =============================== consumers ==========================================================
            uint32_t gp = gpanel;
#ifdef PIPE_TIMING
            long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#endif
            TSTART();
            for (int j = 0; j < nt; ++j, ++gp) {
                const int i_hi = sImax[j];
                const int nreg = 2 * (i_hi - j + 1);
                const int nunits = nreg + (i_hi < nt - 1 ? 1 : 0);
                const int qA = (i_hi < nt - 1) ? nreg : 2 * (nt - 1 - j) + hA;
                const int kA = (j == nt - 1) ? 0 : sKmin[j];
                const int omin_j = sOmin[j];
                const int omax_j = sOmax[j];
                uint32_t deferred = 0;
                int rnd = 0;
                for (int r0 = 0; r0 < nunits; r0 += PIPE_CONSUMERS, ++rnd) {
                    const int q = r0 + warp;
                    const bool active = q < nunits;
                    PIPE_UNIT(q, nreg, j, i, h);
                    const bool is_aug = active && uaug_;
                    const bool first_aug = (r0 == nreg);
                    const int i_first = first_aug ? nt - 1 : j + (r0 >> 1);
                    double acc[2][4][2];
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                        for (int nf = 0; nf < 4; ++nf) acc[mi][nf][0] = acc[mi][nf][1] = 0.0;

                    if (!first_aug && omax_j >= sOmin[i_first]) {
                        const bool need0 = active && !is_aug && omax_j >= sOmin[i];
                        for (int c2 = 0; c2 < 2 * nTc; ++c2, ++it) {
                            const int s = it % PIPE_STAGES;
                            mbar_wait(&bars[BAR_FULL + s], (it / PIPE_STAGES) & 1);
                            TADD(0);
                            if (need0) {
                                const double* Bs = sStage + (size_t)s * STAGE_ELEMS + lane;
                                half_chunk_mma<true>(acc, Bs + (2 + warp) * SUB_ELEMS, Bs, sNegV + c2 * 16 + t);
                            }
                            __syncwarp();
                            if (lane == 0) mbar_arrive(&bars[BAR_EMPTY + s]);
                            TADD(1);
                        }
                        if (need0 && omin_j != sOmax[i]) {
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
                    {
                        const bool a_in_round = (qA >= r0) && (qA < r0 + PIPE_CONSUMERS);
                        const int klo = a_in_round ? kA : sKmin[i_first];
                        const int my_klo = active ? ((q == qA) ? kA : sKmin[i]) : 0x7fffffff;
                        for (int k2 = 2 * klo; k2 < 2 * j; ++k2, ++it) {
                            const int s = it % PIPE_STAGES;
                            mbar_wait(&bars[BAR_FULL + s], (it / PIPE_STAGES) & 1);
                            TADD(0);
                            if ((k2 >> 1) >= my_klo) {
                                const double* Bs = sStage + (size_t)s * STAGE_ELEMS + lane;
                                half_chunk_mma<false>(acc, Bs + (2 + warp) * SUB_ELEMS, Bs, nullptr);
                            }
                            __syncwarp();
                            if (lane == 0) mbar_arrive(&bars[BAR_EMPTY + s]);
                            TADD(1);
                        }
                    }
                    if (!active) { TADD(2); continue; }
                    // ---- C = A_ij - acc   (+ dY in the augmented row)
                    {
                        const double2* At = reinterpret_cast<const double2*>(At_w + ((long long)i * (i + 1) / 2 + j) * TILE_ELEMS);
#pragma unroll
                        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                            for (int nf = 0; nf < 4; ++nf) {
                                const double2 a0 = __ldg(At + ((2 * h + mi) * 4 + nf) * 32 + lane);
                                acc[mi][nf][0] = a0.x - acc[mi][nf][0];
                                acc[mi][nf][1] = a0.y - acc[mi][nf][1];
                            }
                        if (q == qA && g == (ra & 7)) {
                            const int mA = (ra >> 3) & 1;
#pragma unroll
                            for (int mi = 0; mi < 2; ++mi)
                                if (mi == mA) {
#pragma unroll
                                    for (int nf = 0; nf < 4; ++nf)
#pragma unroll
                                        for (int e = 0; e < 2; ++e) acc[mi][nf][e] += sDy[j * 32 + nf * 8 + 2 * t + e];
                                }
                        }
                    }
                    if (i == j) {
#pragma unroll
                        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                            for (int nf = 0; nf < 4; ++nf)
#pragma unroll
                                for (int e = 0; e < 2; ++e) sD[(h * 16 + mi * 8 + g) * 33 + nf * 8 + 2 * t + e] = acc[mi][nf][e];
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&bars[BAR_DIAG]);
                        TADD(2);
                        continue;
                    }
                    double* Ct = wsL + ((long long)i * (i - 1) / 2 + j) * TILE_ELEMS + h * STRIP_ELEMS + lane;
                    if (!mbar_test(&bars[BAR_LINV], gp & 1)) {
                        // factor of the diagonal tile not there yet: park the strip, solve at the end of the panel
#pragma unroll
                        for (int mi = 0; mi < 2; ++mi) {
                            double a8[8];
                            cf_to_of(acc[mi], a8, lane);
#pragma unroll
                            for (int ks = 0; ks < 8; ++ks) Ct[(ks * 2 + mi) * 32] = a8[ks];
                        }
                        deferred |= 1u << rnd;
                        TADD(2);
                        continue;
                    }
                    {
                        double xa[2][8];
                        strip_trsm_blocked(acc, xa, sLinv + lane, sLinvB + lane, lane);
#pragma unroll
                        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                            for (int ks = 0; ks < 8; ++ks) Ct[(ks * 2 + mi) * 32] = xa[mi][ks];
                    }
                    TADD(3);
                }
                // ---- deferred strips of this panel (also: every warp observes every phase of BAR_LINV, in order)
                mbar_wait(&bars[BAR_LINV], gp & 1);
                TADD(4);
                for (int rr = 0; rr < rnd; ++rr) {
                    if (!((deferred >> rr) & 1)) continue;
                    const int q = rr * PIPE_CONSUMERS + warp;
                    PIPE_UNIT(q, nreg, j, i, h);
                    double* Ct = wsL + ((long long)i * (i - 1) / 2 + j) * TILE_ELEMS + h * STRIP_ELEMS + lane;
                    // the strip was parked as operand fragments; rebuild its accumulator fragments (inverse quad shuffle)
                    double x[2][4][2];
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                        for (int nf = 0; nf < 4; ++nf) {
                            const double a0 = Ct[((2 * nf) * 2 + mi) * 32], a1 = Ct[((2 * nf + 1) * 2 + mi) * 32];
                            // element (g, 2t+e) of the 8x8 block sits in operand fragment (2t+e)>>2 at lane (g, (2t+e)&3)
                            const int t_ = lane & 3, base = lane & 28;
                            const double p0 = shfl_d(a0, base | ((2 * t_) & 3)), p1 = shfl_d(a1, base | ((2 * t_) & 3));
                            const double q0 = shfl_d(a0, base | ((2 * t_ + 1) & 3)), q1 = shfl_d(a1, base | ((2 * t_ + 1) & 3));
                            x[mi][nf][0] = (t_ >= 2) ? p1 : p0;
                            x[mi][nf][1] = (t_ >= 2) ? q1 : q0;
                        }
                    double xa[2][8];
                    strip_trsm_blocked(x, xa, sLinv + lane, sLinvB + lane, lane);
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                        for (int ks = 0; ks < 8; ++ks) Ct[(ks * 2 + mi) * 32] = xa[mi][ks];
                }
                // this panel's tiles -> visible to the async proxy (bulk copies of later panels)
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(&bars[BAR_PANEL]);
                TADD(5);
            }
#ifdef PIPE_TIMING
            if (PP.dbg && lane == 0)
                for (int c = 0; c < 8; ++c) atomicAdd((unsigned long long*)&PP.dbg[warp * 8 + c], (unsigned long long)tacc[c]);
#endif
        }
#undef PIPE_UNIT
        gpanel += nt;
        __syncthreads();
        if (tid == 0) {
            const int bad = sFlag[1];
            if (bad) {
                P.logp[w] = __longlong_as_double(0x7ff8000000000000LL);
                if (P.info) P.info[w] = bad > P.F ? P.F : bad;
            } else {
                P.logp[w] = -0.5 * sRed[9] - sRed[8];
                if (P.info) P.info[w] = 0;
            }
        }
    }
}

}  // namespace bi
