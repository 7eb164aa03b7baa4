// chol_pipe.cuh -- K4 (v2): fused covariance assembly + blocked Cholesky + solve + log-det, warp-specialised and fed by
// TMA bulk copies through an mbarrier ring.  Same mathematics and tile formats as chol_loglik.cuh (see there); what
// changes is how operands reach the FP64 tensor pipe:
//
//   warps 0..7  consumers: one 16-row x 32-column strip ("unit") of the panel each per round, DMMA from shared memory
//   warp  8     producer : per K-chunk one cp.async.bulk for the panel's 32x32 B tile (shared by all consumers) and one
//                          per active consumer for its 16x32 A strip, completing on the stage's `full` mbarrier
//   warp  9     factor   : potrf + inverse of the 32x32 diagonal tile in registers while the consumers keep updating
//
// All consumers sweep K in lock step within a round (they share the B tile), the producer runs ahead across rounds and
// panels, and the only cross-panel dependency -- tiles of block column j-1 written by panel j-1's triangular solves --
// is the LAST chunk of every sweep, so the ring never drains at a panel boundary.  A strip is solved against L_jj^-1
// straight out of registers when the inverse is already there, otherwise parked in its own slot of the factor and
// solved at the end of the panel.
//
// Generic-proxy stores of factor tiles are made visible to the async proxy (the bulk copies of later panels) by
// fence.proxy.async before the arrive on `panel_done`.
#pragma once
#include "chol_loglik.cuh"

namespace bi {

constexpr int PIPE_CONSUMERS = 8;
constexpr int PIPE_THREADS = (PIPE_CONSUMERS + 2) * 32;   // + producer + factor warp
constexpr int PIPE_STAGES = 2;
constexpr int STRIP_ELEMS = TILE_ELEMS / 2;                // 16 x 32 doubles = 4 KB
constexpr int STAGE_ELEMS = TILE_ELEMS + PIPE_CONSUMERS * STRIP_ELEMS;   // B tile + 8 A strips = 40 KB

__host__ __device__ inline size_t pipe_smem_bytes(int Fp, int K0p) {
    size_t b = 0;
    b += sizeof(double) * (size_t)PIPE_STAGES * STAGE_ELEMS;
    b += sizeof(double) * TILE_ELEMS;          // sLinv
    b += sizeof(double) * 32 * 33;             // sD
    b += sizeof(double) * 32;                  // sRinv
    b += sizeof(double) * (size_t)Fp;          // sDy
    b += sizeof(double) * (size_t)K0p * 2;     // sNegV, sM
    b += sizeof(double) * 16;                  // sRed (8) + logdet + spare
    b += sizeof(unsigned long long) * 16;      // mbarriers
    b += sizeof(int) * (size_t)Fp;             // sObs
    b += sizeof(int) * 8;                      // flags
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
__device__ __forceinline__ bool mbar_test(unsigned long long* bar, uint32_t parity) {
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
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
    while (!mbar_test(bar, parity)) {
    }
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

struct PipeParams {
    CholParams c;
    const double* TmP;   // T tiles padded to whole 32-wide K chunks: [nt][nTc][1024] operand-fragment-major
    int nTc;             // number of 32-wide chunks of T (ceil(K0 / 32)); K0p = 32 * nTc
};

// mbarrier slots
enum { BAR_FULL = 0, BAR_EMPTY = PIPE_STAGES, BAR_DIAG = 2 * PIPE_STAGES, BAR_LINV, BAR_PANEL, BAR_COUNT };

template <bool EXPLICIT>
__global__ void __launch_bounds__(PIPE_THREADS, 2) chol_pipe_kernel(const PipeParams PP) {
    const CholParams& P = PP.c;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* sStage = reinterpret_cast<double*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
    double* sLinv = sStage + (size_t)PIPE_STAGES * STAGE_ELEMS;
    double* sD = sLinv + TILE_ELEMS;
    double* sRinv = sD + 32 * 33;
    double* sDy = sRinv + 32;
    const int K0p = PP.nTc * 32;
    double* sNegV = sDy + P.Fp;
    double* sM = sNegV + K0p;
    double* sRed = sM + K0p;
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(sRed + 16);
    int* sObs = reinterpret_cast<int*>(bars + 16);
    int* sFlag = sObs + P.Fp;

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    const int g = lane >> 2;
    const int t = lane & 3;
    const int nt = P.nt;
    const int nTc = EXPLICIT ? 0 : PP.nTc;

    double* wsL = P.ws + (long long)blockIdx.x * P.ws_stride;
    double* wsZ = wsL + (long long)nt * (nt - 1) / 2 * TILE_ELEMS;

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
    if (!EXPLICIT)
        for (int r = tid; r < P.Fp; r += PIPE_THREADS) sObs[r] = P.obs_id[r];
    __syncthreads();

    uint32_t it = 0;       // K-chunks issued / consumed so far (identical sequence in producer and consumers)
    uint32_t gpanel = 0;   // panels processed so far by this CTA (parity of the per-panel barriers)

    for (long long w = blockIdx.x; w < P.W; w += gridDim.x) {
        __syncthreads();   // previous walker fully retired
        if (tid == 0) {
            int inside = 1;
            if (!EXPLICIT && P.check_prior) {
                for (int k = 0; k < P.d; ++k) {
                    const double x = P.theta[w * P.d + k];
                    if (!(x > P.prior_min[k] && x < P.prior_max[k])) inside = 0;
                }
            }
            sFlag[0] = inside;
            sFlag[1] = 0;
        }
        if (!EXPLICIT)
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
        for (int r = tid; r < P.Fp; r += PIPE_THREADS) {
            double s;
            if (EXPLICIT) {
                s = (r < P.F) ? P.dy_x[w * P.F + r] : 0.0;
            } else {
                s = __ldg(P.e + r);
                for (int q = 0; q < P.K0; ++q) s = fma(__ldg(P.Tt + (long long)q * P.Fp + r), sM[q], s);
            }
            sDy[r] = s;
        }
        __syncthreads();

        if (warp == PIPE_CONSUMERS) {
            // =============================== producer ==========================================================
            if (lane == 0) {
                uint32_t gp = gpanel;
                for (int j = 0; j < nt; ++j, ++gp) {
                    const int i_hi = P.imax[j];
                    const int nreg = 2 * (i_hi - j + 1);
                    const int nunits = nreg + 1;
                    bool waited = false;
                    for (int r0 = 0; r0 < nunits; r0 += PIPE_CONSUMERS) {
                        const int nact = min(PIPE_CONSUMERS, nunits - r0);
                        const bool first_z = (r0 == nreg);
                        const int i_first = first_z ? nt : j + (r0 >> 1);
                        // ---- T chunks (theta-dependent part of the covariance)
                        if (!EXPLICIT && !first_z && P.tile_omax[j] >= P.tile_omin[i_first]) {
                            for (int c = 0; c < nTc; ++c, ++it) {
                                const int s = it % PIPE_STAGES;
                                mbar_wait(&bars[BAR_EMPTY + s], ((it / PIPE_STAGES) & 1) ^ 1);
                                double* st = sStage + (size_t)s * STAGE_ELEMS;
                                int nstrip = 0;
                                for (int u = 0; u < nact; ++u) {
                                    const int q = r0 + u;
                                    if (q == nreg) break;
                                    const int i = j + (q >> 1);
                                    if (P.tile_omax[j] >= P.tile_omin[i]) ++nstrip;
                                }
                                mbar_expect_tx(&bars[BAR_FULL + s], (uint32_t)((TILE_ELEMS + nstrip * STRIP_ELEMS) * 8));
                                bulk_g2s(st, PP.TmP + ((size_t)j * nTc + c) * TILE_ELEMS, TILE_ELEMS * 8, &bars[BAR_FULL + s]);
                                for (int u = 0; u < nact; ++u) {
                                    const int q = r0 + u;
                                    if (q == nreg) break;
                                    const int i = j + (q >> 1);
                                    if (P.tile_omax[j] >= P.tile_omin[i])
                                        bulk_g2s(st + TILE_ELEMS + u * STRIP_ELEMS,
                                                 PP.TmP + ((size_t)i * nTc + c) * TILE_ELEMS + (q & 1) * STRIP_ELEMS,
                                                 STRIP_ELEMS * 8, &bars[BAR_FULL + s]);
                                }
                            }
                        }
                        // ---- factor chunks k < j
                        const bool z_in_round = nreg < r0 + PIPE_CONSUMERS;
                        const int klo = P.kmin[z_in_round ? j : i_first];
                        for (int k = klo; k < j; ++k, ++it) {
                            if (k == j - 1 && !waited) {
                                // block column j-1 was written by panel j-1's solves: wait for all of them
                                mbar_wait(&bars[BAR_PANEL], (gp - 1) & 1);
                                waited = true;
                            }
                            const int s = it % PIPE_STAGES;
                            mbar_wait(&bars[BAR_EMPTY + s], ((it / PIPE_STAGES) & 1) ^ 1);
                            double* st = sStage + (size_t)s * STAGE_ELEMS;
                            int nstrip = 0;
                            for (int u = 0; u < nact; ++u) {
                                const int q = r0 + u;
                                const int km = P.kmin[(q == nreg) ? j : j + (q >> 1)];
                                if (k >= km) ++nstrip;
                            }
                            mbar_expect_tx(&bars[BAR_FULL + s], (uint32_t)((TILE_ELEMS + nstrip * STRIP_ELEMS) * 8));
                            bulk_g2s(st, wsL + ((size_t)j * (j - 1) / 2 + k) * TILE_ELEMS, TILE_ELEMS * 8, &bars[BAR_FULL + s]);
                            for (int u = 0; u < nact; ++u) {
                                const int q = r0 + u;
                                const bool uz = (q == nreg);
                                const int i = uz ? nt : j + (q >> 1);
                                if (k < P.kmin[uz ? j : i]) continue;
                                const double* src = uz ? wsZ + (size_t)k * TILE_ELEMS
                                                       : wsL + ((size_t)i * (i - 1) / 2 + k) * TILE_ELEMS + (q & 1) * STRIP_ELEMS;
                                bulk_g2s(st + TILE_ELEMS + u * STRIP_ELEMS, src, STRIP_ELEMS * 8, &bars[BAR_FULL + s]);
                            }
                        }
                    }
                    // every phase of BAR_PANEL is observed exactly once, in order (parity waits must not skip a phase)
                    if (!waited && gp > 0) mbar_wait(&bars[BAR_PANEL], (gp - 1) & 1);
                }
            }
            __syncwarp();
        } else if (warp == PIPE_CONSUMERS + 1) {
            // =============================== factor warp ========================================================
            double logdet = 0.0;
            int first_bad = 0;
            uint32_t gp = gpanel;
            for (int j = 0; j < nt; ++j, ++gp) {
                mbar_wait(&bars[BAR_DIAG], gp & 1);
                const int bad = warp_potrf_inv(sD, sLinv, sRinv, lane, logdet);
                if (bad) {
                    if (first_bad == 0) first_bad = j * 32 + bad;
                    // keep the data flow alive (control flow must not depend on the data): poison the inverse
                    for (int n = 0; n < 32; ++n) sLinv[n * 32 + lane] = __longlong_as_double(0x7ff8000000000000LL);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&bars[BAR_LINV]);
            }
            if (lane == 0) {
                sRed[8] = logdet;
                sFlag[1] = first_bad;
            }
        } else {
            // =============================== consumers ==========================================================
            double zpart = 0.0;
            uint32_t gp = gpanel;
            for (int j = 0; j < nt; ++j, ++gp) {
                const int i_hi = P.imax[j];
                const int nreg = 2 * (i_hi - j + 1);
                const int nunits = nreg + 1;
                const int omin_j = EXPLICIT ? 0 : P.tile_omin[j];
                const int omax_j = EXPLICIT ? 0 : P.tile_omax[j];
                uint32_t deferred = 0;
                int rnd = 0;
                for (int r0 = 0; r0 < nunits; r0 += PIPE_CONSUMERS, ++rnd) {
                    const int q = r0 + warp;
                    const bool active = q < nunits;
                    const bool is_z = (q == nreg);
                    const int i = is_z ? nt : j + (q >> 1);
                    const int h = is_z ? 0 : (q & 1);
                    const bool first_z = (r0 == nreg);
                    const int i_first = first_z ? nt : j + (r0 >> 1);
                    double acc[2][4][2];
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                        for (int nf = 0; nf < 4; ++nf) acc[mi][nf][0] = acc[mi][nf][1] = 0.0;

                    if (!EXPLICIT && !first_z && omax_j >= P.tile_omin[i_first]) {
                        const bool need0 = active && !is_z && omax_j >= P.tile_omin[i];
                        for (int c = 0; c < nTc; ++c, ++it) {
                            const int s = it % PIPE_STAGES;
                            mbar_wait(&bars[BAR_FULL + s], (it / PIPE_STAGES) & 1);
                            if (need0) {
                                const double* Bs = sStage + (size_t)s * STAGE_ELEMS + lane;
                                const double* As = Bs + TILE_ELEMS + warp * STRIP_ELEMS;
#pragma unroll
                                for (int ks = 0; ks < 8; ++ks) {
                                    const double nv = sNegV[c * 32 + 4 * ks + t];
                                    double a[2], b[4];
#pragma unroll
                                    for (int mi = 0; mi < 2; ++mi) a[mi] = As[(mi * 8 + ks) * 32];
#pragma unroll
                                    for (int nf = 0; nf < 4; ++nf) b[nf] = Bs[(nf * 8 + ks) * 32] * nv;
#pragma unroll
                                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                                        for (int nf = 0; nf < 4; ++nf) dmma(acc[mi][nf], a[mi], b[nf]);
                                }
                            }
                            __syncwarp();
                            if (lane == 0) mbar_arrive(&bars[BAR_EMPTY + s]);
                        }
                        if (need0 && omin_j != P.tile_omax[i]) {
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
                        const bool z_in_round = nreg < r0 + PIPE_CONSUMERS;
                        const int klo = P.kmin[z_in_round ? j : i_first];
                        const int my_klo = active ? P.kmin[is_z ? j : i] : 0x7fffffff;
                        for (int k = klo; k < j; ++k, ++it) {
                            const int s = it % PIPE_STAGES;
                            mbar_wait(&bars[BAR_FULL + s], (it / PIPE_STAGES) & 1);
                            if (k >= my_klo) {
                                const double* Bs = sStage + (size_t)s * STAGE_ELEMS + lane;
                                const double* As = Bs + TILE_ELEMS + warp * STRIP_ELEMS;
#pragma unroll
                                for (int ks = 0; ks < 8; ++ks) {
                                    double a[2], b[4];
#pragma unroll
                                    for (int mi = 0; mi < 2; ++mi) a[mi] = As[(mi * 8 + ks) * 32];
#pragma unroll
                                    for (int nf = 0; nf < 4; ++nf) b[nf] = Bs[(nf * 8 + ks) * 32];
#pragma unroll
                                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                                        for (int nf = 0; nf < 4; ++nf) dmma(acc[mi][nf], a[mi], b[nf]);
                                }
                            }
                            __syncwarp();
                            if (lane == 0) mbar_arrive(&bars[BAR_EMPTY + s]);
                        }
                    }
                    if (!active) continue;
                    // ---- C = A_ij - acc
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
                                    const int cc = j * 32 + nf * 8 + 2 * t + e;
                                    const double init = (r < P.F && cc < P.F) ? Cw[(long long)r * P.F + cc] : (r == cc ? 1.0 : 0.0);
                                    acc[mi][nf][e] = init - acc[mi][nf][e];
                                }
                        }
                    } else {
                        const double2* At = reinterpret_cast<const double2*>(P.Atiles + ((long long)i * (i + 1) / 2 + j) * TILE_ELEMS);
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
                                for (int e = 0; e < 2; ++e) sD[(h * 16 + mi * 8 + g) * 33 + nf * 8 + 2 * t + e] = acc[mi][nf][e];
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&bars[BAR_DIAG]);
                        continue;
                    }
                    double* Ct = (is_z ? wsZ + (long long)j * TILE_ELEMS : wsL + ((long long)i * (i - 1) / 2 + j) * TILE_ELEMS) +
                                 h * STRIP_ELEMS + lane;
                    double a[2][8];
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi) cf_to_of(acc[mi], a[mi], lane);
                    if (!mbar_test(&bars[BAR_LINV], gp & 1)) {
                        // inverse not there yet: park the strip, solve at the end of the panel
#pragma unroll
                        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                            for (int ks = 0; ks < 8; ++ks) Ct[(mi * 8 + ks) * 32] = a[mi][ks];
                        deferred |= 1u << rnd;
                        continue;
                    }
                    // ---- triangular solve against the 32 x 32 inverse: X = C L_jj^-T
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                        for (int nf = 0; nf < 4; ++nf) acc[mi][nf][0] = acc[mi][nf][1] = 0.0;
#pragma unroll
                    for (int ks = 0; ks < 8; ++ks)
#pragma unroll
                        for (int nf = 0; nf < 4; ++nf)
                            if (ks <= 2 * nf + 1) {
                                const double b = sLinv[(nf * 8 + ks) * 32 + lane];
#pragma unroll
                                for (int mi = 0; mi < 2; ++mi) dmma(acc[mi][nf], a[mi][ks], b);
                            }
                    if (is_z && g == 0) {
#pragma unroll
                        for (int nf = 0; nf < 4; ++nf) zpart += acc[0][nf][0] * acc[0][nf][0] + acc[0][nf][1] * acc[0][nf][1];
                    }
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi) {
                        double a8[8];
                        cf_to_of(acc[mi], a8, lane);
#pragma unroll
                        for (int ks = 0; ks < 8; ++ks) Ct[(mi * 8 + ks) * 32] = a8[ks];
                    }
                }
                // ---- deferred strips of this panel
                if (deferred) {
                    mbar_wait(&bars[BAR_LINV], gp & 1);
                    for (int rr = 0; deferred >> rr; ++rr) {
                        if (!((deferred >> rr) & 1)) continue;
                        const int q = rr * PIPE_CONSUMERS + warp;
                        const bool is_z = (q == nreg);
                        const int i = is_z ? nt : j + (q >> 1);
                        const int h = is_z ? 0 : (q & 1);
                        double* Ct = (is_z ? wsZ + (long long)j * TILE_ELEMS : wsL + ((long long)i * (i - 1) / 2 + j) * TILE_ELEMS) +
                                     h * STRIP_ELEMS + lane;
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
                        for (int ks = 0; ks < 8; ++ks)
#pragma unroll
                            for (int nf = 0; nf < 4; ++nf)
                                if (ks <= 2 * nf + 1) {
                                    const double b = sLinv[(nf * 8 + ks) * 32 + lane];
#pragma unroll
                                    for (int mi = 0; mi < 2; ++mi) dmma(x[mi][nf], a[mi][ks], b);
                                }
                        if (is_z && g == 0) {
#pragma unroll
                            for (int nf = 0; nf < 4; ++nf) zpart += x[0][nf][0] * x[0][nf][0] + x[0][nf][1] * x[0][nf][1];
                        }
#pragma unroll
                        for (int mi = 0; mi < 2; ++mi) {
                            double a8[8];
                            cf_to_of(x[mi], a8, lane);
#pragma unroll
                            for (int ks = 0; ks < 8; ++ks) Ct[(mi * 8 + ks) * 32] = a8[ks];
                        }
                    }
                } else {
                    // every warp observes the phase so that the parity sequence stays aligned
                    mbar_wait(&bars[BAR_LINV], gp & 1);
                }
                // this panel's tiles -> visible to the async proxy (bulk copies of later panels)
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(&bars[BAR_PANEL]);
            }
            const double zs = warp_sum(zpart);
            if (lane == 0) sRed[warp] = zs;
        }
        gpanel += nt;
        __syncthreads();
        if (tid == 0) {
            const int bad = sFlag[1];
            if (bad) {
                P.logp[w] = __longlong_as_double(0x7ff8000000000000LL);
                if (P.info) P.info[w] = bad > P.F ? P.F : bad;
            } else {
                double tot = 0.0;
                for (int k = 0; k < PIPE_CONSUMERS; ++k) tot += sRed[k];
                P.logp[w] = -0.5 * tot - sRed[8];
                if (P.info) P.info[w] = 0;
            }
        }
    }
}

}  // namespace bi
