// bi_api.cu -- the C ABI of include/bi_b200.h: device image construction and kernel launches.  No torch, no CPU path.
#include <cuda_runtime.h>

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

#include "../../include/bi_b200.h"
#include "block_direct.cuh"
#include "chol_loglik.cuh"
#include "chol_pipe.cuh"
#include "common.cuh"
#include "cov_assemble.cuh"
#include "ensemble.cuh"
#include "gp_predict.cuh"
#include "gp_train.cuh"
#include "setup_kernels.cuh"
#include "woodbury.cuh"

using namespace bi;

namespace {

// The dynamic shared-memory limit of a kernel is an attribute of the FUNCTION on the current device, shared by every handle
// of the process.  Setting it to what the handle being created needs would lower it under an older handle of a larger
// problem (its next launch then fails with "invalid argument"), so it is always raised to the device's opt-in maximum;
// what a launch actually asks for is checked against that maximum where the handle is built.
template <class Kernel>
cudaError_t allow_max_smem(Kernel* kernel) {
    int dev = 0, optin = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, optin);
    return e;
}

// Launch with (pdl) or without the programmatic-serialisation attribute (common.cuh pdl_wait / pdl_trigger): the kernels of a
// sampler half-step -- propose, K1+K2, K5 direct, combine, accept -- start while their predecessor drains and run their
// theta-independent prologue (staging of X / ell, of T_b and A_b) ahead of it.  Captured as a programmatic edge in a CUDA graph.
template <typename... KArgs, typename... Args>
cudaError_t launch_k(bool pdl, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

struct GroupImg {
    int N = 0, P = 0, Fg = 0, R = 0, nks = 0, ntn = 0, K0g = 0, pc_off = 0;
    double* dXsT = nullptr;    // (P, d, N)
    double* dEll = nullptr;    // (P, d)
    double* dLinvF = nullptr;  // (P, ntn, 4, nks, 32)
    double* dTg = nullptr;     // (Fg, K0g) row-major
    double* dUs = nullptr;     // (Fg, Fg)
    double* dMean = nullptr;   // (Fg)
};

}  // namespace

struct bi_handle_s {
    int device = 0, d = 0, G = 0, B = 0, F = 0, Fp = 0, nt = 0, Ptot = 0, K0 = 0, nks0 = 0, Nmax = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    std::vector<GroupImg> groups;
    std::vector<void*> allocs;
    bool guard = false;                                        // BI_DEBUG_GUARD=1: canary bands + poisoned payloads
    std::vector<std::pair<unsigned char*, size_t>> guards;     // (base of the allocation, payload bytes)
    PcDesc* dPcs = nullptr;
    double *dMin = nullptr, *dMax = nullptr;
    std::vector<double> hMin, hMax;
    bool has_data = false, sigma_blockdiag = true;
    double *dTrow = nullptr, *dTt = nullptr, *dTm = nullptr, *dE = nullptr, *dMeanMerged = nullptr, *dUblk = nullptr;
    double *dSig = nullptr, *dYerr = nullptr, *dATiles = nullptr, *dTmP = nullptr;
    int nTc = 0, chol_impl = 2;
    size_t pipe_smem = 0, pipe_smem_wide = 0;
    // 4-consumer instantiation of K4 (three CTAs per SM) for few-tile problems and launches of more than two walkers per SM:
    // -1 = choose per launch, 0 = never, 1 = always (developer knob BI_PIPE_SMALL)
    int pipe_small = -1;
    size_t pipe_smem_small = 0;
    int chol_ctas_small = 0;
    int pipe_wide = 0;        // 16-consumer instantiation of K4 available (tail of a launch, launches of <= one walker per SM)
    // A tiles (Sigma_exp + blockdiag(U s s^T) / n_div) per divisor: slot 0 is n_div = 1 (what an MCMC evaluates; built
    // on its first use and never rebuilt, so a captured CUDA graph that reads it stays valid whatever else the handle is asked
    // afterwards), slots 1.. rotate among the other divisors (log_posterior(X) with the batch-count divisor, predict-style calls).
    static constexpr int A_SLOTS = 4;
    double* dASlot[A_SLOTS] = {nullptr, nullptr, nullptr, nullptr};
    double a_slot_ndiv[A_SLOTS] = {-1.0, -1.0, -1.0, -1.0};
    unsigned long long a_slot_stamp[A_SLOTS] = {0, 0, 0, 0}, a_clock = 0;
    // experimental data sets for batched closure tests (bi_set_data_sets)
    int n_sets = 0;
    double *dEsets = nullptr, *dYerrSets = nullptr, *dATilesSets = nullptr;
    double sets_ndiv = -1.0;
    std::vector<double> hMeanMerged;
    // Woodbury mode (allocated on first use)
    std::vector<std::array<long long, 4>> blk;   // observable blocks sorted by out_start: [group, out, in, len]
    std::vector<WbBlock> wbBlocks;
    WbBlock* dWbBlocks = nullptr;
    int wbPmax = 0, wbNt = 0;
    long long wbChunk = 0, wbStride = 0;
    int wbCtas = 0;
    size_t wbSmem = 0;
    double wb_ndiv = -1.0;
    double *dG = nullptr, *dH = nullptr, *dQ0 = nullptr, *dLdA = nullptr;
    double *dMw = nullptr, *dUw = nullptr, *dQuadA = nullptr, *dQuadM = nullptr, *dLdM = nullptr, *dLpW = nullptr, *dWbWs = nullptr;
    int *dInfoW = nullptr, *dWbK = nullptr;
    // block-direct evaluation of BI_MODE_STRUCTURED (every observable block fits one warp: f_b <= 31)
    int bd_state = 0;             // 0 not looked at yet, 1 usable, -1 not applicable (falls back to the tile-skipping kernel)
    int bdFB = 0, bdNks = 0;      // row length padded to a multiple of 8; k-steps of 4 principal components (max over blocks)
    bool bdAugIn = false;         // every block leaves row FB - 1 free: the augmented row rides inside the last fragment
    int* dBdCount = nullptr;      // (max_batch) per-walker count of delivered block terms: the combine step fused into the kernel
    bool bdFuse = true;           // developer knob BI_BD_FUSE=0: separate block_combine_kernel launch
    std::vector<BdBlock> bdBlocks;
    BdBlock* dBdBlocks = nullptr;
    double* dBdA[2] = {nullptr, nullptr};      // [0] n_div = 1 (never rewritten), [1] the last other divisor
    double bd_ndiv[2] = {-1.0, -1.0};
    double* dBdPart = nullptr;
    int* dBdBad = nullptr;
    double* dBdASets = nullptr;   // (n_sets, sum f_b^2): A_b per experimental data set (bi_set_data_sets)
    double bd_sets_ndiv = -1.0;
    long long bd_a_tot = 0;
    int *dObs = nullptr, *dTileOmin = nullptr, *dTileOmax = nullptr;
    int *dKminDense = nullptr, *dImaxDense = nullptr, *dKminStruct = nullptr, *dImaxStruct = nullptr;
    std::vector<int> hOmin, hOmax, hKmin[2], hImax[2];   // host copies of the tile tables ([0] dense, [1] structured)
    long long max_batch = 4096;
    double *dTheta = nullptr, *dM = nullptr, *dV = nullptr, *dLogp = nullptr;
    int* dInfo = nullptr;
    unsigned long long* dCount = nullptr;
    double* dWs = nullptr;
    int chol_ctas = 0;
    long long ws_stride = 0;
    size_t chol_smem = 0, gp_smem = 0;
    int gp_wt = 32;
    int gp_small = 1;         // 8-walker tiles for launches of less than one wave (developer knob BI_GP_SMALL=0 switches it off)
    bool pdl = true;          // programmatic dependent launch for the half-step kernels (developer knob BI_PDL=0 switches it off)
    int gp_mixed = 1;         // 32 + 16-walker tiles in one launch when the last wave is partly empty (BI_GP_MIXED: 0 off,
                              // 1 auto, -n forces n 32-walker tiles)
    int k3_persist = -1;      // covariance assembly: -1 = choose per layout, 0 / 1 = developer override (BI_K3_PERSIST)
    double k3_block_share = 1.0;
    double* dOutScratch = nullptr;   // staging for *_host outputs (mean / cov / k)
    size_t out_scratch_bytes = 0;
    long long launches = 0;
    size_t scratch_bytes = 0;
    bool profiling = false;
    std::vector<std::array<cudaEvent_t, 2>> prof_events[2];   // [0] gp_predict, [1] chol_loglik
    double prof_ms[2] = {0.0, 0.0};
    long long prof_n[2] = {0, 0};
    std::string err;
};

namespace {

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            if (h) h->err = std::string(#call) + ": " + cudaGetErrorString(e_);                    \
            return (int)e_;                                                                        \
        }                                                                                          \
    } while (0)

int fail(bi_handle h, int code, const char* msg) {
    if (h) h->err = msg;
    return code;
}

// Guard mode (BI_DEBUG_GUARD=1, developer / test knob; compute-sanitizer is closed on this pool): every buffer of the
// handle is allocated between two 4 KB canary bands and its payload is poisoned with 0xFF bytes (a NaN for doubles), so
// that a store outside a buffer shows up in bi_query(h, 32) and a read of never-written scratch shows up as NaN in the
// results.  No effect on the kernels themselves.
constexpr size_t GUARD_BYTES = 4096;
constexpr unsigned char GUARD_PATTERN = 0xA5;

template <typename T>
int dalloc(bi_handle h, T** p, size_t n) {
    void* q = nullptr;
    const size_t bytes = std::max<size_t>(n, 1) * sizeof(T);
    const size_t pad = h->guard ? GUARD_BYTES : 0;
    cudaError_t e = cudaMalloc(&q, bytes + 2 * pad);
    if (e == cudaSuccess && h->guard) {
        // on the handle's own stream: the uploads that follow (cudaMemcpyAsync on h->stream, a non-blocking stream) must
        // not be overtaken by the poison
        e = cudaMemsetAsync(q, GUARD_PATTERN, bytes + 2 * pad, h->stream);
        if (e == cudaSuccess) e = cudaMemsetAsync(static_cast<unsigned char*>(q) + pad, 0xFF, bytes, h->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
        h->guards.push_back({static_cast<unsigned char*>(q), bytes});
    }
    if (e != cudaSuccess) {
        h->err = std::string("cudaMalloc: ") + cudaGetErrorString(e);
        return (int)e;
    }
    h->allocs.push_back(q);
    h->scratch_bytes += n * sizeof(T);
    *p = reinterpret_cast<T*>(static_cast<unsigned char*>(q) + pad);
    return 0;
}

// number of canary bands that no longer hold their pattern (guard mode), -1 when guard mode is off
int64_t check_guards(bi_handle h) {
    if (!h->guard) return -1;
    cudaSetDevice(h->device);
    cudaDeviceSynchronize();
    std::vector<unsigned char> buf(GUARD_BYTES);
    int64_t broken = 0;
    for (auto& g : h->guards) {
        for (int side = 0; side < 2; ++side) {
            const unsigned char* src = side == 0 ? g.first : g.first + GUARD_BYTES + g.second;
            if (cudaMemcpy(buf.data(), src, GUARD_BYTES, cudaMemcpyDeviceToHost) != cudaSuccess) return -2;
            bool ok = true;
            for (unsigned char b : buf) ok = ok && b == GUARD_PATTERN;
            broken += !ok;
        }
    }
    return broken;
}

template <typename T>
int upload(bi_handle h, T** p, const T* src, size_t n) {
    int rc = dalloc(h, p, n);
    if (rc) return rc;
    CK(cudaMemcpyAsync(*p, src, n * sizeof(T), cudaMemcpyHostToDevice, h->stream));
    return 0;
}

int ensure_out_scratch(bi_handle h, size_t bytes) {
    if (bytes <= h->out_scratch_bytes) return 0;
    if (h->dOutScratch) {
        CK(cudaStreamSynchronize(h->stream));
        CK(cudaFree(h->dOutScratch));
        h->dOutScratch = nullptr;
        h->out_scratch_bytes = 0;
    }
    CK(cudaMalloc(&h->dOutScratch, bytes));
    h->out_scratch_bytes = bytes;
    return 0;
}

// Selects (building it on stream s when needed) the A-tile buffer of this divisor and makes it h->dATiles.  A handle is
// single-stream: calls that overlap on different streams are not supported (include/bi_b200.h).
int build_a_tiles(bi_handle h, double n_div, cudaStream_t s) {
    int slot = -1;
    if (n_div == 1.0) slot = 0;
    else
        for (int i = 1; i < bi_handle_s::A_SLOTS; ++i)
            if (h->a_slot_ndiv[i] == n_div) slot = i;
    if (slot < 0) {                                   // least recently used of the rotating slots
        slot = 1;
        for (int i = 2; i < bi_handle_s::A_SLOTS; ++i)
            if (h->a_slot_stamp[i] < h->a_slot_stamp[slot]) slot = i;
        h->a_slot_ndiv[slot] = -1.0;
    }
    h->a_slot_stamp[slot] = ++h->a_clock;
    h->dATiles = h->dASlot[slot];
    if (h->a_slot_ndiv[slot] == n_div) return 0;
    const int ntl = h->nt * (h->nt + 1) / 2;
    build_a_tiles_kernel<<<ntl, 256, 0, s>>>(h->dSig, h->dYerr, h->dUblk, n_div, h->F, h->chol_impl == 2 ? 1 : 0, h->dATiles);
    h->launches++;
    CK(cudaGetLastError());
    h->a_slot_ndiv[slot] = n_div;
    return 0;
}

struct ProfScope {
    bi_handle h; int k; cudaStream_t s; cudaEvent_t e1 = nullptr;
    ProfScope(bi_handle h_, int k_, cudaStream_t s_) : h(h_), k(k_), s(s_) {
        if (!h->profiling) return;
        cudaEvent_t e0;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        cudaEventRecord(e0, s);
        h->prof_events[k].push_back({e0, e1});
    }
    ~ProfScope() { if (e1) cudaEventRecord(e1, s); }
};

int launch_gp(bi_handle h, const double* dTheta, long long W, double* dM, double* dV, cudaStream_t s) {
    int wt = h->gp_wt;
    // small batches (a sampler half-step of a 100-walker ensemble): fewer than one wave of 32-walker CTAs -> 8-walker tiles
    const bool small_batch = wt == 32 && h->gp_small && ((W + 31) / 32) * h->Ptot < 2ll * h->sm_count;
    if (small_batch && h->gp_small != 16) wt = 8;
    dim3 grid((unsigned)((W + wt - 1) / wt), (unsigned)h->Ptot);
    // Wave tail: more than one wave of 32-walker CTAs whose last wave is partly empty -> the first n_big tiles stay 32 wide
    // (as many as fill whole waves), the remaining walkers go in 16-walker tiles in the same launch (gp_predict_mixed_kernel)
    long long n_big = -1;
    if (wt == 32 && h->gp_mixed != 0) {
        const long long S = 2ll * h->sm_count, P = h->Ptot, tiles = (W + 31) / 32, units = tiles * P;
        if (small_batch) n_big = 0;                                                        // BI_GP_SMALL=16: 16-walker tiles only
        else if (h->gp_mixed < 0) n_big = std::min<long long>(-h->gp_mixed, (W - 1) / 32);  // developer knob: forced
        else if (units > S && units % S != 0 && (units % S) * 5 <= S * 4) {
            // as many 32-walker tiles as fill whole waves.  Measured (C3a, L2 flushed): 1024 walkers 113.6 -> 105.5 us,
            // 700 walkers 86.0 -> 76.8, 512 walkers 69.7 -> 64.5; insensitive to +-3 tiles; a last wave that is more than
            // 80 % full is left alone
            const long long cand = std::min<long long>((units / S) * S / P, (W - 1) / 32);
            if (cand > 0) n_big = cand;
        }
    }
    {
        ProfScope ps(h, 0, s);
#ifdef GP_TIMING
        { unsigned long long z[8] = {0}; cudaMemcpyToSymbolAsync(gp_dbg, z, sizeof(z), 0, cudaMemcpyHostToDevice, s); }
#endif
        if (n_big >= 0) {
            const long long n_small = (W - 32 * n_big + GP_WT_TAIL - 1) / GP_WT_TAIL;
            grid = dim3((unsigned)((n_big + n_small) * h->Ptot), 1);
            CK(launch_k(h->pdl, gp_predict_mixed_kernel, grid, dim3(256), h->gp_smem, s, h->dPcs, h->Ptot, h->d, dTheta, W, (int)n_big,
                        (int)n_small, dM, dV, h->Ptot));
        } else if (wt == 64)
            gp_predict_kernel<true, 64><<<grid, 512, h->gp_smem, s>>>(h->dPcs, 0, h->d, dTheta, W, nullptr, 0, dM, dV, h->Ptot);
        else if (wt == 8)
            CK(launch_k(h->pdl, gp_predict_kernel<true, 8>, grid, dim3(gp_threads(8)), gp_smem_bytes(h->d, h->Nmax, 8), s, h->dPcs, 0, h->d,
                        dTheta, W, (double*)nullptr, 0, dM, dV, h->Ptot));
        else
            CK(launch_k(h->pdl, gp_predict_kernel<true, 32>, grid, dim3(256), h->gp_smem, s, h->dPcs, 0, h->d, dTheta, W, (double*)nullptr, 0,
                        dM, dV, h->Ptot));
#ifdef GP_TIMING
        {
            unsigned long long z[8];
            cudaMemcpyFromSymbolAsync(z, gp_dbg, sizeof(z), 0, cudaMemcpyDeviceToHost, s);
            cudaStreamSynchronize(s);
            const double nc = (double)grid.x * grid.y;
            std::fprintf(stderr, "[gp timing] ctas %.0f: load %.0f  k1 %.0f  sync %.0f  k2 %.0f  reduce+sync %.0f cycles per CTA (thread 0)\n", nc,
                         z[0] / nc, z[1] / nc, z[2] / nc, z[3] / nc, z[4] / nc);
        }
#endif
    }
    h->launches++;
    CK(cudaGetLastError());
    return 0;
}

int launch_chol(bi_handle h, const double* dTheta, const double* dM, const double* dV, long long W, int mode,
                double* dLogp, int* dInfo, cudaStream_t s, const int* dSetId = nullptr) {
    CholParams P;
    P.F = h->F; P.Fp = h->Fp; P.nt = h->nt; P.K0 = h->K0; P.nks0 = h->nks0; P.Ptot = h->Ptot; P.d = h->d;
    P.W = W; P.theta = dTheta; P.m = dM; P.v = dV; P.prior_min = h->dMin; P.prior_max = h->dMax;
    P.Tt = h->dTt; P.Tm = h->dTm; P.e = h->dE; P.Atiles = h->dATiles; P.obs_id = h->dObs;
    P.tile_omin = h->dTileOmin; P.tile_omax = h->dTileOmax;
    P.kmin = (mode == BI_MODE_DENSE) ? h->dKminDense : h->dKminStruct;
    P.imax = (mode == BI_MODE_DENSE) ? h->dImaxDense : h->dImaxStruct;
    P.ws = h->dWs; P.ws_stride = h->ws_stride; P.logp = dLogp; P.info = dInfo; P.check_prior = 1;
    P.cov_x = nullptr; P.dy_x = nullptr; P.quad_out = nullptr; P.logdet_out = nullptr;
    P.set_id = dSetId; P.e_stride = 0; P.a_stride = 0;
    if (dSetId) {
        P.e = h->dEsets; P.Atiles = h->dATilesSets;
        P.e_stride = h->Fp; P.a_stride = (long long)h->nt * (h->nt + 1) / 2 * TILE_ELEMS;
    }
    const int grid = (int)std::min<long long>(W, h->chol_ctas);
    {
        ProfScope ps(h, 1, s);
        if (h->chol_impl == 2) {
            PipeParams PP;
            PP.c = P; PP.TmP = h->dTmP; PP.nTc = h->nTc; PP.dbg = nullptr;
#ifdef PIPE_TIMING
            static long long* dbg = nullptr;
            if (!dbg) cudaMalloc(&dbg, 64 * 8);
            cudaMemsetAsync(dbg, 0, 64 * 8, s);
            PP.dbg = dbg;
#endif
            // Walkers of a last wave that fills at most half of the resident CTA slots (one walker per SM) go to the
            // 16-consumer instantiation: alone on its SM a CTA of 8 consumers is no faster than next to a second one.
            // Measured (C3a, 1024 walkers = 3 full waves + 136): 3.62 -> 3.30 ms.  Only where a panel has more strips than
            // 8 warps take in a round: with tile skipping (structured mode) or F <= 288 it changes nothing or loses 1-2 %.
            long long w_tail = 0;
#ifndef PIPE_TIMING
            if (h->pipe_wide && mode == BI_MODE_DENSE && h->nt >= 10) {
                const long long rem = W % h->chol_ctas;
                if (rem > 0 && rem <= h->sm_count) w_tail = rem;
            }
#endif
            const long long w_main = W - w_tail;
            if (w_main > 0) {
                PP.c.W = w_main;
                // few tiles (F <= 287): a walker's time is the chain potrf -> solve -> next diagonal tile, so with more than
                // two walkers per SM on offer a third CTA per SM (4 consumers each) beats wider rounds: C1 shapes, 1200
                // walkers 0.489 -> 0.408 ms; 100 walkers 0.093 -> 0.100 ms (profiles/r02_notes.md).  Same arithmetic per strip.
                const bool small = h->chol_ctas_small > 0 &&
                                   (h->pipe_small == 1 || (h->pipe_small < 0 && h->nt < 10 && w_main > 2ll * h->sm_count));
                if (small)
                    chol_pipe_kernel<PIPE_CONSUMERS_SMALL><<<(int)std::min<long long>(w_main, h->chol_ctas_small),
                                                             pipe_threads(PIPE_CONSUMERS_SMALL), h->pipe_smem_small, s>>>(PP);
                else
                    chol_pipe_kernel<PIPE_CONSUMERS_MAIN><<<(int)std::min<long long>(w_main, h->chol_ctas),
                                                            pipe_threads(PIPE_CONSUMERS_MAIN), h->pipe_smem, s>>>(PP);
            }
            if (w_tail > 0) {
                PipeParams PT = PP;
                PT.c.W = w_tail;
                PT.c.theta += w_main * h->d; PT.c.m += w_main * h->Ptot; PT.c.v += w_main * h->Ptot; PT.c.logp += w_main;
                if (PT.c.info) PT.c.info += w_main;
                if (PT.c.set_id) PT.c.set_id += w_main;
                chol_pipe_kernel<PIPE_CONSUMERS_WIDE><<<(int)w_tail, pipe_threads(PIPE_CONSUMERS_WIDE), h->pipe_smem_wide, s>>>(PT);
                if (w_main > 0) h->launches++;
            }
#ifdef PIPE_TIMING
            {
                long long hb[64];
                cudaMemcpyAsync(hb, dbg, sizeof(hb), cudaMemcpyDeviceToHost, s);
                cudaStreamSynchronize(s);
                const char* nm[6] = {"wait_full", "sweep_mma", "epi_park/diag", "epi_fast_trsm", "wait_linv", "deferred+fence"};
                double tot = 0;
                for (int c = 0; c < 6; ++c) for (int wq = 0; wq < 8; ++wq) tot += (double)hb[wq * 8 + c];
                std::fprintf(stderr, "[pipe timing] W=%lld grid=%d: ", (long long)W, grid);
                for (int c = 0; c < 6; ++c) {
                    double v = 0;
                    for (int wq = 0; wq < 8; ++wq) v += (double)hb[wq * 8 + c];
                    std::fprintf(stderr, "%s %.1f%%  ", nm[c], 100.0 * v / tot);
                }
                {
                    unsigned long long pz[8];
                    cudaMemcpyFromSymbol(pz, potrf_dbg, sizeof(pz));
                    const double np_ = (double)W * h->nt;
                    std::fprintf(stderr, "| potrf phases (cycles/panel): load %.0f chain %.0f store+trailing %.0f inverse %.0f Ljj %.0f ",
                                 pz[0] / np_, pz[1] / np_, pz[2] / np_, pz[3] / np_, pz[4] / np_);
                    unsigned long long zz8[8] = {0};
                    cudaMemcpyToSymbol(potrf_dbg, zz8, sizeof(zz8));
                }
                std::fprintf(stderr, "| wait_full by warp:");
                for (int wq = 0; wq < 8; ++wq) {
                    double tw = 0;
                    for (int c = 0; c < 6; ++c) tw += (double)hb[wq * 8 + c];
                    std::fprintf(stderr, " %.0f", 100.0 * hb[wq * 8 + 0] / tw);
                }
                std::fprintf(stderr, " | cycles/walker/warp %.0f | factor warp: potrf %.0f cyc/panel, idle %.0f cyc/panel\n", tot / 8.0 / (double)W,
                             (double)hb[6] / (double)W / h->nt, (double)hb[7] / (double)W / h->nt);
            }
#endif
        } else {
            chol_loglik_kernel<false><<<grid, CHOL_THREADS, h->chol_smem, s>>>(P);
        }
    }
    h->launches++;
    CK(cudaGetLastError());
    return 0;
}

}  // namespace

extern "C" {

int bi_abi_version(void) { return BI_ABI_VERSION; }

const char* bi_last_error(bi_handle h) { return h ? h->err.c_str() : "null handle"; }

// DMMA (mma.sync m8n8k4 f64 = 512 flop) instructions chol_pipe_kernel issues for ONE walker inside the prior box, from the
// same tile tables the kernel walks (kmin / imax / tile_omin / tile_omax): what bench.py reports as `flops_executed`.
static int64_t k4_dmma_per_walker(bi_handle h, int mode) {
    if (!h->has_data || h->chol_impl != 2 || mode < 0 || mode > 1 || (int)h->hOmin.size() != h->nt) return -1;
    const int nt = h->nt, nTc = h->nTc;
    const std::vector<int>&kmin = h->hKmin[mode], &imax = h->hImax[mode], &omin = h->hOmin, &omax = h->hOmax;
    const int ra = h->F - (nt - 1) * 32, hA = ra >> 4;
    int64_t n = 0;
    for (int j = 0; j < nt; ++j) {
        const int i_hi = imax[j];
        const int nreg = 2 * (i_hi - j + 1);
        const int nunits = nreg + (i_hi < nt - 1 ? 1 : 0);
        const int qA = (i_hi < nt - 1) ? nreg : 2 * (nt - 1 - j) + hA;
        const int kA = (j == nt - 1) ? 0 : kmin[j];
        for (int q = 0; q < nunits; ++q) {
            const bool aug = q >= nreg;
            const int i = aug ? nt - 1 : j + (q >> 1);
            if (!aug && omax[j] >= omin[i]) n += 2ll * nTc * 32;           // theta-dependent T D T^T chunks of this strip
            const int klo = (q == qA) ? kA : kmin[i];
            if (j > klo) n += (int64_t)(j - klo) * 64;                       // L_ik L_jk^T chunks
            if (i != j) n += 40;                                             // blocked triangular solve of the strip
        }
        n += 20;                                                             // trailing updates inside the diagonal tile's potrf
    }
    return n;
}

static int ensure_block_direct(bi_handle h);

int64_t bi_query(bi_handle h, int what) {
    if (!h) return -1;
    if (what == 16 || what == 17) return k4_dmma_per_walker(h, what - 16);
    if (what == 32) return check_guards(h);
    if (what == 18) {
        if (!h->has_data || !h->sigma_blockdiag) return 0;
        cudaSetDevice(h->device);
        return (ensure_block_direct(h) == 0 && h->bd_state == 1) ? 1 : 0;
    }
    switch (what) {
        case 0: return h->F;
        case 1: return h->Ptot;
        case 2: return h->d;
        case 3: return h->G;
        case 4: return h->sm_count;
        case 5: return h->chol_ctas;
        case 6: return (int64_t)h->scratch_bytes;
        case 7: return h->launches;
        default: return -1;
    }
}

int bi_profile(bi_handle h, int enable) {
    if (!h) return BI_E_ARG;
    h->profiling = enable != 0;
    return 0;
}

int bi_profile_read(bi_handle h, int kernel, double* total_ms, int64_t* n_launches) {
    if (!h || kernel < 0 || kernel > 1 || !total_ms || !n_launches) return BI_E_ARG;
    CK(cudaSetDevice(h->device));
    CK(cudaDeviceSynchronize());
    for (auto& ev : h->prof_events[kernel]) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, ev[0], ev[1]) == cudaSuccess) { h->prof_ms[kernel] += ms; h->prof_n[kernel]++; }
        cudaEventDestroy(ev[0]); cudaEventDestroy(ev[1]);
    }
    h->prof_events[kernel].clear();
    *total_ms = h->prof_ms[kernel];
    *n_launches = h->prof_n[kernel];
    h->prof_ms[kernel] = 0.0; h->prof_n[kernel] = 0;
    return 0;
}

int bi_sync(bi_handle h) {
    if (!h) return BI_E_ARG;
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    return 0;
}

void bi_destroy(bi_handle h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    for (void* p : h->allocs) cudaFree(p);
    if (h->dOutScratch) cudaFree(h->dOutScratch);
    if (h->dEsets) cudaFree(h->dEsets);
    if (h->dYerrSets) cudaFree(h->dYerrSets);
    if (h->dATilesSets) cudaFree(h->dATilesSets);
    if (h->dBdASets) cudaFree(h->dBdASets);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

int bi_create(const bi_desc* D, bi_handle* out) {
    if (!D || !out) return BI_E_ARG;
    *out = nullptr;
    if (D->abi_version != BI_ABI_VERSION) return BI_E_ARG;
    if (D->d <= 0 || D->n_groups <= 0 || D->n_blocks <= 0 || D->n_features <= 0 || !D->groups || !D->blocks ||
        !D->prior_min || !D->prior_max)
        return BI_E_ARG;
    bi_handle h = new bi_handle_s();
    if (const char* e = std::getenv("BI_DEBUG_GUARD")) h->guard = std::atoi(e) != 0;
    auto bail = [&](int rc) {
        std::fprintf(stderr, "bi_create failed (%d): %s\n", rc, h->err.c_str());
        bi_destroy(h);
        return rc;
    };
#define TRY(x)                      \
    do {                            \
        int rc_ = (x);              \
        if (rc_) return bail(rc_);  \
    } while (0)
#define CKB(call)                                                              \
    do {                                                                       \
        cudaError_t e_ = (call);                                               \
        if (e_ != cudaSuccess) {                                               \
            h->err = std::string(#call) + ": " + cudaGetErrorString(e_);       \
            return bail((int)e_);                                              \
        }                                                                      \
    } while (0)

    h->device = D->device;
    CKB(cudaSetDevice(h->device));
    cudaDeviceProp prop;
    CKB(cudaGetDeviceProperties(&prop, h->device));
    if (prop.major < 10) {
        h->err = "this library is built for sm_100a (B200) only";
        return bail(BI_E_UNSUPPORTED);
    }
    h->sm_count = prop.multiProcessorCount;
    CKB(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    h->d = D->d; h->G = D->n_groups; h->B = D->n_blocks; h->F = D->n_features;
    // one more row than F: row F of the factorised matrix is the augmented dY row of the pipelined kernel (chol_pipe.cuh);
    // the v1 kernel sees it as one more identity padding row
    h->nt = (h->F + 1 + 31) / 32; h->Fp = h->nt * 32;
    if (D->max_batch > 0) h->max_batch = D->max_batch;
    const int d = h->d, F = h->F, Fp = h->Fp, nt = h->nt;
    h->hMin.assign(D->prior_min, D->prior_min + d);
    h->hMax.assign(D->prior_max, D->prior_max + d);
    TRY(upload(h, &h->dMin, D->prior_min, (size_t)d));
    TRY(upload(h, &h->dMax, D->prior_max, (size_t)d));

    // ---- per-group images -------------------------------------------------------------------------------------
    h->groups.resize(h->G);
    int pc_off = 0;
    for (int gi = 0; gi < h->G; ++gi) {
        const bi_group_desc& g = D->groups[gi];
        GroupImg& I = h->groups[gi];
        if (g.n_design <= 0 || g.n_pc <= 0 || g.n_features <= 0 || g.n_rest < 0 || !g.X || !g.kind || !g.ell ||
            !g.constant || !g.noise || !g.alpha || !g.L || !g.S || !g.scale || !g.mean ||
            (g.n_rest > 0 && (!g.S_rest || !g.ev_rest))) {
            h->err = "group descriptor has a null or non-positive field";
            return bail(BI_E_ARG);
        }
        I.N = g.n_design; I.P = g.n_pc; I.Fg = g.n_features; I.R = g.n_rest;
        I.nks = ((I.N + 15) / 16) * 4; I.ntn = (I.N + 1 + 31) / 32; I.K0g = ((I.P + 3) / 4) * 4; I.pc_off = pc_off;
        pc_off += I.P;
        h->Nmax = std::max(h->Nmax, I.N);
        const int N = I.N, Pg = I.P, Fg = I.Fg;
        for (int p = 0; p < Pg; ++p)
            if (g.kind[p] < 0 || g.kind[p] > 3) {
                h->err = "unknown kernel kind";
                return bail(BI_E_ARG);
            }
        // X / ell, transposed (sklearn divides both operands by length_scale before cdist, kernels.py:1688)
        std::vector<double> xs((size_t)Pg * d * N);
        for (int p = 0; p < Pg; ++p)
            for (int k = 0; k < d; ++k)
                for (int j = 0; j < N; ++j) xs[((size_t)p * d + k) * N + j] = g.X[(size_t)j * d + k] / g.ell[(size_t)p * d + k];
        TRY(upload(h, &I.dXsT, xs.data(), xs.size()));
        TRY(upload(h, &I.dEll, g.ell, (size_t)Pg * d));
        // L^-1 (+ alpha row), operand-fragment-major
        double *dL = nullptr, *dX = nullptr, *dAlpha = nullptr;
        CKB(cudaMalloc(&dL, (size_t)Pg * N * N * sizeof(double)));
        CKB(cudaMalloc(&dX, (size_t)Pg * N * N * sizeof(double)));
        CKB(cudaMalloc(&dAlpha, (size_t)Pg * N * sizeof(double)));
        CKB(cudaMemcpyAsync(dL, g.L, (size_t)Pg * N * N * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        CKB(cudaMemcpyAsync(dAlpha, g.alpha, (size_t)Pg * N * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        CKB(cudaMemsetAsync(dX, 0, (size_t)Pg * N * N * sizeof(double), h->stream));
        const size_t inv_smem = (size_t)N * 32 * sizeof(double);
        if (inv_smem > 227 * 1024) {
            h->err = "n_design too large for the triangular-inverse setup kernel (max 908)";
            return bail(BI_E_UNSUPPORTED);
        }
        CKB(allow_max_smem(tri_invert_kernel));
        tri_invert_kernel<<<dim3((N + 31) / 32, Pg), 32, inv_smem, h->stream>>>(dL, dX, N);
        CKB(cudaGetLastError());
        const size_t per_pc = (size_t)I.ntn * 4 * I.nks * 32;
        TRY(dalloc(h, &I.dLinvF, per_pc * Pg));
        pack_linv_kernel<<<dim3((unsigned)((per_pc + 255) / 256), Pg), 256, 0, h->stream>>>(dX, dAlpha, I.dLinvF, N, I.nks, I.ntn);
        CKB(cudaGetLastError());
        // T_g = scale * S^T (Fg, K0g), unexplained covariance, scaler mean
        std::vector<double> tg((size_t)Fg * I.K0g, 0.0);
        for (int r = 0; r < Fg; ++r)
            for (int q = 0; q < Pg; ++q) tg[(size_t)r * I.K0g + q] = g.scale[r] * g.S[(size_t)q * Fg + r];
        TRY(upload(h, &I.dTg, tg.data(), tg.size()));
        TRY(upload(h, &I.dMean, g.mean, (size_t)Fg));
        TRY(dalloc(h, &I.dUs, (size_t)Fg * Fg));
        if (I.R > 0) {
            double *dSr = nullptr, *dEv = nullptr, *dSc = nullptr;
            CKB(cudaMalloc(&dSr, (size_t)I.R * Fg * sizeof(double)));
            CKB(cudaMalloc(&dEv, (size_t)I.R * sizeof(double)));
            CKB(cudaMalloc(&dSc, (size_t)Fg * sizeof(double)));
            CKB(cudaMemcpyAsync(dSr, g.S_rest, (size_t)I.R * Fg * sizeof(double), cudaMemcpyHostToDevice, h->stream));
            CKB(cudaMemcpyAsync(dEv, g.ev_rest, (size_t)I.R * sizeof(double), cudaMemcpyHostToDevice, h->stream));
            CKB(cudaMemcpyAsync(dSc, g.scale, (size_t)Fg * sizeof(double), cudaMemcpyHostToDevice, h->stream));
            unexplained_kernel<<<dim3((Fg + 31) / 32, (Fg + 7) / 8), 256, 0, h->stream>>>(dSr, dEv, dSc, I.dUs, Fg, I.R);
            CKB(cudaGetLastError());
            CKB(cudaStreamSynchronize(h->stream));
            cudaFree(dSr); cudaFree(dEv); cudaFree(dSc);
        } else {
            CKB(cudaMemsetAsync(I.dUs, 0, (size_t)Fg * Fg * sizeof(double), h->stream));
        }
        CKB(cudaStreamSynchronize(h->stream));
        cudaFree(dL); cudaFree(dX); cudaFree(dAlpha);
    }
    h->Ptot = pc_off;
    h->K0 = ((h->Ptot + 3) / 4) * 4;
    h->nks0 = h->K0 / 4;

    // ---- PC descriptors ---------------------------------------------------------------------------------------
    {
        std::vector<PcDesc> pcs((size_t)h->Ptot);
        for (int gi = 0; gi < h->G; ++gi) {
            const bi_group_desc& g = D->groups[gi];
            const GroupImg& I = h->groups[gi];
            for (int p = 0; p < I.P; ++p) {
                PcDesc& pc = pcs[(size_t)I.pc_off + p];
                pc.XsT = I.dXsT + (size_t)p * d * I.N;
                pc.ell = I.dEll + (size_t)p * d;
                pc.LinvF = I.dLinvF + (size_t)p * I.ntn * 4 * I.nks * 32;
                pc.N = I.N; pc.nks = I.nks; pc.ntn = I.ntn; pc.kind = g.kind[p];
                pc.cst = g.constant[p];
                pc.diag = (1.0 + g.constant[p]) + g.noise[p];
                pc.group = gi; pc.local = p;
            }
        }
        TRY(upload(h, &h->dPcs, pcs.data(), pcs.size()));
    }

    // ---- merged layout (emulation.py:346-406) ---------------------------------------------------------------
    std::vector<int> row_group((size_t)F, -1), row_in((size_t)F, -1), obs((size_t)Fp, 0);
    {
        std::vector<std::array<long long, 4>> blocks((size_t)h->B);
        for (int b = 0; b < h->B; ++b)
            for (int k = 0; k < 4; ++k) blocks[b][k] = D->blocks[(size_t)b * 4 + k];
        std::sort(blocks.begin(), blocks.end(), [](const auto& a, const auto& b) { return a[1] < b[1]; });
        h->blk = blocks;
        for (int b = 0; b < h->B; ++b) {
            const long long gi = blocks[b][0], so = blocks[b][1], si = blocks[b][2], ln = blocks[b][3];
            if (gi < 0 || gi >= h->G || ln <= 0 || so < 0 || so + ln > F || si < 0 || si + ln > h->groups[gi].Fg) {
                h->err = "observable block outside its matrix";
                return bail(BI_E_ARG);
            }
            for (long long r = 0; r < ln; ++r) {
                if (row_group[so + r] != -1) {
                    h->err = "observable blocks overlap";
                    return bail(BI_E_ARG);
                }
                row_group[so + r] = (int)gi; row_in[so + r] = (int)(si + r); obs[so + r] = b;
            }
        }
        for (int r = 0; r < F; ++r)
            if (row_group[r] < 0) {
                h->err = "observable blocks do not cover the merged feature axis";
                return bail(BI_E_ARG);
            }
        for (int r = F; r < Fp; ++r) obs[r] = h->B + (r - F);
        TRY(dalloc(h, &h->dUblk, (size_t)F * F));
        CKB(cudaMemsetAsync(h->dUblk, 0, (size_t)F * F * sizeof(double), h->stream));
        for (int b = 0; b < h->B; ++b) {
            const long long gi = blocks[b][0], so = blocks[b][1], si = blocks[b][2], ln = blocks[b][3];
            const GroupImg& I = h->groups[gi];
            CKB(cudaMemcpy2DAsync(h->dUblk + (size_t)so * F + so, (size_t)F * sizeof(double),
                                  I.dUs + (size_t)si * I.Fg + si, (size_t)I.Fg * sizeof(double),
                                  (size_t)ln * sizeof(double), (size_t)ln, cudaMemcpyDeviceToDevice, h->stream));
        }
    }
    {
        const int K0 = h->K0, nks0 = h->nks0;
        std::vector<double> trow((size_t)F * K0, 0.0), tt((size_t)K0 * Fp, 0.0), tm((size_t)nt * 4 * nks0 * 32, 0.0),
            mean((size_t)F, 0.0), ev((size_t)Fp, 0.0);
        for (int r = 0; r < F; ++r) {
            const int gi = row_group[r], ri = row_in[r];
            const bi_group_desc& g = D->groups[gi];
            const GroupImg& I = h->groups[gi];
            for (int p = 0; p < I.P; ++p) {
                const double tv = g.scale[ri] * g.S[(size_t)p * I.Fg + ri];
                const int q = I.pc_off + p;
                trow[(size_t)r * K0 + q] = tv;
                tt[(size_t)q * Fp + r] = tv;
                tm[(size_t)(r / 32) * 4 * nks0 * 32 + of_index(r % 32, q, nks0)] = tv;
            }
            mean[r] = g.mean[ri];
            if (D->y) ev[r] = g.mean[ri] - D->y[r];
        }
        TRY(upload(h, &h->dTrow, trow.data(), trow.size()));
        TRY(upload(h, &h->dTt, tt.data(), tt.size()));
        TRY(upload(h, &h->dTm, tm.data(), tm.size()));
        h->nTc = (K0 + 31) / 32;
        {
            std::vector<double> tmp((size_t)nt * h->nTc * TILE_ELEMS, 0.0);
            for (int r = 0; r < F; ++r)
                for (int q = 0; q < h->Ptot; ++q)
                    tmp[((size_t)(r / 32) * h->nTc + q / 32) * TILE_ELEMS + of2_index(r % 32, q % 32)] = trow[(size_t)r * K0 + q];
            TRY(upload(h, &h->dTmP, tmp.data(), tmp.size()));
        }
        TRY(upload(h, &h->dMeanMerged, mean.data(), mean.size()));
        h->hMeanMerged = mean;
        TRY(upload(h, &h->dE, ev.data(), ev.size()));
        CKB(cudaStreamSynchronize(h->stream));
    }
    {
        std::vector<int> omin((size_t)nt), omax((size_t)nt), kminD((size_t)nt, 0), imaxD((size_t)nt, nt - 1),
            kminS((size_t)nt), imaxS((size_t)nt);
        for (int i = 0; i < nt; ++i) { omin[i] = obs[(size_t)i * 32]; omax[i] = obs[(size_t)i * 32 + 31]; }
        {   // share of the merged covariance's lower tiles that hold an observable block (the rest is zero-filled by K3)
            const int nt3 = (F + 31) / 32;
            long long hit = 0;
            for (int i = 0; i < nt3; ++i)
                for (int j = 0; j <= i; ++j) hit += omax[j] >= omin[i];
            h->k3_block_share = (double)hit / ((double)nt3 * (nt3 + 1) / 2);
        }
        for (int i = 0; i < nt; ++i) {
            int k = 0;
            while (omax[k] < omin[i]) ++k;
            kminS[i] = k;
        }
        for (int j = 0; j < nt; ++j) {
            int i = j;
            while (i + 1 < nt && kminS[i + 1] <= j) ++i;
            imaxS[j] = i;
        }
        h->hOmin = omin; h->hOmax = omax;
        h->hKmin[0] = kminD; h->hImax[0] = imaxD; h->hKmin[1] = kminS; h->hImax[1] = imaxS;
        TRY(upload(h, &h->dObs, obs.data(), obs.size()));
        TRY(upload(h, &h->dTileOmin, omin.data(), omin.size()));
        TRY(upload(h, &h->dTileOmax, omax.data(), omax.size()));
        TRY(upload(h, &h->dKminDense, kminD.data(), kminD.size()));
        TRY(upload(h, &h->dImaxDense, imaxD.data(), imaxD.size()));
        TRY(upload(h, &h->dKminStruct, kminS.data(), kminS.size()));
        TRY(upload(h, &h->dImaxStruct, imaxS.data(), imaxS.size()));
        CKB(cudaStreamSynchronize(h->stream));
    }

    // ---- experimental data ------------------------------------------------------------------------------------
    h->has_data = (D->y != nullptr) && (D->y_err != nullptr || D->sigma_exp != nullptr);
    if (h->has_data) {
        if (D->sigma_exp) {
            TRY(upload(h, &h->dSig, D->sigma_exp, (size_t)F * F));
            for (int r = 0; r < F && h->sigma_blockdiag; ++r)
                for (int c = 0; c < F; ++c)
                    if (obs[r] != obs[c] && D->sigma_exp[(size_t)r * F + c] != 0.0) { h->sigma_blockdiag = false; break; }
        } else {
            TRY(upload(h, &h->dYerr, D->y_err, (size_t)F));
        }
        for (int i = 0; i < bi_handle_s::A_SLOTS; ++i) TRY(dalloc(h, &h->dASlot[i], (size_t)nt * (nt + 1) / 2 * TILE_ELEMS));
        h->dATiles = h->dASlot[0];
        CKB(cudaStreamSynchronize(h->stream));
    }

    // ---- per-call scratch ---------------------------------------------------------------------------------------
    const long long mb = h->max_batch;
    TRY(dalloc(h, &h->dTheta, (size_t)mb * d));
    TRY(dalloc(h, &h->dM, (size_t)mb * h->Ptot));
    TRY(dalloc(h, &h->dV, (size_t)mb * h->Ptot));
    TRY(dalloc(h, &h->dLogp, (size_t)mb));
    TRY(dalloc(h, &h->dInfo, (size_t)mb));
    TRY(dalloc(h, &h->dCount, (size_t)1));

    h->gp_wt = 32;   // measured (C3, B200): 32 walkers x 2 CTAs/SM 0.124 ms, 64 walkers x 1 CTA/SM 0.138 ms per 1024 walkers
    if (const char* e = std::getenv("BI_GP_WT")) h->gp_wt = std::atoi(e) == 64 ? 64 : 32;   // developer A/B knob
    if (const char* e = std::getenv("BI_GP_SMALL")) h->gp_small = std::atoi(e);   // 0 off, 16: 16-walker tiles, else 8-walker tiles
    if (const char* e = std::getenv("BI_GP_MIXED")) h->gp_mixed = std::atoi(e);
    if (const char* e = std::getenv("BI_PDL")) h->pdl = std::atoi(e) != 0;
    if (gp_smem_bytes(d, h->Nmax, h->gp_wt) > 227 * 1024) h->gp_wt = 32;
    h->gp_smem = gp_smem_bytes(d, h->Nmax, h->gp_wt);
    if (h->gp_smem > 227 * 1024) {
        h->err = "n_design too large for the GP kernel's shared-memory operand tile";
        return bail(BI_E_UNSUPPORTED);
    }
    if (h->gp_wt == 64) {
        CKB(allow_max_smem(gp_predict_kernel<true, 64>));
        CKB(allow_max_smem(gp_predict_kernel<false, 64>));
    } else {
        CKB(allow_max_smem(gp_predict_kernel<true, 32>));
        CKB(allow_max_smem(gp_predict_kernel<false, 32>));
        CKB(allow_max_smem(gp_predict_kernel<true, 8>));
        CKB(allow_max_smem(gp_predict_mixed_kernel));
    }
    CKB(allow_max_smem(cov_persist_kernel));
    if (const char* e = std::getenv("BI_K3_PERSIST")) h->k3_persist = std::atoi(e) != 0;   // developer A/B knob
    {
        const size_t cs = (size_t)std::max(h->K0, 4) * 33 * sizeof(double);
        if (cs > 48 * 1024) CKB(allow_max_smem(cov_assemble_kernel));
    }
    if (h->has_data) {
        h->chol_smem = chol_smem_bytes(Fp, h->K0);
        CKB(allow_max_smem(chol_loglik_kernel<false>));
        int occ = 0;
        CKB(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, chol_loglik_kernel<false>, CHOL_THREADS, h->chol_smem));
        if (occ < 1) {
            h->err = "Cholesky kernel does not fit on an SM";
            return bail(BI_E_UNSUPPORTED);
        }
        if (const char* e = std::getenv("BI_CHOL_IMPL")) h->chol_impl = std::atoi(e) == 1 ? 1 : 2;
        if (const char* e = std::getenv("BI_PIPE_SMALL")) h->pipe_small = std::atoi(e) != 0;   // developer A/B knob
        h->pipe_smem = pipe_smem_bytes(Fp, h->nTc * 32);
        if (h->chol_impl == 2 && (h->pipe_smem > 227 * 1024 || nt > 120)) h->chol_impl = 1;
        if (h->chol_impl == 2) {
            CKB(allow_max_smem(chol_pipe_kernel<PIPE_CONSUMERS_MAIN>));
            CKB(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, chol_pipe_kernel<PIPE_CONSUMERS_MAIN>, pipe_threads(PIPE_CONSUMERS_MAIN),
                                                              h->pipe_smem));
            if (h->pipe_small != 0) {
                int occ_s = 0;
                h->pipe_smem_small = pipe_smem_bytes(Fp, h->nTc * 32, PIPE_CONSUMERS_SMALL);
                CKB(allow_max_smem(chol_pipe_kernel<PIPE_CONSUMERS_SMALL>));
                CKB(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_s, chol_pipe_kernel<PIPE_CONSUMERS_SMALL>,
                                                                  pipe_threads(PIPE_CONSUMERS_SMALL), h->pipe_smem_small));
                h->chol_ctas_small = (nt < 10 || h->pipe_small == 1) ? h->sm_count * occ_s : 0;   // scratch is sized by it
            }
            h->pipe_smem_wide = pipe_smem_bytes(Fp, h->nTc * 32, PIPE_CONSUMERS_WIDE);
            h->pipe_wide = h->pipe_smem_wide <= 227 * 1024;
            if (const char* e = std::getenv("BI_PIPE_WIDE")) { if (std::atoi(e) == 0) h->pipe_wide = 0; }   // developer A/B knob
            if (h->pipe_wide)
                CKB(allow_max_smem(chol_pipe_kernel<PIPE_CONSUMERS_WIDE>));
            if (occ < 1) {
                h->err = "pipelined Cholesky kernel does not fit on an SM";
                return bail(BI_E_UNSUPPORTED);
            }
        }
        if (const char* e = std::getenv("BI_CHOL_OCC")) occ = std::max(1, std::min(occ, std::atoi(e)));   // developer A/B knob
        h->chol_ctas = h->sm_count * occ;
        h->ws_stride = chol_ws_doubles(nt);
        TRY(dalloc(h, &h->dWs, (size_t)std::max(h->chol_ctas, h->chol_ctas_small) * h->ws_stride));
    }
    CKB(cudaStreamSynchronize(h->stream));
    *out = h;
    return 0;
#undef TRY
#undef CKB
}

// ------------------------------------------------------------------------------------------------------------------
int bi_kernel_vectors(bi_handle h, int group, const double* theta, int64_t W, double* k_out, void* stream) {
    if (!h || !theta || !k_out || W < 0 || group < 0 || group >= h->G) return fail(h, BI_E_ARG, "bad argument");
    if (W == 0) return 0;
    CK(cudaSetDevice(h->device));
    cudaStream_t s = (cudaStream_t)stream;
    const GroupImg& I = h->groups[group];
    const int wt = h->gp_wt;
    dim3 grid((unsigned)((W + wt - 1) / wt), (unsigned)I.P);
    if (wt == 64)
        gp_predict_kernel<false, 64><<<grid, 512, h->gp_smem, s>>>(h->dPcs, I.pc_off, h->d, theta, W, k_out, I.P, nullptr, nullptr, h->Ptot);
    else
        gp_predict_kernel<false, 32><<<grid, 256, h->gp_smem, s>>>(h->dPcs, I.pc_off, h->d, theta, W, k_out, I.P, nullptr, nullptr, h->Ptot);
    h->launches++;
    CK(cudaGetLastError());
    return 0;
}

int bi_kernel_vectors_host(bi_handle h, int group, const double* theta, int64_t W, double* k_out) {
    if (!h || !theta || !k_out || W < 0 || group < 0 || group >= h->G) return fail(h, BI_E_ARG, "bad argument");
    CK(cudaSetDevice(h->device));
    const GroupImg& I = h->groups[group];
    const size_t per = (size_t)I.P * I.N;
    const long long chunk = std::min<long long>(h->max_batch, std::max<long long>(1, (256ll << 20) / (long long)(per * 8)));
    int rc = ensure_out_scratch(h, (size_t)chunk * per * 8);
    if (rc) return rc;
    for (long long w0 = 0; w0 < W; w0 += chunk) {
        const long long wb = std::min<long long>(chunk, W - w0);
        CK(cudaMemcpyAsync(h->dTheta, theta + w0 * h->d, (size_t)wb * h->d * 8, cudaMemcpyHostToDevice, h->stream));
        rc = bi_kernel_vectors(h, group, h->dTheta, wb, h->dOutScratch, h->stream);
        if (rc) return rc;
        CK(cudaMemcpyAsync(k_out + (size_t)w0 * per, h->dOutScratch, (size_t)wb * per * 8, cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
    }
    return 0;
}

int bi_predict_pc(bi_handle h, const double* theta, int64_t W, double* m_out, double* v_out, void* stream) {
    if (!h || !theta || !m_out || !v_out || W < 0) return fail(h, BI_E_ARG, "bad argument");
    if (W == 0) return 0;
    CK(cudaSetDevice(h->device));
    return launch_gp(h, theta, W, m_out, v_out, (cudaStream_t)stream);
}

int bi_predict_pc_host(bi_handle h, const double* theta, int64_t W, double* m_out, double* v_out) {
    if (!h || !theta || !m_out || !v_out || W < 0) return fail(h, BI_E_ARG, "bad argument");
    CK(cudaSetDevice(h->device));
    for (long long w0 = 0; w0 < W; w0 += h->max_batch) {
        const long long wb = std::min<long long>(h->max_batch, W - w0);
        CK(cudaMemcpyAsync(h->dTheta, theta + w0 * h->d, (size_t)wb * h->d * 8, cudaMemcpyHostToDevice, h->stream));
        int rc = launch_gp(h, h->dTheta, wb, h->dM, h->dV, h->stream);
        if (rc) return rc;
        CK(cudaMemcpyAsync(m_out + w0 * h->Ptot, h->dM, (size_t)wb * h->Ptot * 8, cudaMemcpyDeviceToHost, h->stream));
        CK(cudaMemcpyAsync(v_out + w0 * h->Ptot, h->dV, (size_t)wb * h->Ptot * 8, cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
    }
    return 0;
}

static int predict_any(bi_handle h, int group, const double* theta, int64_t W, double n_div, double* mean_out,
                       double* cov_out, cudaStream_t s, double* dM, double* dV) {
    // theta, outputs: device pointers; dM/dV: (W, Ptot) scratch
    int rc = launch_gp(h, theta, W, dM, dV, s);
    if (rc) return rc;
    CovParams P;
    if (group < 0) {
        P.Fo = h->F; P.K0 = h->K0; P.voff = 0; P.nq = h->Ptot; P.Trow = h->dTrow; P.Uc = h->dUblk;
        P.mu = h->dMeanMerged; P.obs = h->dObs; P.tile_omin = h->dTileOmin; P.tile_omax = h->dTileOmax;
    } else {
        const GroupImg& I = h->groups[group];
        P.Fo = I.Fg; P.K0 = I.K0g; P.voff = I.pc_off; P.nq = I.P; P.Trow = I.dTg; P.Uc = I.dUs; P.mu = I.dMean;
        P.obs = nullptr; P.tile_omin = nullptr; P.tile_omax = nullptr;
    }
    P.Ptot = h->Ptot; P.m = dM; P.v = dV; P.n_div = n_div; P.mean_out = mean_out; P.cov_out = cov_out; P.W = W;
    if (mean_out) {
        const long long n = W * P.Fo;
        mean_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(P);
        h->launches++;
        CK(cudaGetLastError());
    }
    if (cov_out) {
        // Measured (1491 walkers, B200): one dense 600-bin block 1.36 ms with the (row tile, walker) grid, 1.03 ms persistent;
        // 20 observables x 10 bins (46 % of the tiles hold a block) 0.32 vs 0.21 ms; 30 x 20 bins (19 %) 1.00 vs 1.12 ms --
        // the persistent kernel wins wherever contraction matters, the grid kernel where the launch is mostly zero fill.
        const size_t psmem = k3p_smem_bytes(P.Fo, P.K0);
        const bool dense_layout = group >= 0 || h->k3_block_share >= 0.3;
        if ((h->k3_persist < 0 ? dense_layout : h->k3_persist != 0) && psmem <= K3P_SMEM_MAX) {
            const long long nt3 = (P.Fo + 31) / 32, items = W * (nt3 * (nt3 + 1) / 2);
            const long long ctas = std::min<long long>(h->sm_count, (items + K3P_THREADS / 32 - 1) / (K3P_THREADS / 32));
            cov_persist_kernel<<<(unsigned)ctas, K3P_THREADS, psmem, s>>>(P);
        } else {
            const size_t smem = (size_t)P.K0 * 33 * sizeof(double);     // v + the 32 x K0 operand tile (<= 48 KB for K0 <= 180)
            cov_assemble_kernel<<<dim3((P.Fo + 31) / 32, (unsigned)W), 256, smem, s>>>(P);
        }
        h->launches++;
        CK(cudaGetLastError());
    }
    return 0;
}

static int predict_dev(bi_handle h, int group, const double* theta, int64_t W, double n_div, double* mean_out,
                       double* cov_out, void* stream) {
    if (!h || !theta || W < 0 || group >= h->G || !(n_div > 0.0)) return fail(h, BI_E_ARG, "bad argument");
    if (W == 0) return 0;
    CK(cudaSetDevice(h->device));
    cudaStream_t s = (cudaStream_t)stream;
    const int Fo = group < 0 ? h->F : h->groups[group].Fg;
    for (long long w0 = 0; w0 < W; w0 += h->max_batch) {
        const long long wb = std::min<long long>(h->max_batch, W - w0);
        int rc = predict_any(h, group, theta + w0 * h->d, wb, n_div, mean_out ? mean_out + w0 * Fo : nullptr,
                             cov_out ? cov_out + (size_t)w0 * Fo * Fo : nullptr, s, h->dM, h->dV);
        if (rc) return rc;
    }
    return 0;
}

static int predict_host(bi_handle h, int group, const double* theta, int64_t W, double n_div, double* mean_out,
                        double* cov_out) {
    if (!h || !theta || W < 0 || group >= h->G || !(n_div > 0.0)) return fail(h, BI_E_ARG, "bad argument");
    CK(cudaSetDevice(h->device));
    const int Fo = group < 0 ? h->F : h->groups[group].Fg;
    const size_t per = (size_t)Fo * (cov_out ? (size_t)Fo + 1 : 1) * 8;
    const long long chunk = std::min<long long>(h->max_batch, std::max<long long>(1, (long long)((512ull << 20) / per)));
    int rc = ensure_out_scratch(h, (size_t)chunk * per);
    if (rc) return rc;
    double* dMean = h->dOutScratch;
    double* dCov = cov_out ? h->dOutScratch + (size_t)chunk * Fo : nullptr;
    for (long long w0 = 0; w0 < W; w0 += chunk) {
        const long long wb = std::min<long long>(chunk, W - w0);
        CK(cudaMemcpyAsync(h->dTheta, theta + w0 * h->d, (size_t)wb * h->d * 8, cudaMemcpyHostToDevice, h->stream));
        rc = predict_any(h, group, h->dTheta, wb, n_div, dMean, dCov, h->stream, h->dM, h->dV);
        if (rc) return rc;
        if (mean_out) CK(cudaMemcpyAsync(mean_out + w0 * Fo, dMean, (size_t)wb * Fo * 8, cudaMemcpyDeviceToHost, h->stream));
        if (cov_out)
            CK(cudaMemcpyAsync(cov_out + (size_t)w0 * Fo * Fo, dCov, (size_t)wb * Fo * Fo * 8, cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
    }
    return 0;
}

int bi_predict_group(bi_handle h, int group, const double* theta, int64_t W, double n_div, double* mean_out,
                     double* cov_out, void* stream) {
    if (group < 0) return fail(h, BI_E_ARG, "bad group");
    return predict_dev(h, group, theta, W, n_div, mean_out, cov_out, stream);
}
int bi_predict_group_host(bi_handle h, int group, const double* theta, int64_t W, double n_div, double* mean_out,
                          double* cov_out) {
    if (group < 0) return fail(h, BI_E_ARG, "bad group");
    return predict_host(h, group, theta, W, n_div, mean_out, cov_out);
}
int bi_predict(bi_handle h, const double* theta, int64_t W, double n_div, double* mean_out, double* cov_out, void* stream) {
    return predict_dev(h, -1, theta, W, n_div, mean_out, cov_out, stream);
}
int bi_predict_host(bi_handle h, const double* theta, int64_t W, double n_div, double* mean_out, double* cov_out) {
    return predict_host(h, -1, theta, W, n_div, mean_out, cov_out);
}

// the explicit-matrix Cholesky kernel is shared by _loglikelihood and Woodbury.  The limit is an attribute of the function
// on the CURRENT device (the caller has done cudaSetDevice): it is set unconditionally -- a process-wide high-water mark
// would skip it on a second device.
static cudaError_t raise_explicit_smem(size_t need) {
    if (need > 227 * 1024) return cudaErrorInvalidValue;
    return allow_max_smem(chol_loglik_kernel<true>);
}

// ---- Woodbury mode ---------------------------------------------------------------------------------------------------
static int ensure_woodbury(bi_handle h) {
    if (h->dWbBlocks) return 0;
    const int nB = (int)h->blk.size();
    h->wbBlocks.resize(nB);
    int Pmax = 0;
    for (int b = 0; b < nB; ++b) Pmax = std::max(Pmax, h->groups[h->blk[b][0]].P);
    if (Pmax > 64) return fail(h, BI_E_UNSUPPORTED, "Woodbury mode supports n_pc <= 64 per emulation group");
    h->wbPmax = Pmax;
    for (int b = 0; b < nB; ++b) {
        const GroupImg& I = h->groups[h->blk[b][0]];
        WbBlock& B = h->wbBlocks[b];
        B.row0 = (int)h->blk[b][1]; B.f = (int)h->blk[b][3]; B.col0 = I.pc_off; B.P = I.P;
        B.a_off = 0; B.y_off = 0; B.g_off = (long long)b * Pmax * Pmax;
    }
    int rc = upload(h, &h->dWbBlocks, h->wbBlocks.data(), h->wbBlocks.size());
    if (rc) return rc;
    if ((rc = dalloc(h, &h->dG, (size_t)nB * Pmax * Pmax))) return rc;
    if ((rc = dalloc(h, &h->dH, (size_t)nB * Pmax))) return rc;
    if ((rc = dalloc(h, &h->dQ0, (size_t)nB))) return rc;
    if ((rc = dalloc(h, &h->dLdA, (size_t)nB))) return rc;
    // per-call scratch: M matrices of a chunk of walkers
    const long long per = (long long)nB * Pmax * Pmax * 8;
    h->wbChunk = std::max<long long>(1, std::min<long long>(h->max_batch, (256ll << 20) / per));
    const long long slots = h->wbChunk * nB;
    if ((rc = dalloc(h, &h->dMw, (size_t)slots * Pmax * Pmax))) return rc;
    if ((rc = dalloc(h, &h->dUw, (size_t)slots * Pmax))) return rc;
    if ((rc = dalloc(h, &h->dQuadA, (size_t)slots))) return rc;
    if ((rc = dalloc(h, &h->dQuadM, (size_t)slots))) return rc;
    if ((rc = dalloc(h, &h->dLdM, (size_t)slots))) return rc;
    if ((rc = dalloc(h, &h->dLpW, (size_t)slots))) return rc;
    if ((rc = dalloc(h, &h->dInfoW, (size_t)slots))) return rc;
    h->wbNt = (Pmax + 31) / 32;
    std::vector<int> kd((size_t)2 * h->wbNt, 0);
    for (int i = 0; i < h->wbNt; ++i) kd[(size_t)h->wbNt + i] = h->wbNt - 1;
    if ((rc = upload(h, &h->dWbK, kd.data(), kd.size()))) return rc;
    h->wbSmem = chol_smem_bytes(h->wbNt * 32, 4);
    CK(raise_explicit_smem(h->wbSmem));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, chol_loglik_kernel<true>, CHOL_THREADS, h->wbSmem));
    if (occ < 1) return fail(h, BI_E_UNSUPPORTED, "Woodbury: n_pc too large for the Cholesky kernel");
    h->wbCtas = (int)std::min<long long>(slots, (long long)h->sm_count * occ);
    h->wbStride = chol_ws_doubles(h->wbNt);
    if ((rc = dalloc(h, &h->dWbWs, (size_t)h->wbCtas * h->wbStride))) return rc;
    CK(cudaStreamSynchronize(h->stream));
    return 0;
}

static int woodbury_setup(bi_handle h, double n_div, cudaStream_t s) {
    if (h->wb_ndiv == n_div) return 0;
    const int nB = (int)h->wbBlocks.size();
    long long a_tot = 0, y_tot = 0;
    for (int b = 0; b < nB; ++b) {
        WbBlock& B = h->wbBlocks[b];
        B.a_off = a_tot; B.y_off = y_tot;
        a_tot += (long long)B.f * B.f; y_tot += (long long)B.f * (B.P + 1);
    }
    double *dA = nullptr, *dY = nullptr;
    CK(cudaMalloc(&dA, (size_t)a_tot * 8));
    CK(cudaMalloc(&dY, (size_t)y_tot * 8));
    CK(cudaMemcpyAsync(h->dWbBlocks, h->wbBlocks.data(), sizeof(WbBlock) * nB, cudaMemcpyHostToDevice, s));
    woodbury_setup_kernel<<<nB, 256, 0, s>>>(h->dWbBlocks, h->F, h->K0, h->wbPmax, h->dSig, h->dYerr, h->dUblk, n_div, h->dTrow,
                                             h->dE, dA, dY, h->dG, h->dH, h->dQ0, h->dLdA);
    h->launches++;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaFree(dA); cudaFree(dY);
    if (e != cudaSuccess) { h->err = std::string("woodbury_setup: ") + cudaGetErrorString(e); return (int)e; }
    h->wb_ndiv = n_div;
    return 0;
}

// theta, logp, info: device pointers; one pass of gp_predict + build + P x P Cholesky + combine per chunk of walkers
static int run_woodbury(bi_handle h, const double* dTheta, long long W, double n_div, double* dLogp, int* dInfo, cudaStream_t s) {
    int rc = ensure_woodbury(h);
    if (rc) return rc;
    if ((rc = woodbury_setup(h, n_div, s))) return rc;
    const int nB = (int)h->wbBlocks.size(), Pm = h->wbPmax;
    for (long long w0 = 0; w0 < W; w0 += h->wbChunk) {
        const long long wb = std::min<long long>(h->wbChunk, W - w0);
        if ((rc = launch_gp(h, dTheta + w0 * h->d, wb, h->dM, h->dV, s))) return rc;
        woodbury_build_kernel<<<dim3(nB, (unsigned)wb), 64, 0, s>>>(h->dWbBlocks, nB, Pm, h->Ptot, h->dM, h->dV, h->dG, h->dH, h->dQ0,
                                                                     h->dMw, h->dUw, h->dQuadA);
        h->launches++;
        CK(cudaGetLastError());
        CholParams P;
        std::memset(&P, 0, sizeof(P));
        const long long slots = wb * nB;
        P.F = Pm; P.Fp = h->wbNt * 32; P.nt = h->wbNt; P.K0 = 4; P.nks0 = 1; P.d = h->d; P.W = slots;
        P.kmin = h->dWbK; P.imax = h->dWbK + h->wbNt; P.ws = h->dWbWs; P.ws_stride = h->wbStride;
        P.logp = h->dLpW; P.info = h->dInfoW; P.cov_x = h->dMw; P.dy_x = h->dUw; P.quad_out = h->dQuadM; P.logdet_out = h->dLdM;
        {
            ProfScope ps(h, 1, s);
            chol_loglik_kernel<true><<<(int)std::min<long long>(slots, h->wbCtas), CHOL_THREADS, h->wbSmem, s>>>(P);
        }
        h->launches++;
        CK(cudaGetLastError());
        woodbury_combine_kernel<<<(unsigned)((wb + 127) / 128), 128, 0, s>>>(nB, wb, h->d, dTheta + w0 * h->d, h->dMin, h->dMax, h->dQuadA,
                                                                             h->dQuadM, h->dLdM, h->dLdA, h->dInfoW, dLogp + w0,
                                                                             dInfo ? dInfo + w0 : nullptr);
        h->launches++;
        CK(cudaGetLastError());
    }
    return 0;
}

// ---- block-direct structured mode ---------------------------------------------------------------------------------------
// BI_MODE_STRUCTURED is evaluated observable block by observable block in registers (block_direct.cuh) when every block
// has at most 31 rows and its group at most BD_PMAX principal components; otherwise (or with BI_STRUCT_IMPL=tiles, the
// developer A/B knob) by the fused Cholesky kernel skipping structurally-zero tiles.
static int ensure_block_direct(bi_handle h) {
    if (h->bd_state != 0) return 0;
    h->bd_state = -1;
    if (const char* e = std::getenv("BI_STRUCT_IMPL")) { if (std::string(e) == "tiles") return 0; }
    const int nB = (int)h->blk.size();
    int fmax = 0;
    long long a_tot = 0;
    h->bdBlocks.resize(nB);
    for (int b = 0; b < nB; ++b) {
        const GroupImg& I = h->groups[h->blk[b][0]];
        BdBlock& B = h->bdBlocks[b];
        B.row0 = (int)h->blk[b][1]; B.f = (int)h->blk[b][3]; B.col0 = I.pc_off; B.P = I.P; B.a_off = a_tot;
        a_tot += (long long)B.f * B.f;
        fmax = std::max(fmax, B.f);
        if (B.f > 31 || B.P > BD_PMAX) return 0;
    }
    h->bdFB = (fmax + 7) / 8 * 8;
    h->bd_a_tot = a_tot;
    int pmax = 0;
    for (const BdBlock& B : h->bdBlocks) pmax = std::max(pmax, B.P);
    h->bdNks = (pmax + 3) / 4;
    if (block_direct_smem_bytes(h->bdFB, h->bdNks) > 113 * 1024) return 0;   // two CTAs per SM must fit (108.7 KB at FB = 32, P = 64)
    {
        cudaError_t e_ = cudaSuccess;
        h->bdAugIn = fmax < h->bdFB;   // a free row in the last fragment for the augmented row
        if (const char* e = std::getenv("BI_BD_AUGIN")) h->bdAugIn = h->bdAugIn && std::atoi(e) != 0;   // developer A/B knob
        switch (h->bdFB + (h->bdAugIn ? 1 : 0)) {
            case 8: e_ = allow_max_smem(block_direct_kernel<8, false>); break;
            case 9: e_ = allow_max_smem(block_direct_kernel<8, true>); break;
            case 16: e_ = allow_max_smem(block_direct_kernel<16, false>); break;
            case 17: e_ = allow_max_smem(block_direct_kernel<16, true>); break;
            case 24: e_ = allow_max_smem(block_direct_kernel<24, false>); break;
            case 25: e_ = allow_max_smem(block_direct_kernel<24, true>); break;
            default: e_ = allow_max_smem(block_direct_kernel<32, true>); break;
        }
        if (e_ != cudaSuccess) { h->err = std::string("block_direct smem: ") + cudaGetErrorString(e_); return (int)e_; }
    }
    int rc = upload(h, &h->dBdBlocks, h->bdBlocks.data(), h->bdBlocks.size());
    if (rc) return rc;
    for (int i = 0; i < 2; ++i)
        if ((rc = dalloc(h, &h->dBdA[i], (size_t)a_tot))) return rc;
    if ((rc = dalloc(h, &h->dBdPart, (size_t)h->max_batch * nB))) return rc;
    if ((rc = dalloc(h, &h->dBdBad, (size_t)h->max_batch * nB))) return rc;
    if ((rc = dalloc(h, &h->dBdCount, (size_t)h->max_batch))) return rc;
    CK(cudaMemsetAsync(h->dBdCount, 0, (size_t)h->max_batch * sizeof(int), h->stream));   // after the guard-mode poison, same stream
    if (const char* e = std::getenv("BI_BD_FUSE")) h->bdFuse = std::atoi(e) != 0;
    CK(cudaStreamSynchronize(h->stream));
    h->bd_state = 1;
    return 0;
}

static int run_block_direct(bi_handle h, const double* dTheta, long long W, double n_div, double* dLogp, int* dInfo, cudaStream_t s,
                            const int* dSetId = nullptr) {
    const int nB = (int)h->bdBlocks.size();
    const int slot = (n_div == 1.0) ? 0 : 1;
    const double *dAall = h->dBdA[slot], *dEall = h->dE;
    long long e_stride = 0, a_stride = 0;
    if (dSetId) {
        if (!h->dBdASets) CK(cudaMalloc(&h->dBdASets, (size_t)h->n_sets * h->bd_a_tot * 8));
        if (h->bd_sets_ndiv != n_div) {
            for (int c = 0; c < h->n_sets; ++c) {
                block_a_kernel<<<nB, 256, 0, s>>>(h->dBdBlocks, h->F, nullptr, h->dYerrSets + (size_t)c * h->F, h->dUblk, n_div,
                                                  h->dBdASets + (size_t)c * h->bd_a_tot);
                h->launches++;
            }
            CK(cudaGetLastError());
            h->bd_sets_ndiv = n_div;
        }
        dAall = h->dBdASets; dEall = h->dEsets; e_stride = h->Fp; a_stride = h->bd_a_tot;
    } else if (h->bd_ndiv[slot] != n_div) {
        block_a_kernel<<<nB, 256, 0, s>>>(h->dBdBlocks, h->F, h->dSig, h->dYerr, h->dUblk, n_div, h->dBdA[slot]);
        h->launches++;
        CK(cudaGetLastError());
        h->bd_ndiv[slot] = n_div;
    }
    const size_t smem = block_direct_smem_bytes(h->bdFB, h->bdNks);
    for (long long w0 = 0; w0 < W; w0 += h->max_batch) {
        const long long wb = std::min<long long>(h->max_batch, W - w0);
        int rc = launch_gp(h, dTheta + w0 * h->d, wb, h->dM, h->dV, s);
        if (rc) return rc;
        // walkers per warp: enough (walker, block) tasks per launch to fill the machine a few times over, at most 8
        const long long tasks = wb * nB;
        int wpw = (int)std::max<long long>(1, std::min<long long>(8, tasks / ((long long)h->sm_count * 2 * BD_WARPS * 4)));
        const dim3 grid((unsigned)nB, (unsigned)((wb + (long long)BD_WARPS * wpw - 1) / ((long long)BD_WARPS * wpw)));
        // The combine step rides in the block kernel only for launches of at most one wave of (walker, block) tasks (a sampler
        // half-step of a 100-walker ensemble): there it saves a launch (C1, 100 walkers: 18.9 -> 15.9 us); on a full batch
        // the serial tail and the fences cost more than the 3 us kernel they replace (C3a, 1024 walkers: 0.103 -> 0.113 ms).
        const bool fuse_now = h->bdFuse && tasks <= (long long)h->sm_count * 3 * BD_WARPS;
        BdFuse fuse;
        fuse.counter = fuse_now ? h->dBdCount : nullptr;
        fuse.theta = dTheta + w0 * h->d; fuse.lo = h->dMin; fuse.hi = h->dMax; fuse.logp = dLogp + w0;
        fuse.info = dInfo ? dInfo + w0 : nullptr; fuse.d = h->d;
        {
            ProfScope ps(h, 1, s);
#define BD_LAUNCH(FBV, INV)                                                                                                  \
    CK(launch_k(h->pdl, block_direct_kernel<FBV, INV>, grid, dim3(BD_THREADS), smem, s, h->dBdBlocks, nB, wb, wpw, h->Ptot, h->K0,       \
                h->bdNks, h->dM, h->dV, h->dTrow, dEall, dAall, h->dBdPart, h->dBdBad, dSetId ? dSetId + w0 : (const int*)nullptr,      \
                e_stride, a_stride, fuse))
            switch (h->bdFB + (h->bdAugIn ? 1 : 0)) {
                case 8: BD_LAUNCH(8, false); break;
                case 9: BD_LAUNCH(8, true); break;
                case 16: BD_LAUNCH(16, false); break;
                case 17: BD_LAUNCH(16, true); break;
                case 24: BD_LAUNCH(24, false); break;
                case 25: BD_LAUNCH(24, true); break;
                default: BD_LAUNCH(32, true); break;
            }
#undef BD_LAUNCH
            if (!fuse_now)
                CK(launch_k(h->pdl, block_combine_kernel, dim3((unsigned)((wb + 127) / 128)), dim3(128), 0, s, nB, wb, h->d,
                            dTheta + w0 * h->d, h->dMin, h->dMax, h->dBdPart, h->dBdBad, dLogp + w0, dInfo ? dInfo + w0 : (int*)nullptr));
        }
        h->launches += fuse_now ? 1 : 2;
        CK(cudaGetLastError());
    }
    return 0;
}

// ------------------------------------------------------------------------------------------------------------------
static int check_lp(bi_handle h, const double* theta, int64_t W, int mode, double* logp_out) {
    if (!h || !theta || !logp_out || W < 0) return fail(h, BI_E_ARG, "bad argument");
    if (!h->has_data) return fail(h, BI_E_STATE, "handle was created without experimental data");
    if (mode != BI_MODE_DENSE && mode != BI_MODE_STRUCTURED && mode != BI_MODE_WOODBURY) return fail(h, BI_E_ARG, "unknown mode");
    if (mode != BI_MODE_DENSE && !h->sigma_blockdiag)
        return fail(h, BI_E_STRUCTURE, "Sigma_exp couples observable blocks: structured / Woodbury mode is not valid");
    return 0;
}

int bi_log_posterior(bi_handle h, const double* theta, int64_t W, double n_div, int mode, double* logp_out,
                     int32_t* info_out, void* stream) {
    int rc = check_lp(h, theta, W, mode, logp_out);
    if (rc) return rc;
    if (W == 0) return 0;
    CK(cudaSetDevice(h->device));
    cudaStream_t s = (cudaStream_t)stream;
    if (!(n_div > 0.0)) {
        // rows inside the prior box in this call (log_posterior.py:67 -> emulation.py:493); needs one readback
        CK(cudaMemsetAsync(h->dCount, 0, sizeof(unsigned long long), s));
        count_inside_kernel<<<(unsigned)((W + 255) / 256), 256, 0, s>>>(theta, W, h->d, h->dMin, h->dMax, h->dCount);
        h->launches++;
        unsigned long long cnt = 0;
        CK(cudaMemcpyAsync(&cnt, h->dCount, sizeof(cnt), cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
        n_div = cnt > 0 ? (double)cnt : 1.0;
    }
    if (mode == BI_MODE_WOODBURY) return run_woodbury(h, theta, W, n_div, logp_out, info_out, s);
    if (mode == BI_MODE_STRUCTURED) {
        if ((rc = ensure_block_direct(h))) return rc;
        if (h->bd_state == 1) return run_block_direct(h, theta, W, n_div, logp_out, info_out, s);
    }
    rc = build_a_tiles(h, n_div, s);
    if (rc) return rc;
    for (long long w0 = 0; w0 < W; w0 += h->max_batch) {
        const long long wb = std::min<long long>(h->max_batch, W - w0);
        rc = launch_gp(h, theta + w0 * h->d, wb, h->dM, h->dV, s);
        if (rc) return rc;
        rc = launch_chol(h, theta + w0 * h->d, h->dM, h->dV, wb, mode, logp_out + w0, info_out ? info_out + w0 : nullptr, s);
        if (rc) return rc;
    }
    return 0;
}

int bi_log_posterior_host(bi_handle h, const double* theta, int64_t W, double n_div, int mode, double* logp_out,
                          int32_t* info_out) {
    int rc = check_lp(h, theta, W, mode, logp_out);
    if (rc) return rc;
    if (W == 0) return 0;
    CK(cudaSetDevice(h->device));
    if (!(n_div > 0.0)) {
        long long cnt = 0;
        for (long long w = 0; w < W; ++w) {
            bool in = true;
            for (int k = 0; k < h->d; ++k) {
                const double x = theta[w * h->d + k];
                if (!(x > h->hMin[k] && x < h->hMax[k])) in = false;
            }
            cnt += in;
        }
        n_div = cnt > 0 ? (double)cnt : 1.0;
    }
    bool direct = false;
    if (mode == BI_MODE_STRUCTURED) {
        if ((rc = ensure_block_direct(h))) return rc;
        direct = h->bd_state == 1;
    }
    if (mode != BI_MODE_WOODBURY && !direct) {
        rc = build_a_tiles(h, n_div, h->stream);
        if (rc) return rc;
    }
    for (long long w0 = 0; w0 < W; w0 += h->max_batch) {
        const long long wb = std::min<long long>(h->max_batch, W - w0);
        CK(cudaMemcpyAsync(h->dTheta, theta + w0 * h->d, (size_t)wb * h->d * 8, cudaMemcpyHostToDevice, h->stream));
        if (mode == BI_MODE_WOODBURY) {
            rc = run_woodbury(h, h->dTheta, wb, n_div, h->dLogp, h->dInfo, h->stream);
            if (rc) return rc;
        } else if (direct) {
            rc = run_block_direct(h, h->dTheta, wb, n_div, h->dLogp, h->dInfo, h->stream);
            if (rc) return rc;
        } else {
            rc = launch_gp(h, h->dTheta, wb, h->dM, h->dV, h->stream);
            if (rc) return rc;
            rc = launch_chol(h, h->dTheta, h->dM, h->dV, wb, mode, h->dLogp, h->dInfo, h->stream);
            if (rc) return rc;
        }
        CK(cudaMemcpyAsync(logp_out + w0, h->dLogp, (size_t)wb * 8, cudaMemcpyDeviceToHost, h->stream));
        if (info_out) CK(cudaMemcpyAsync(info_out + w0, h->dInfo, (size_t)wb * 4, cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
    }
    return 0;
}

int bi_loglikelihood_host(bi_handle h, const double* dy, const double* cov, int64_t n, int32_t Fdim, double* out,
                          int32_t* info_out) {
    if (!h || !dy || !cov || !out || n < 0 || Fdim <= 0) return fail(h, BI_E_ARG, "bad argument");
    if (n == 0) return 0;
    CK(cudaSetDevice(h->device));
    const int F = Fdim, nt = (F + 31) / 32, Fp = nt * 32;
    const size_t smem = chol_smem_bytes(Fp, 4);
    CK(raise_explicit_smem(smem));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, chol_loglik_kernel<true>, CHOL_THREADS, smem));
    if (occ < 1) return fail(h, BI_E_UNSUPPORTED, "matrix too large for the Cholesky kernel's shared memory");
    const long long per = (long long)F * F * 8;
    const long long chunk = std::max<long long>(1, std::min<long long>(n, (512ll << 20) / per));
    const int ctas = (int)std::min<long long>(chunk, (long long)h->sm_count * occ);
    const long long stride = chol_ws_doubles(nt);
    double *dCov = nullptr, *dDy = nullptr, *dWs = nullptr, *dLp = nullptr;
    int *dInfo = nullptr, *dK = nullptr;
    std::vector<int> kd((size_t)2 * nt, 0);
    for (int i = 0; i < nt; ++i) kd[(size_t)nt + i] = nt - 1;
    cudaError_t e = cudaSuccess;
    auto A = [&](void** p, size_t b) { if (e == cudaSuccess) e = cudaMalloc(p, b); };
    A((void**)&dCov, (size_t)chunk * per); A((void**)&dDy, (size_t)chunk * F * 8); A((void**)&dWs, (size_t)ctas * stride * 8);
    A((void**)&dLp, (size_t)chunk * 8); A((void**)&dInfo, (size_t)chunk * 4); A((void**)&dK, kd.size() * 4);
    int rc = 0;
    if (e == cudaSuccess) e = cudaMemcpyAsync(dK, kd.data(), kd.size() * 4, cudaMemcpyHostToDevice, h->stream);
    for (long long w0 = 0; w0 < n && e == cudaSuccess; w0 += chunk) {
        const long long wb = std::min<long long>(chunk, n - w0);
        e = cudaMemcpyAsync(dCov, cov + (size_t)w0 * F * F, (size_t)wb * per, cudaMemcpyHostToDevice, h->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(dDy, dy + (size_t)w0 * F, (size_t)wb * F * 8, cudaMemcpyHostToDevice, h->stream);
        if (e != cudaSuccess) break;
        CholParams P;
        std::memset(&P, 0, sizeof(P));
        P.F = F; P.Fp = Fp; P.nt = nt; P.K0 = 4; P.nks0 = 1; P.Ptot = 0; P.d = h->d; P.W = wb;
        P.kmin = dK; P.imax = dK + nt; P.ws = dWs; P.ws_stride = stride; P.logp = dLp; P.info = dInfo;
        P.cov_x = dCov; P.dy_x = dDy;
        chol_loglik_kernel<true><<<(int)std::min<long long>(wb, ctas), CHOL_THREADS, smem, h->stream>>>(P);
        h->launches++;
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaMemcpyAsync(out + w0, dLp, (size_t)wb * 8, cudaMemcpyDeviceToHost, h->stream);
        if (e == cudaSuccess && info_out) e = cudaMemcpyAsync(info_out + w0, dInfo, (size_t)wb * 4, cudaMemcpyDeviceToHost, h->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    }
    if (e != cudaSuccess) { h->err = std::string("bi_loglikelihood_host: ") + cudaGetErrorString(e); rc = (int)e; }
    cudaFree(dCov); cudaFree(dDy); cudaFree(dWs); cudaFree(dLp); cudaFree(dInfo); cudaFree(dK);
    return rc;
}

// ---- stretch-move half-step around the log-posterior ---------------------------------------------------
int bi_stretch_propose(bi_handle h, const double* coords, const int32_t* s_idx, const int32_t* c_idx, const double* zz,
                       int64_t ns, double* q_out, void* stream) {
    if (!h || !coords || !s_idx || !c_idx || !zz || !q_out || ns < 0) return fail(h, BI_E_ARG, "bad argument");
    if (ns == 0) return 0;
    CK(cudaSetDevice(h->device));
    const long long n = (long long)ns * h->d;
    CK(launch_k(h->pdl, stretch_propose_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, (cudaStream_t)stream, coords, s_idx, c_idx,
                zz, (long long)ns, h->d, q_out));
    h->launches++;
    CK(cudaGetLastError());
    return 0;
}

int bi_stretch_accept(bi_handle h, double* coords, double* logp, const int32_t* s_idx, const double* q, const double* new_lp,
                      const double* factors, const double* logu, int64_t ns, int64_t* n_accepted, int32_t* nan_flag,
                      void* stream) {
    if (!h || !coords || !logp || !s_idx || !q || !new_lp || !factors || !logu || !n_accepted || !nan_flag || ns < 0)
        return fail(h, BI_E_ARG, "bad argument");
    if (ns == 0) return 0;
    CK(cudaSetDevice(h->device));
    CK(launch_k(h->pdl, stretch_accept_kernel, dim3((unsigned)((ns + 127) / 128)), dim3(128), 0, (cudaStream_t)stream, coords, logp, s_idx, q,
                new_lp, factors, logu, (long long)ns, h->d, reinterpret_cast<long long*>(n_accepted), nan_flag, (const int*)nullptr,
                (double*)nullptr, (double*)nullptr));
    h->launches++;
    CK(cudaGetLastError());
    return 0;
}

int bi_stretch_accept_record(bi_handle h, double* coords, double* logp, const int32_t* s_idx, const double* q, const double* new_lp,
                             const double* factors, const double* logu, int64_t ns, int64_t* n_accepted, int32_t* nan_flag,
                             const int32_t* other_idx, double* chain_out, double* lp_out, void* stream) {
    if (!h || !coords || !logp || !s_idx || !q || !new_lp || !factors || !logu || !n_accepted || !nan_flag || !other_idx ||
        !chain_out || !lp_out || ns < 0)
        return fail(h, BI_E_ARG, "bad argument");
    if (ns == 0) return 0;
    CK(cudaSetDevice(h->device));
    CK(launch_k(h->pdl, stretch_accept_kernel, dim3((unsigned)((ns + 127) / 128)), dim3(128), 0, (cudaStream_t)stream, coords, logp, s_idx, q,
                new_lp, factors, logu, (long long)ns, h->d, reinterpret_cast<long long*>(n_accepted), nan_flag, other_idx, chain_out,
                lp_out));
    h->launches++;
    CK(cudaGetLastError());
    return 0;
}

int bi_stretch_randoms_philox(bi_handle h, uint64_t seed, uint64_t step0, int32_t nsteps, int32_t n_half, double a,
                              int32_t* s_idx, int32_t* c_idx, double* zz, double* factors, double* logu, void* stream) {
    if (!h || !s_idx || !c_idx || !zz || !factors || !logu || nsteps < 0 || n_half < 1 || !(a > 1.0))
        return fail(h, BI_E_ARG, "bad argument");
    if (nsteps == 0) return 0;
    CK(cudaSetDevice(h->device));
    const long long n = (long long)nsteps * 2 * n_half;
    stretch_randoms_philox_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        seed, step0, nsteps, n_half, h->d, a, s_idx, c_idx, zz, factors, logu);
    h->launches++;
    CK(cudaGetLastError());
    return 0;
}

// ---- several experimental data sets on one emulator image (closure tests) --------------------------------
int bi_set_data_sets(bi_handle h, int32_t n_sets, const double* y, const double* y_err) {
    if (!h || n_sets <= 0 || !y || !y_err) return fail(h, BI_E_ARG, "bad argument");
    if (!h->has_data) return fail(h, BI_E_STATE, "handle was created without experimental data (scratch is sized by it)");
    CK(cudaSetDevice(h->device));
    CK(cudaDeviceSynchronize());          // earlier calls may still read the previous sets on a caller's stream
    for (double** p : {&h->dEsets, &h->dYerrSets, &h->dATilesSets, &h->dBdASets})
        if (*p) { cudaFree(*p); *p = nullptr; }
    h->bd_sets_ndiv = -1.0;
    const int F = h->F, Fp = h->Fp;
    std::vector<double> e((size_t)n_sets * Fp, 0.0);
    for (int c = 0; c < n_sets; ++c)
        for (int r = 0; r < F; ++r) e[(size_t)c * Fp + r] = h->hMeanMerged[r] - y[(size_t)c * F + r];
    const size_t atile = (size_t)h->nt * (h->nt + 1) / 2 * TILE_ELEMS;
    CK(cudaMalloc(&h->dEsets, e.size() * 8));
    CK(cudaMalloc(&h->dYerrSets, (size_t)n_sets * F * 8));
    CK(cudaMalloc(&h->dATilesSets, (size_t)n_sets * atile * 8));
    CK(cudaMemcpy(h->dEsets, e.data(), e.size() * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(h->dYerrSets, y_err, (size_t)n_sets * F * 8, cudaMemcpyHostToDevice));
    h->n_sets = n_sets;
    h->sets_ndiv = -1.0;
    return 0;
}

int bi_log_posterior_sets(bi_handle h, const double* theta, const int32_t* set_id, int64_t W, double n_div, int mode,
                          double* logp_out, int32_t* info_out, void* stream) {
    if (!h || !theta || !set_id || !logp_out || W < 0) return fail(h, BI_E_ARG, "bad argument");
    if (h->n_sets <= 0) return fail(h, BI_E_STATE, "bi_set_data_sets has not been called");
    if (!(n_div > 0.0)) return fail(h, BI_E_ARG, "data sets need an explicit n_div (1 for MCMC)");
    if (mode != BI_MODE_DENSE && mode != BI_MODE_STRUCTURED)
        return fail(h, BI_E_UNSUPPORTED, "data sets support the dense and structured modes");
    if (W == 0) return 0;
    CK(cudaSetDevice(h->device));
    cudaStream_t s = (cudaStream_t)stream;
    if (mode == BI_MODE_STRUCTURED) {
        if (!h->sigma_blockdiag) return fail(h, BI_E_STRUCTURE, "Sigma_exp couples observable blocks: structured mode is not valid");
        int rc = ensure_block_direct(h);
        if (rc) return rc;
        if (h->bd_state == 1) return run_block_direct(h, theta, W, n_div, logp_out, info_out, s, set_id);
    }
    if (h->sets_ndiv != n_div) {
        const int ntl = h->nt * (h->nt + 1) / 2;
        for (int c = 0; c < h->n_sets; ++c) {
            build_a_tiles_kernel<<<ntl, 256, 0, s>>>(nullptr, h->dYerrSets + (size_t)c * h->F, h->dUblk, n_div, h->F,
                                                     h->chol_impl == 2 ? 1 : 0, h->dATilesSets + (size_t)c * ntl * TILE_ELEMS);
            h->launches++;
        }
        CK(cudaGetLastError());
        h->sets_ndiv = n_div;
    }
    for (long long w0 = 0; w0 < W; w0 += h->max_batch) {
        const long long wb = std::min<long long>(h->max_batch, W - w0);
        int rc = launch_gp(h, theta + w0 * h->d, wb, h->dM, h->dV, s);
        if (rc) return rc;
        rc = launch_chol(h, theta + w0 * h->d, h->dM, h->dV, wb, mode, logp_out + w0, info_out ? info_out + w0 : nullptr, s,
                         set_id + w0);
        if (rc) return rc;
    }
    return 0;
}

// ---- emulator training objective: batched log-marginal-likelihood + gradient ---------------------------
struct bi_trainer_s {
    int device = 0, N = 0, d = 0, n_y = 0, kind = 0, has_const = 0, has_white = 0, nh = 0, sms = 0, nwarps = 0, occ = 1;
    double alpha = 0.0;
    double *dX = nullptr, *dY = nullptr, *dScratch = nullptr, *dKt = nullptr, *dGt = nullptr, *dAb = nullptr;
    size_t smem_factor = 0, smem_grad = 0, slot = 0, tiles = 0;
    long long cap = 0, chunk_max = 0;      // problems the K / alpha buffers hold now; upper limit per launch group
    int nslots = 0, ntri = 0, Np = 0;
    long long launches = 0;
    // staging buffers of the _host calls (theta, y_index, lml, grad), grown on demand and kept: a fit makes ~1e3 calls
    double *hTh = nullptr, *hLml = nullptr, *hGrad = nullptr;
    int32_t* hIdx = nullptr;
    long long hcap = 0;
    std::string err;
};

static int tfail(bi_trainer t, int code, const std::string& msg) {
    if (t) t->err = msg;
    return code;
}
#define TCK(call)                                                                      \
    do {                                                                               \
        cudaError_t e_ = (call);                                                       \
        if (e_ != cudaSuccess) return tfail(t, (int)e_, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

}  // extern "C"
template <int KIND>
static cudaError_t trainer_attrs(bi_trainer t) {
    cudaError_t e = cudaSuccess;
    if (t->smem_grad > 48 * 1024) {
        e = allow_max_smem(gp_grad_kernel<KIND, 8>);
        if (e == cudaSuccess)
            e = allow_max_smem(gp_grad_kernel<KIND, 16>);
    }
    return e;
}
template <int KIND>
static void trainer_kernels(bi_trainer t, const LmlParams& P, bool with_grad, cudaStream_t s) {
    gp_kbuild_kernel<KIND><<<dim3((unsigned)t->ntri, (unsigned)P.B), 256, 0, s>>>(P);
    const unsigned grid = (unsigned)std::min<long long>(P.B, t->nslots);
    gp_factor_kernel<<<grid, t->nwarps * 32, t->smem_factor, s>>>(P);
    if (with_grad) {
        if (t->d <= 8) gp_grad_kernel<KIND, 8><<<(unsigned)P.B, 512, t->smem_grad, s>>>(P);
        else gp_grad_kernel<KIND, 16><<<(unsigned)P.B, 512, t->smem_grad, s>>>(P);
    }
    t->launches += with_grad ? 3 : 2;
}
extern "C" {

int bi_trainer_create(const double* X, int32_t N, int32_t d, const double* Y, int32_t n_y, int32_t kind, int32_t has_const,
                      int32_t has_white, double alpha, int32_t device, bi_trainer* out) {
    if (!out) return BI_E_ARG;
    *out = nullptr;
    if (!X || !Y || N < 1 || N + 1 > 32 * TRAIN_MAX_NT || d < 1 || d > TRAIN_MAX_D || n_y < 1 || kind < 0 || kind > 3 || !(alpha >= 0.0))
        return BI_E_ARG;
    bi_trainer t = new bi_trainer_s();
    t->device = device; t->N = N; t->d = d; t->n_y = n_y; t->kind = kind;
    t->has_const = has_const ? 1 : 0; t->has_white = has_white ? 1 : 0; t->alpha = alpha;
    t->nh = d + t->has_const + t->has_white;
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&t->sms, cudaDevAttrMultiProcessorCount, device);
    if (e == cudaSuccess) e = cudaMalloc(&t->dX, sizeof(double) * N * d);
    if (e == cudaSuccess) e = cudaMalloc(&t->dY, sizeof(double) * (size_t)n_y * N);
    if (e == cudaSuccess) e = cudaMemcpy(t->dX, X, sizeof(double) * N * d, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(t->dY, Y, sizeof(double) * (size_t)n_y * N, cudaMemcpyHostToDevice);
    const int nt = train_nt(N);
    t->ntri = nt * (nt + 1) / 2;
    t->Np = nt * 32;
    t->nwarps = std::max(2, std::min(nt, 16));          // one warp per unit of a panel (nt units)
    t->occ = std::max(1, 16 / t->nwarps);               // 512 threads per SM at 128 registers
    if (const char* s = getenv("BI_TRAIN_WARPS")) t->nwarps = std::max(2, std::min(16, atoi(s)));
    if (const char* s = getenv("BI_TRAIN_OCC")) t->occ = std::max(1, std::min(8, atoi(s)));
    t->smem_factor = lml_factor_smem_bytes(N, t->nwarps);
    t->smem_grad = lml_grad_smem_bytes(N, d);
    if (e == cudaSuccess && (t->smem_factor > 227 * 1024 || t->smem_grad > 227 * 1024)) e = cudaErrorInvalidValue;   // n_design too large
    t->tiles = lml_tiles_doubles(N);
    t->slot = t->tiles;                                  // M = L^-T has as many tiles as the lower triangle
    t->nslots = t->sms * t->occ;
    // the K / K^-1 tiles of a launch group live in HBM: at most ~4 GB of them at a time
    t->chunk_max = std::max<long long>(t->nslots, (long long)((3ull << 30) / (t->tiles * sizeof(double))));     // x2: K and G tiles
    if (e == cudaSuccess && t->smem_factor > 48 * 1024)
        e = allow_max_smem(gp_factor_kernel);
    if (e == cudaSuccess) e = kind == 0 ? trainer_attrs<0>(t) : kind == 1 ? trainer_attrs<1>(t) : kind == 2 ? trainer_attrs<2>(t) : trainer_attrs<3>(t);
    if (e == cudaSuccess) e = cudaMalloc(&t->dScratch, sizeof(double) * t->slot * t->nslots);
    if (e != cudaSuccess) {
        const int rc = (int)e;
        cudaFree(t->dX); cudaFree(t->dY); cudaFree(t->dScratch);
        delete t;
        return rc;
    }
    *out = t;
    return 0;
}

void bi_trainer_destroy(bi_trainer t) {
    if (!t) return;
    cudaSetDevice(t->device);
    cudaFree(t->dX); cudaFree(t->dY); cudaFree(t->dScratch); cudaFree(t->dKt); cudaFree(t->dGt); cudaFree(t->dAb);
    cudaFree(t->hTh); cudaFree(t->hLml); cudaFree(t->hGrad); cudaFree(t->hIdx);
    delete t;
}

const char* bi_trainer_last_error(bi_trainer t) { return t ? t->err.c_str() : "null trainer"; }
int64_t bi_trainer_launches(bi_trainer t) { return t ? t->launches : 0; }

static int trainer_launch(bi_trainer t, const double* theta, const int32_t* y_index, int64_t B, double* lml, double* grad,
                          double* L_out, double* alpha_out, cudaStream_t s) {
    const long long need = std::min<long long>(B, t->chunk_max);
    if (need > t->cap) {                               // grow the K / alpha buffers (stream-ordered free of the old ones)
        TCK(cudaStreamSynchronize(s));
        cudaFree(t->dKt); cudaFree(t->dGt); cudaFree(t->dAb);
        t->dKt = t->dGt = t->dAb = nullptr;
        t->cap = 0;
        TCK(cudaMalloc(&t->dKt, sizeof(double) * t->tiles * need));
        TCK(cudaMalloc(&t->dGt, sizeof(double) * t->tiles * need));
        TCK(cudaMalloc(&t->dAb, sizeof(double) * (size_t)t->Np * need));
        t->cap = need;
    }
    for (int64_t b0 = 0; b0 < B; b0 += t->cap) {
        LmlParams P;
        P.B = std::min<long long>(t->cap, B - b0);
        P.X = t->dX; P.Y = t->dY; P.y_index = y_index ? y_index + b0 : nullptr; P.theta = theta + b0 * t->nh;
        if (!y_index) P.Y = t->dY + b0 * t->N;         // problem b uses row b
        P.N = t->N; P.d = t->d; P.nh = t->nh; P.has_const = t->has_const; P.has_white = t->has_white;
        P.alpha = t->alpha; P.Kt = t->dKt; P.Gt = grad ? t->dGt : nullptr; P.Ab = t->dAb; P.scratch = t->dScratch; P.scratch_stride = (long long)t->slot;
        P.lml = lml + b0; P.grad = grad ? grad + b0 * t->nh : nullptr;
        P.L_out = L_out ? L_out + (size_t)b0 * t->N * t->N : nullptr;
        P.alpha_out = alpha_out ? alpha_out + (size_t)b0 * t->N : nullptr;
#ifdef TRAIN_TIMING
        { unsigned long long z[12] = {0}; cudaMemcpyToSymbolAsync(train_dbg, z, sizeof(z), 0, cudaMemcpyHostToDevice, s); }
#endif
        const bool wg = grad != nullptr;
        switch (t->kind) {
            case 0: trainer_kernels<0>(t, P, wg, s); break;
            case 1: trainer_kernels<1>(t, P, wg, s); break;
            case 2: trainer_kernels<2>(t, P, wg, s); break;
            default: trainer_kernels<3>(t, P, wg, s); break;
        }
#ifdef TRAIN_TIMING
        {
            unsigned long long z[12];
            cudaMemcpyFromSymbolAsync(z, train_dbg, sizeof(z), 0, cudaMemcpyDeviceToHost, s);
            cudaStreamSynchronize(s);
            const double nb = (double)P.B;
            std::fprintf(stderr, "[train timing] cycles per problem (thread 0): load %.0f  diag update %.0f  potrf+inv %.0f  wait updates %.0f  "
                         "solves %.0f  alpha %.0f  Kinv %.0f\n", z[0] / nb, z[1] / nb, z[2] / nb, z[3] / nb,
                         z[4] / nb, z[5] / nb, z[7] / nb);
        }
#endif
        TCK(cudaGetLastError());
    }
    return 0;
}

int bi_trainer_lml(bi_trainer t, const double* theta, const int32_t* y_index, int64_t B, double* lml_out, double* grad_out,
                   void* stream) {
    if (!t || !theta || !lml_out || !grad_out || B < 0) return tfail(t, BI_E_ARG, "bad argument");
    if (!y_index && B > t->n_y) return tfail(t, BI_E_ARG, "y_index required when B exceeds the number of target rows");
    if (B == 0) return 0;
    TCK(cudaSetDevice(t->device));
    return trainer_launch(t, theta, y_index, B, lml_out, grad_out, nullptr, nullptr, (cudaStream_t)stream);
}

int bi_trainer_factor(bi_trainer t, const double* theta, const int32_t* y_index, int64_t B, double* lml_out, double* L_out,
                      double* alpha_out, void* stream) {
    if (!t || !theta || !lml_out || !L_out || !alpha_out || B < 0) return tfail(t, BI_E_ARG, "bad argument");
    if (!y_index && B > t->n_y) return tfail(t, BI_E_ARG, "y_index required when B exceeds the number of target rows");
    if (B == 0) return 0;
    TCK(cudaSetDevice(t->device));
    return trainer_launch(t, theta, y_index, B, lml_out, nullptr, L_out, alpha_out, (cudaStream_t)stream);
}

static int trainer_host(bi_trainer t, const double* theta, const int32_t* y_index, int64_t B, double* lml_out, double* grad_out,
                        double* L_out, double* alpha_out) {
    if (!t || !theta || !lml_out || B < 0) return tfail(t, BI_E_ARG, "bad argument");
    if (!y_index && B > t->n_y) return tfail(t, BI_E_ARG, "y_index required when B exceeds the number of target rows");
    if (y_index)
        for (int64_t b = 0; b < B; ++b)
            if (y_index[b] < 0 || y_index[b] >= t->n_y) return tfail(t, BI_E_ARG, "y_index out of range");
    if (B == 0) return 0;
    TCK(cudaSetDevice(t->device));
    const size_t NN = (size_t)t->N * t->N;
    if (B > t->hcap) {
        cudaFree(t->hTh); cudaFree(t->hLml); cudaFree(t->hGrad); cudaFree(t->hIdx);
        t->hTh = t->hLml = t->hGrad = nullptr;
        t->hIdx = nullptr;
        t->hcap = 0;
        const long long cap = std::max<long long>(B, 256);
        TCK(cudaMalloc(&t->hTh, sizeof(double) * cap * t->nh));
        TCK(cudaMalloc(&t->hLml, sizeof(double) * cap));
        TCK(cudaMalloc(&t->hGrad, sizeof(double) * cap * t->nh));
        TCK(cudaMalloc(&t->hIdx, sizeof(int32_t) * cap));
        t->hcap = cap;
    }
    double *dTh = t->hTh, *dLml = t->hLml, *dGrad = grad_out ? t->hGrad : nullptr, *dL = nullptr, *dAl = nullptr;
    int32_t* dIdx = y_index ? t->hIdx : nullptr;
    int rc = 0;
    cudaError_t e = cudaSuccess;
    if (L_out) e = cudaMalloc(&dL, sizeof(double) * B * NN);
    if (e == cudaSuccess && L_out) e = cudaMalloc(&dAl, sizeof(double) * B * t->N);
    if (e == cudaSuccess) e = cudaMemcpy(dTh, theta, sizeof(double) * B * t->nh, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && y_index) e = cudaMemcpy(dIdx, y_index, sizeof(int32_t) * B, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) rc = trainer_launch(t, dTh, dIdx, B, dLml, dGrad, dL, dAl, nullptr);
    if (e == cudaSuccess && rc == 0) e = cudaMemcpy(lml_out, dLml, sizeof(double) * B, cudaMemcpyDeviceToHost);     // default stream: ordered after the kernels
    if (e == cudaSuccess && rc == 0 && grad_out) e = cudaMemcpy(grad_out, dGrad, sizeof(double) * B * t->nh, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && rc == 0 && L_out) e = cudaMemcpy(L_out, dL, sizeof(double) * B * NN, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && rc == 0 && L_out) e = cudaMemcpy(alpha_out, dAl, sizeof(double) * B * t->N, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) rc = tfail(t, (int)e, std::string("bi_trainer host call: ") + cudaGetErrorString(e));
    cudaFree(dL); cudaFree(dAl);
    return rc;
}

int bi_trainer_lml_host(bi_trainer t, const double* theta, const int32_t* y_index, int64_t B, double* lml_out, double* grad_out) {
    if (!grad_out) return tfail(t, BI_E_ARG, "bad argument");
    return trainer_host(t, theta, y_index, B, lml_out, grad_out, nullptr, nullptr);
}

int bi_trainer_factor_host(bi_trainer t, const double* theta, const int32_t* y_index, int64_t B, double* lml_out, double* L_out,
                           double* alpha_out) {
    if (!L_out || !alpha_out) return tfail(t, BI_E_ARG, "bad argument");
    return trainer_host(t, theta, y_index, B, lml_out, nullptr, L_out, alpha_out);
}

}  // extern "C"

// ---- bi_loglikelihood_noalloc: the explicit (dy, C) log-likelihood for callers that may not allocate ----------------------
// No counterpart in the reference; added at the request of the run's operators (callers under CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc).  Same kernel as bi_loglikelihood_host (chol_loglik_kernel<true>), but
// its per-CTA scratch (the factor's strictly-lower tiles + one z tile per panel) and the kmin / imax table live in ONE
// file-scope device array of fixed size instead of cudaMalloc'd buffers.  Every GPU that loads the library holds
// BI_NOALLOC_SCRATCH_BYTES (8 MiB + 256 B) until the process exits.  All calls share the array: the kernel writes every
// scratch tile before it reads it and the table is rewritten by every call, so nothing has to start at zero, but two
// calls that overlap on the GPU (different streams, no ordering) race on it.
// BI_NOALLOC_MAX_F (256: 8 x 8 tiles of 32 x 32) comes from include/bi_b200.h
#define BI_NOALLOC_WS_DOUBLES (1 << 20)           /* 8 MiB of factor scratch */
#define BI_NOALLOC_TABLE_DOUBLES 32               /* 256 B: kmin[8], imax[8] as int32 */
#define BI_NOALLOC_SCRATCH_BYTES (8 * (BI_NOALLOC_WS_DOUBLES + BI_NOALLOC_TABLE_DOUBLES))
static __device__ __align__(256) double g_noalloc_scratch[BI_NOALLOC_TABLE_DOUBLES + BI_NOALLOC_WS_DOUBLES];

namespace {
// device code may name the array directly: dense table (kmin = 0, imax = nt - 1) for this call's tile count
__global__ void noalloc_table_kernel(int nt) {
    int* tab = reinterpret_cast<int*>(g_noalloc_scratch);
    const int i = threadIdx.x;
    if (i < nt) {
        tab[i] = 0;
        tab[nt + i] = nt - 1;
    }
}

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
double* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};   // DEVICE address of g_noalloc_scratch, per device (the symbol's name is
                                                   // only a host-side stand-in in host code)

// once per device, with that device current
cudaError_t noalloc_scratch_address(int device, double** out) {
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        double* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_noalloc_ptr[device] = p;
    }
    *out = g_noalloc_ptr[device];
    return cudaSuccess;
}
}  // namespace

extern "C" int bi_loglikelihood_noalloc(bi_handle h, const double* dy, const double* cov, int64_t n, int32_t Fdim, double* out,
                                        int32_t* info_out, void* stream) {
    if (!h || !dy || !cov || !out || n < 0 || Fdim <= 0) return fail(h, BI_E_ARG, "bad argument");
    if (Fdim > BI_NOALLOC_MAX_F)
        return fail(h, BI_E_UNSUPPORTED, "bi_loglikelihood_noalloc: F exceeds BI_NOALLOC_MAX_F = 256 (static scratch); use bi_loglikelihood_host");
    if (n > (int64_t)1 << 40) return fail(h, BI_E_UNSUPPORTED, "bi_loglikelihood_noalloc: n exceeds 2^40 rows");
    if (n == 0) return 0;
    CK(cudaSetDevice(h->device));
    const int F = Fdim, nt = (F + 31) / 32, Fp = nt * 32;
    const size_t smem = chol_smem_bytes(Fp, 4);
    CK(raise_explicit_smem(smem));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, chol_loglik_kernel<true>, CHOL_THREADS, smem));
    if (occ < 1) return fail(h, BI_E_UNSUPPORTED, "matrix too large for the Cholesky kernel's shared memory");
    double* scratch = nullptr;
    CK(noalloc_scratch_address(h->device, &scratch));
    const long long stride = chol_ws_doubles(nt);              // <= 36,864 doubles at nt = 8
    const long long fit = BI_NOALLOC_WS_DOUBLES / stride;      // CTAs the fixed scratch can serve (>= 28)
    const int ctas = (int)std::min<long long>(std::min<long long>(n, fit), (long long)h->sm_count * occ);
    cudaStream_t s = (cudaStream_t)stream;
    noalloc_table_kernel<<<1, 32, 0, s>>>(nt);
    h->launches++;
    CK(cudaGetLastError());
    int* tab = reinterpret_cast<int*>(scratch);
    CholParams P;
    std::memset(&P, 0, sizeof(P));
    P.F = F; P.Fp = Fp; P.nt = nt; P.K0 = 4; P.nks0 = 1; P.Ptot = 0; P.d = h->d; P.W = n;
    P.kmin = tab; P.imax = tab + nt; P.ws = scratch + BI_NOALLOC_TABLE_DOUBLES; P.ws_stride = stride;
    P.logp = out; P.info = info_out; P.cov_x = cov; P.dy_x = dy;
    chol_loglik_kernel<true><<<ctas, CHOL_THREADS, smem, s>>>(P);
    h->launches++;
    CK(cudaGetLastError());
    return 0;
}
