// ensemble.cuh -- the two state-dependent pieces of an affine-invariant stretch-move half-step, on the device.
//
// The reference drives log_posterior through emcee.EnsembleSampler (mcmc.py:83-107; emcee 3.1.4 RedBlueMove.propose +
// StretchMove.get_proposal, restated host-side in bayesian_inference_b200/ensemble.py).  Everything a half-step draws
// from the random stream -- the red/blue split, the stretch factors zz, the partner indices, the acceptance uniforms --
// is independent of the chain state, so the host generates it ahead of time in emcee's consumption order and uploads
// it; what remains per half-step is
//     q_k      = c_k - (c_k - s_k) * zz_k                              (stretch_propose_kernel)
//     accept_k = (factors_k + logp_new_k) - logp_old_k > log u_k        (stretch_accept_kernel)
// with the log-posterior kernels in between, all on one stream without a host round trip.  The arithmetic is written
// with explicitly rounded operations (no FMA contraction) in numpy's evaluation order, so the accepted chain is
// bit-identical to the host-side sampler's.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "common.cuh"

namespace bi {

// one thread per (proposal k, parameter) element; s_idx[k] = walker being moved, c_idx[k] = its partner
__global__ void stretch_propose_kernel(const double* __restrict__ coords, const int* __restrict__ s_idx,
                                       const int* __restrict__ c_idx, const double* __restrict__ zz, long long ns, int d,
                                       double* __restrict__ q) {
    pdl_wait();
    pdl_trigger();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= ns * d) return;
    const long long k = idx / d;
    const int p = (int)(idx - k * d);
    const double c = coords[(long long)c_idx[k] * d + p];
    const double s = coords[(long long)s_idx[k] * d + p];
    // numpy: c[rint] - (c[rint] - s) * zz[:, None]   -- three separately rounded operations
    q[idx] = __dsub_rn(c, __dmul_rn(__dsub_rn(c, s), zz[k]));
}

// one thread per proposal; walkers of one half-step are distinct, so the scatter needs no atomics.
// chain_out != nullptr (the second half-step of a step): the state of the ensemble after this step is recorded as well --
// thread k writes the row of its own walker after its accept decision and the row of other_idx[k], a walker of the half
// that does not move in this half-step (other_idx = the s_idx of the first half-step: together they are all W walkers).
__global__ void stretch_accept_kernel(double* __restrict__ coords, double* __restrict__ logp, const int* __restrict__ s_idx,
                                      const double* __restrict__ q, const double* __restrict__ new_lp,
                                      const double* __restrict__ factors, const double* __restrict__ logu, long long ns,
                                      int d, long long* __restrict__ n_accepted, int* __restrict__ nan_flag,
                                      const int* __restrict__ other_idx, double* __restrict__ chain_out,
                                      double* __restrict__ lp_out) {
    pdl_wait();
    pdl_trigger();
    const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= ns) return;
    const double nl = new_lp[k];
    if (nl != nl) {           // emcee: ValueError("Probability function returned NaN"); the host raises it
        atomicExch(nan_flag, 1);
        return;
    }
    const int w = s_idx[k];
    // numpy: lnpdiff = factors + new_log_probs - state.log_prob[all_inds[S1]]   (left to right)
    const double lnpdiff = __dsub_rn(__dadd_rn(factors[k], nl), logp[w]);
    if (lnpdiff > logu[k]) {
        for (int p = 0; p < d; ++p) coords[(long long)w * d + p] = q[k * d + p];
        logp[w] = nl;
        n_accepted[w] += 1;
    }
    if (chain_out != nullptr) {
        const int o = other_idx[k];
        for (int p = 0; p < d; ++p) {
            chain_out[(long long)w * d + p] = coords[(long long)w * d + p];
            chain_out[(long long)o * d + p] = coords[(long long)o * d + p];
        }
        lp_out[w] = logp[w];
        lp_out[o] = logp[o];
    }
}

// ---- counter-based random numbers for the device-resident sampler (no host stream, SURVEY 8 f1 second half) -----------
// Philox4x32-10 (Salmon et al., SC'11) keyed by the seed; the counter is (proposal k, half | draw << 1, step lo, step hi),
// so any step of any chain can be regenerated and the numbers do not depend on how steps are grouped into blocks.
// This is NOT emcee's RandomState stream: the ensemble is split into even / odd walkers instead of a fresh permutation
// per step (Goodman & Weare's original fixed halves), and the chain differs from the reference's for the same seed --
// it is the variant "reported separately"; tests pin the generator against a numpy restatement bit for bit.
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u;
        k.y += 0xBB67AE85u;
    }
    return c;
}
// 52 random bits -> a double strictly inside (0, 1); every operation is exact
__device__ __forceinline__ double u52(uint32_t hi, uint32_t lo) {
    return ((double)(hi >> 6) * 67108864.0 + (double)(lo >> 6) + 0.5) * (1.0 / 4503599627370496.0);
}

// one thread per (step t, half h, proposal k); outputs shaped (nsteps, 2, ns) like ensemble.draw_stretch_randoms
__global__ void stretch_randoms_philox_kernel(unsigned long long seed, unsigned long long step0, int nsteps, int ns, int ndim,
                                              double a, int* __restrict__ s_idx, int* __restrict__ c_idx,
                                              double* __restrict__ zz, double* __restrict__ factors, double* __restrict__ logu) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)nsteps * 2 * ns) return;
    const int k = (int)(idx % ns), h = (int)((idx / ns) & 1);
    const unsigned long long step = step0 + (unsigned long long)(idx / (2LL * ns));
    const uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
    const uint4 r0 = philox4x32_10(make_uint4((uint32_t)k, (uint32_t)h, (uint32_t)step, (uint32_t)(step >> 32)), key);
    const uint4 r1 = philox4x32_10(make_uint4((uint32_t)k, (uint32_t)(h | 2), (uint32_t)step, (uint32_t)(step >> 32)), key);
    const double u1 = u52(r0.x, r0.y), u2 = u52(r0.z, r0.w), u3 = u52(r1.x, r1.y);
    // emcee StretchMove: zz = ((a - 1) u + 1)^2 / a, factors = (ndim - 1) log zz; accept test against log u
    const double t = __dadd_rn(__dmul_rn(a - 1.0, u1), 1.0);
    const double z = __ddiv_rn(__dmul_rn(t, t), a);
    const int j = min((int)(u2 * (double)ns), ns - 1);
    s_idx[idx] = 2 * k + h;                 // the half being moved: even walkers first, then odd
    c_idx[idx] = 2 * j + (1 - h);           // partner from the other half
    zz[idx] = z;
    factors[idx] = (double)(ndim - 1) * log(z);
    logu[idx] = log(u3);
}

}  // namespace bi
