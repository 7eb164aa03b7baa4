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

namespace bi {

// one thread per (proposal k, parameter) element; s_idx[k] = walker being moved, c_idx[k] = its partner
__global__ void stretch_propose_kernel(const double* __restrict__ coords, const int* __restrict__ s_idx,
                                       const int* __restrict__ c_idx, const double* __restrict__ zz, long long ns, int d,
                                       double* __restrict__ q) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= ns * d) return;
    const long long k = idx / d;
    const int p = (int)(idx - k * d);
    const double c = coords[(long long)c_idx[k] * d + p];
    const double s = coords[(long long)s_idx[k] * d + p];
    // numpy: c[rint] - (c[rint] - s) * zz[:, None]   -- three separately rounded operations
    q[idx] = __dsub_rn(c, __dmul_rn(__dsub_rn(c, s), zz[k]));
}

// one thread per proposal; walkers of one half-step are distinct, so the scatter needs no atomics
__global__ void stretch_accept_kernel(double* __restrict__ coords, double* __restrict__ logp, const int* __restrict__ s_idx,
                                      const double* __restrict__ q, const double* __restrict__ new_lp,
                                      const double* __restrict__ factors, const double* __restrict__ logu, long long ns,
                                      int d, long long* __restrict__ n_accepted, int* __restrict__ nan_flag) {
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
}

}  // namespace bi
