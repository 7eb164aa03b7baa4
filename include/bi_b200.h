/* bi_b200.h -- C ABI of the B200-native MCMC log-posterior path.
 *
 * One shared library (libbi_b200.so, built from bayesian_inference_b200/csrc/ by nvcc for sm_100a) exports the
 * functions below.  They are what a foreign-function binding of the reference's hot path would bind; each one cites
 * the reference interface it replaces (paths relative to the reference checkout, src/bayesian_inference/).
 *
 *   reference                                                     this ABI
 *   ------------------------------------------------------------  -------------------------------------------
 *   log_posterior.initialize_pool_variables  log_posterior.py:26   bi_create            (load once into HBM)
 *   emulator.predict(X, return_std=True) xP  emulation.py:496-499  bi_predict_pc[_host]
 *   sklearn kernel_(X, X_train_)             _gpr.py:446           bi_kernel_vectors[_host]
 *   emulation.predict_emulation_group        emulation.py:466-548  bi_predict_group[_host]
 *   emulation.predict (merged)               emulation.py:410-462  bi_predict[_host]
 *   log_posterior.log_posterior              log_posterior.py:42   bi_log_posterior[_host]
 *   log_posterior._loglikelihood             log_posterior.py:104  bi_loglikelihood_host
 *   (none: operators' request, allocation-free variant)            bi_loglikelihood_noalloc
 *   emcee stretch move driven by mcmc.run_mcmc  mcmc.py:83-107     bi_stretch_propose / bi_stretch_accept
 *   GaussianProcessRegressor.fit objective   emulation.py:169-172  bi_trainer_lml / bi_trainer_factor
 *
 * Conventions
 *   - every matrix is row-major float64; every index is 0-based;
 *   - bi_desc / bi_group_desc carry HOST pointers; bi_create copies them to the device and builds the device image
 *     (L^-1 per principal component, tiled Sigma_exp / unexplained covariance / inverse-PCA matrices);
 *   - functions without the _host suffix take DEVICE pointers on the handle's device, enqueue on `stream`
 *     (a cudaStream_t passed as void*; NULL is CUDA's default stream, as everywhere in the runtime API) and return
 *     without synchronising; their scratch buffers are per handle, so calls on one handle must be stream-ordered
 *     (one stream at a time).  The theta-independent tiles Sigma_exp + U / n_div are kept per divisor: the buffer of
 *     n_div = 1 is built once and never rewritten, so a CUDA graph captured around bi_log_posterior(n_div = 1) stays
 *     valid when the same handle is later asked for other divisors;
 *   - functions with the _host suffix take HOST pointers, copy in, run, copy out and synchronise before returning;
 *   - return value: 0 = ok, < 0 = BI_E_* (bad argument / state), > 0 = a cudaError_t value.
 *     bi_last_error(h) returns a static description of the last failure on that handle;
 *   - a handle is not re-entrant: one host thread per handle (the reference: one process per pool worker).
 *
 * There is no CPU fallback behind this ABI.
 */
#ifndef BI_B200_H
#define BI_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BI_ABI_VERSION 1

/* stationary kernel of one principal component's GP (sklearn kernels.py:1685-1712 Matern, :1530-1545 RBF) */
#define BI_KERNEL_MATERN12 0 /* exp(-r)                               */
#define BI_KERNEL_MATERN32 1 /* (1 + sqrt3 r) exp(-sqrt3 r)           */
#define BI_KERNEL_MATERN52 2 /* (1 + sqrt5 r + 5 r^2 / 3) exp(-sqrt5 r) */
#define BI_KERNEL_RBF      3 /* exp(-r^2 / 2)                         */

/* how the per-walker covariance is factorised in bi_log_posterior */
#define BI_MODE_DENSE      0 /* full F x F blocked Cholesky, no structure assumed (what the reference's dpotrf does)  */
#define BI_MODE_STRUCTURED 1 /* skips tiles that are structurally zero (block-diagonal-per-observable covariance);   */
                             /* falls back to nothing: refused with BI_E_STRUCTURE if Sigma_exp couples observables  */
#define BI_MODE_WOODBURY   2 /* matrix-determinant-lemma / Woodbury: n_pc x n_pc factorisation per observable block   */

#define BI_E_ARG        (-1) /* NULL / out-of-range argument            */
#define BI_E_STATE      (-2) /* handle lacks what the call needs (e.g. no experimental data for log_posterior) */
#define BI_E_STRUCTURE  (-3) /* structured / Woodbury mode requested but Sigma_exp couples observable blocks   */
#define BI_E_UNSUPPORTED (-4)
#define BI_E_NOMEM      (-5)

/* One emulation group: the fitted objects of emulation.py:181-190, flattened. */
typedef struct bi_group_desc {
    int32_t n_design;        /* N   rows of X_train_                                           */
    int32_t n_pc;            /* P   retained principal components = number of GPs              */
    int32_t n_features;      /* F_g columns of pca.components_                                 */
    int32_t n_rest;          /* Q-P truncated principal components (rows of S_rest)            */
    const double*  X;        /* (N, d)    gpr.X_train_ (shared by all PCs, copy_X_train=False) */
    const int32_t* kind;     /* (P)       BI_KERNEL_*                                          */
    const double*  ell;      /* (P, d)    anisotropic length scales                            */
    const double*  constant; /* (P)       ConstantKernel value, 0 if absent                    */
    const double*  noise;    /* (P)       WhiteKernel level, 0 if absent                       */
    const double*  alpha;    /* (P, N)    gpr.alpha_                                           */
    const double*  L;        /* (P, N, N) gpr.L_, lower triangle used                          */
    const double*  S;        /* (P, F_g)  pca.components_[:P]                                  */
    const double*  S_rest;   /* (Q-P, F_g) pca.components_[P:]   (may be NULL when n_rest==0)  */
    const double*  ev_rest;  /* (Q-P)     pca.explained_variance_[P:]                          */
    const double*  scale;    /* (F_g)     scaler.scale_                                        */
    const double*  mean;     /* (F_g)     scaler.mean_                                         */
} bi_group_desc;

/* Everything initialize_pool_variables (log_posterior.py:26-38) installs, flattened. */
typedef struct bi_desc {
    int32_t abi_version;     /* BI_ABI_VERSION                                                              */
    int32_t device;          /* CUDA device ordinal                                                         */
    int32_t d;               /* parameter dimension                                                         */
    int32_t n_groups;
    int32_t n_blocks;        /* B observable blocks                                                         */
    int32_t n_features;      /* F width of the merged matrices                                              */
    const bi_group_desc* groups;
    const int64_t* blocks;   /* (B, 4) rows [group, out_start, in_start, len] (emulation.py:362-404)        */
    const double*  prior_min;/* (d) strict bounds (log_posterior.py:63)                                     */
    const double*  prior_max;/* (d)                                                                         */
    const double*  y;        /* (F) experimental central values, NULL for predict-only handles              */
    const double*  y_err;    /* (F) Sigma_exp = diag(y_err^2) (log_posterior.py:94) unless sigma_exp given  */
    const double*  sigma_exp;/* (F, F) dense experimental covariance or NULL                                */
    int32_t max_batch;       /* walkers processed per internal pass (0 = default 4096); bounds scratch HBM  */
    int32_t reserved;
} bi_desc;

typedef struct bi_handle_s* bi_handle;

int  bi_abi_version(void);
int  bi_create(const bi_desc* desc, bi_handle* out);
void bi_destroy(bi_handle h);
const char* bi_last_error(bi_handle h);
/* query: what = 0 F, 1 total n_pc, 2 d, 3 n_groups, 4 SM count, 5 Cholesky CTAs resident, 6 scratch bytes,
 *        7 kernels launched since create (launch counter used by bench.py's gpu_launches),
 *        16 / 17 FP64 tensor instructions (DMMA m8n8k4 = 512 flop each) the fused Cholesky kernel issues per walker in
 *        BI_MODE_DENSE / BI_MODE_STRUCTURED, counted from the tile tables the kernel walks (bench.py's flops_executed),
 *        18 whether BI_MODE_STRUCTURED runs block by block in registers (1: every observable block has <= 31 rows) or on
 *        the fused Cholesky kernel with tile skipping (0),
 *        32 guard mode (environment BI_DEBUG_GUARD=1 at bi_create: canary bands around every device buffer of the handle,
 *        payloads poisoned with NaN): number of canary bands overwritten so far, -1 when guard mode is off             */
int64_t bi_query(bi_handle h, int what);
/* synchronise the handle's own stream */
int  bi_sync(bi_handle h);
/* per-kernel device timing with CUDA events recorded on the launching stream (used by bench.py's roofline):
 * bi_profile(h, 1) starts recording; bi_profile_read synchronises the device, returns and resets the summed duration
 * and launch count of kernel 0 = gp_predict (K1+K2) or 1 = chol_loglik (K4). */
int  bi_profile(bi_handle h, int enable);
int  bi_profile_read(bi_handle h, int kernel, double* total_ms, int64_t* n_launches);

/* k[w][p][j] = kernel_p(theta_w, X_j) of group g: out (W, P_g, N_g).  sklearn _gpr.py:446. */
int bi_kernel_vectors(bi_handle h, int group, const double* theta, int64_t W, double* k_out, void* stream);
int bi_kernel_vectors_host(bi_handle h, int group, const double* theta, int64_t W, double* k_out);

/* GP mean and variance of every principal component of every group, concatenated in group order:
 * m_out, v_out (W, P_total).  emulation.py:496-499 (v = std**2 after sklearn's clip at 0). */
int bi_predict_pc(bi_handle h, const double* theta, int64_t W, double* m_out, double* v_out, void* stream);
int bi_predict_pc_host(bi_handle h, const double* theta, int64_t W, double* m_out, double* v_out);

/* One group in its own feature order: mean_out (W, F_g), cov_out (W, F_g, F_g) or NULL.
 * n_div divides the unexplained covariance (emulation.py:493,531-532: rows in the call); pass 1 for MCMC. */
int bi_predict_group(bi_handle h, int group, const double* theta, int64_t W, double n_div,
                     double* mean_out, double* cov_out, void* stream);
int bi_predict_group_host(bi_handle h, int group, const double* theta, int64_t W, double n_div,
                          double* mean_out, double* cov_out);

/* Merged prediction in sorted-observable order: mean_out (W, F); cov_out (W, F, F) block-diagonal per observable,
 * or NULL.  emulation.py:410-462 + :346-406. */
int bi_predict(bi_handle h, const double* theta, int64_t W, double n_div, double* mean_out, double* cov_out,
               void* stream);
int bi_predict_host(bi_handle h, const double* theta, int64_t W, double n_div, double* mean_out, double* cov_out);

/* log-posterior of W walkers: logp_out (W) (-inf outside the strict prior box, NaN where the covariance is not
 * positive definite), info_out (W) int32 or NULL (0 ok, k>0 = first non-positive pivot, 1-based, -1 outside prior).
 * n_div <= 0 means "rows inside the prior box in this call" (what a batched reference call does,
 * log_posterior.py:67,80 + emulation.py:493); MCMC passes 1. */
int bi_log_posterior(bi_handle h, const double* theta, int64_t W, double n_div, int mode,
                     double* logp_out, int32_t* info_out, void* stream);
int bi_log_posterior_host(bi_handle h, const double* theta, int64_t W, double n_div, int mode,
                          double* logp_out, int32_t* info_out);

/* -1/2 dy^T C^-1 dy - sum log diag chol(C) for n explicit (dy, C) pairs: dy (n, F'), cov (n, F', F').
 * log_posterior.py:104-146.  F' is arbitrary (need not equal the handle's F). */
int bi_loglikelihood_host(bi_handle h, const double* dy, const double* cov, int64_t n, int32_t Fdim,
                          double* out, int32_t* info_out);

/* The same log-likelihood for callers that may not allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference: added at the request of the run's operators.
 * dy (n, F'), cov (n, F', F'), out (n) and info_out (n, or NULL) are DEVICE pointers owned by the caller on the handle's
 * device; the call enqueues two kernels on `stream` and returns without synchronising.  It calls no cudaMalloc /
 * cudaFree: the kernel's scratch is a file-scope device array of BI_NOALLOC_SCRATCH_BYTES (8 MiB + 256 B) that every
 * GPU loading the library holds until exit.  Limits: F' <= BI_NOALLOC_MAX_F (256), else BI_E_UNSUPPORTED; n is only
 * bounded by the caller's buffers (the fixed scratch serves at most 8 MiB / scratch-per-CTA resident CTAs, which loop
 * over the rows).  All calls, from every handle on a device, share the one array: calls that overlap on the GPU
 * (different streams without an ordering between them) race; calls ordered on one stream are safe, and nothing in the
 * array has to be cleared between calls (every tile is written before it is read). */
#define BI_NOALLOC_MAX_F 256
int bi_loglikelihood_noalloc(bi_handle h, const double* dy, const double* cov, int64_t n, int32_t Fdim,
                             double* out, int32_t* info_out, void* stream);

/* ---- the sampler step around the log-posterior (mcmc.py:83-107 -> emcee 3.1.4 RedBlueMove.propose / StretchMove) ----
 * One half-step of the stretch move on device-resident state; everything drawn from the random stream (split, zz,
 * partners, acceptance uniforms) is chain-independent and is generated by the host in emcee's order.
 * All pointers are DEVICE pointers; both calls enqueue on `stream` and return without synchronising.
 *
 * bi_stretch_propose: q_out[k][:] = c - (c - s) * zz[k] with s = coords[s_idx[k]], c = coords[c_idx[k]]; q_out (ns, d).
 * bi_stretch_accept : for each k, if (factors[k] + new_lp[k]) - logp[s_idx[k]] > logu[k]: coords[s_idx[k]] = q[k],
 *                     logp[s_idx[k]] = new_lp[k], n_accepted[s_idx[k]] += 1.  A NaN in new_lp sets *nan_flag = 1 and
 *                     accepts nothing for that row (emcee raises ValueError; the host mirror does so when it reads
 *                     the flag).  s_idx must not contain duplicates (one half of the ensemble). */
int bi_stretch_propose(bi_handle h, const double* coords, const int32_t* s_idx, const int32_t* c_idx, const double* zz,
                       int64_t ns, double* q_out, void* stream);
int bi_stretch_accept(bi_handle h, double* coords, double* logp, const int32_t* s_idx, const double* q,
                      const double* new_lp, const double* factors, const double* logu, int64_t ns,
                      int64_t* n_accepted, int32_t* nan_flag, void* stream);
/* bi_stretch_accept_record: the accept of the SECOND half-step of a step, which also stores the ensemble after the step the
 * way emcee's backend keeps every iteration (the chain and log_prob mcmc.py:111-116 reads back): chain_out (W, d) and
 * lp_out (W) receive the rows of the ns walkers moved here (after their accept decision) and of the ns walkers named by
 * other_idx (the s_idx of the first half-step, unchanged in this one); W = 2 ns.  Otherwise identical to bi_stretch_accept. */
int bi_stretch_accept_record(bi_handle h, double* coords, double* logp, const int32_t* s_idx, const double* q,
                             const double* new_lp, const double* factors, const double* logu, int64_t ns,
                             int64_t* n_accepted, int32_t* nan_flag, const int32_t* other_idx, double* chain_out,
                             double* lp_out, void* stream);

/* bi_stretch_randoms_philox: the chain-independent numbers of `nsteps` steps generated ON the device (Philox4x32-10,
 * counter = (proposal, half, step), key = seed), shaped (nsteps, 2, n_half) like the host-drawn arrays the two calls
 * above consume: s_idx / c_idx (int32), zz, factors = (d - 1) log zz, logu.  Step `step0 + t` always yields the same
 * numbers whatever the block size.  The ensemble is split into even / odd walkers (fixed halves), so this is not
 * emcee's stream: the chain is a valid stretch-move chain but differs from the reference's for the same seed. */
int bi_stretch_randoms_philox(bi_handle h, uint64_t seed, uint64_t step0, int32_t nsteps, int32_t n_half, double a,
                              int32_t* s_idx, int32_t* c_idx, double* zz, double* factors, double* logu, void* stream);

/* ---- several experimental data sets on one emulator image (closure tests, steer_analysis.py:168-183) ----------------
 * The reference validates an analysis by running one MCMC per validation design point, each against that point's
 * pseudo-data (data_IO.py:362-372) and the SAME emulators, one after the other.  Here the pseudo-data sets are loaded
 * side by side and every walker names the set it is evaluated against, so all closure chains advance in one call.
 *
 * bi_set_data_sets: y, y_err (n_sets, F) HOST pointers, Sigma_exp = diag(y_err^2) per set; replaces earlier sets.
 *                   The handle must have been created with experimental data (it sizes the scratch buffers).
 * bi_log_posterior_sets: as bi_log_posterior with set_id (W) int32 DEVICE pointer, values in [0, n_sets);
 *                   n_div must be given (> 0; 1 for MCMC); modes BI_MODE_DENSE and BI_MODE_STRUCTURED. */
int bi_set_data_sets(bi_handle h, int32_t n_sets, const double* y, const double* y_err);
int bi_log_posterior_sets(bi_handle h, const double* theta, const int32_t* set_id, int64_t W, double n_div, int mode,
                          double* logp_out, int32_t* info_out, void* stream);

/* ---- emulator training objective (emulation.py:128-172 -> sklearn GaussianProcessRegressor.fit, _gpr.py:296-366) -------
 * The reference fits one GP per principal component with n_restarts optimiser starts; each objective evaluation is
 * log_marginal_likelihood(theta, eval_gradient=True) (_gpr.py:541-639).  A trainer holds the design X (N, d) and the
 * targets Y (n_y, N) (rows = principal-component scores, Y_pca_truncated.T) on the device and evaluates any number of
 * hyper-parameter vectors in one launch.  kind = BI_KERNEL_* (Matern 1/2, 3/2, 5/2, RBF); has_const / has_white say
 * whether the kernel is "+ ConstantKernel" / "+ WhiteKernel"; alpha is the jitter GaussianProcessRegressor(alpha=...)
 * adds to the diagonal.  theta (B, n_hyper) is scikit-learn's kernel.theta: log length scales [d], then
 * log constant_value, then log noise_level (those that exist).  y_index (B) names the row of Y per vector
 * (NULL: vector b uses row b, B <= n_y).
 *
 * bi_trainer_lml   : lml_out (B), grad_out (B, n_hyper).  A matrix that is not positive definite gives -inf and a zero
 *                    gradient, as scikit-learn does (_gpr.py:586-589).
 * bi_trainer_factor: what fit() keeps for predict (_gpr.py:335-366): L_out (B, N, N) = chol(K + alpha I) lower,
 *                    alpha_out (B, N) = K^-1 y, lml_out (B).
 * The plain calls take DEVICE pointers and enqueue on `stream`; the _host calls take HOST pointers and synchronise.
 * N <= 991, d <= 16. */
typedef struct bi_trainer_s* bi_trainer;
int bi_trainer_create(const double* X, int32_t N, int32_t d, const double* Y, int32_t n_y, int32_t kind,
                      int32_t has_const, int32_t has_white, double alpha, int32_t device, bi_trainer* out);
void bi_trainer_destroy(bi_trainer t);
const char* bi_trainer_last_error(bi_trainer t);
int64_t bi_trainer_launches(bi_trainer t);
int bi_trainer_lml(bi_trainer t, const double* theta, const int32_t* y_index, int64_t B, double* lml_out,
                   double* grad_out, void* stream);
int bi_trainer_lml_host(bi_trainer t, const double* theta, const int32_t* y_index, int64_t B, double* lml_out,
                        double* grad_out);
int bi_trainer_factor(bi_trainer t, const double* theta, const int32_t* y_index, int64_t B, double* lml_out,
                      double* L_out, double* alpha_out, void* stream);
int bi_trainer_factor_host(bi_trainer t, const double* theta, const int32_t* y_index, int64_t B, double* lml_out,
                           double* L_out, double* alpha_out);

#ifdef __cplusplus
}
#endif
#endif /* BI_B200_H */
