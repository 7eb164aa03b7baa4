"""ORACLE (test infrastructure, not product code) -- numpy restatement of the GP hyper-parameter objective the
reference optimises when it trains its emulators.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` legs of the benchmark tools may import this
module.  The product path (``bayesian_inference_b200.training``) never does.

The reference builds ``Matern | RBF [+ ConstantKernel] [+ WhiteKernel]`` and calls
``GaussianProcessRegressor(kernel, alpha, n_restarts_optimizer).fit(design, y)`` once per principal component
(reference emulation.py:128-172).  scikit-learn is a dependency, not part of /root/reference; what is restated here,
from scikit-learn 1.9.0 as installed in the build container:

* objective + gradient   ``GaussianProcessRegressor.log_marginal_likelihood`` _gpr.py:541-639 (Cholesky of K + alpha I,
                         ``cho_solve`` for alpha and for K^-1, ``0.5 * einsum(alpha alpha^T - K^-1, dK/dtheta)``)
* kernels + gradients    ``Matern.__call__`` kernels.py (nu = 0.5 / 1.5 / 2.5, gradients w.r.t. log length scales),
                         ``RBF.__call__``, ``ConstantKernel.__call__`` (gradient = constant_value everywhere),
                         ``WhiteKernel.__call__`` (gradient = noise_level on the diagonal), ``Sum`` (concatenated theta)
* the fit                ``GaussianProcessRegressor.fit`` _gpr.py:296-366 + ``_constrained_optimization``: L-BFGS-B on
                         -LML in log-parameter space within ``kernel.bounds``, first start = the kernel's initial
                         theta, restarts drawn log-uniformly from the bounds, best LML kept

theta is in scikit-learn's order: ``log(length_scale[0..d-1])``, then ``log(constant_value)`` when the kernel has a
ConstantKernel, then ``log(noise_level)`` when it has a WhiteKernel.

Parity status: PINNED against scikit-learn executed in the build container (tests/golden/make_gp_train_golden.py stores
``log_marginal_likelihood(theta, eval_gradient=True)`` of the reference's kernel families in tests/golden/gp_train.npz).
"""
from __future__ import annotations

import numpy as np
from scipy.linalg import cho_solve, cholesky
from scipy.optimize import minimize

KIND_MATERN12, KIND_MATERN32, KIND_MATERN52, KIND_RBF = 0, 1, 2, 3


def n_hyper(d: int, has_const: bool, has_white: bool) -> int:
    return d + int(bool(has_const)) + int(bool(has_white))


def kernel_matrix(X: np.ndarray, theta: np.ndarray, kind: int, has_const: bool, has_white: bool, eval_gradient: bool = True):
    """K (N, N) and dK/dtheta (N, N, n_hyper) exactly as ``kernel(X, eval_gradient=True)`` lays them out."""
    X = np.asarray(X, float)
    N, d = X.shape
    ell = np.exp(theta[:d])
    Xs = X / ell
    diff = Xs[:, None, :] - Xs[None, :, :]
    D = diff * diff                                  # (N, N, d)   squared scaled differences
    sq = D.sum(-1)
    r = np.sqrt(sq)
    if kind == KIND_MATERN12:
        K = np.exp(-r)
        with np.errstate(divide="ignore", invalid="ignore"):
            G = K[..., None] * D / r[..., None]
        G[~np.isfinite(G)] = 0.0
    elif kind == KIND_MATERN32:
        t = r * np.sqrt(3.0)
        K = (1.0 + t) * np.exp(-t)
        G = 3.0 * D * np.exp(-t)[..., None]
    elif kind == KIND_MATERN52:
        t = r * np.sqrt(5.0)
        K = (1.0 + t + t * t / 3.0) * np.exp(-t)
        G = 5.0 / 3.0 * D * ((t + 1.0) * np.exp(-t))[..., None]
    elif kind == KIND_RBF:
        K = np.exp(-0.5 * sq)
        G = D * K[..., None]
    else:
        raise ValueError(f"unknown kernel kind {kind}")
    K = K.copy()
    grads = [G]
    k = d
    if has_const:
        c = np.exp(theta[k]); k += 1
        K += c
        grads.append(np.full((N, N, 1), c))
    if has_white:
        eta = np.exp(theta[k]); k += 1
        K[np.diag_indices(N)] += eta
        grads.append((eta * np.eye(N))[..., None])
    if not eval_gradient:
        return K, None
    return K, np.concatenate(grads, axis=-1)


def log_marginal_likelihood(X, y, theta, kind, has_const, has_white, alpha):
    """(lml, grad) for one hyper-parameter vector.  A matrix that is not positive definite gives (-inf, zeros)."""
    theta = np.asarray(theta, float)
    K, Kg = kernel_matrix(X, theta, kind, has_const, has_white)
    N = K.shape[0]
    K[np.diag_indices(N)] += alpha
    try:
        L = cholesky(K, lower=True, check_finite=False)
    except np.linalg.LinAlgError:
        return -np.inf, np.zeros_like(theta)
    y = np.asarray(y, float)
    a = cho_solve((L, True), y, check_finite=False)
    lml = -0.5 * (y @ a) - np.log(np.diag(L)).sum() - N / 2.0 * np.log(2.0 * np.pi)
    Kinv = cho_solve((L, True), np.eye(N), check_finite=False)
    inner = np.outer(a, a) - Kinv
    grad = 0.5 * np.einsum("ij,jik->k", inner, Kg)
    return float(lml), grad


def log_marginal_likelihood_batch(X, Y, thetas, kind, has_const, has_white, alpha, y_index=None):
    """Rows of ``thetas`` (B, n_hyper) evaluated against ``Y[y_index[b]]`` (default: row b of Y)."""
    thetas = np.atleast_2d(thetas)
    B = thetas.shape[0]
    y_index = np.arange(B) if y_index is None else np.asarray(y_index)
    lml = np.empty(B)
    grad = np.empty_like(thetas, dtype=float)
    for b in range(B):
        lml[b], grad[b] = log_marginal_likelihood(X, Y[y_index[b]], thetas[b], kind, has_const, has_white, alpha)
    return lml, grad


def fit(X, y, theta0, bounds, kind, has_const, has_white, alpha, n_restarts=0, rng=None):
    """scikit-learn's fit loop: L-BFGS-B from ``theta0`` and from ``n_restarts`` log-uniform draws, best LML wins.
    Returns (theta_best, lml_best, n_objective_evaluations)."""
    bounds = np.asarray(bounds, float)
    rng = np.random.RandomState(0) if rng is None else rng
    nfev = 0

    def obj(th):
        nonlocal nfev
        nfev += 1
        lml, g = log_marginal_likelihood(X, y, th, kind, has_const, has_white, alpha)
        return -lml, -g

    starts = [np.asarray(theta0, float)] + [rng.uniform(bounds[:, 0], bounds[:, 1]) for _ in range(n_restarts)]
    best = (None, np.inf)
    for s in starts:
        res = minimize(obj, s, method="L-BFGS-B", jac=True, bounds=bounds)
        if res.fun < best[1]:
            best = (res.x, res.fun)
    return best[0], -best[1], nfev
