"""ORACLE (test infrastructure, not product code) -- closed-form numpy restatement of the MCMC log-posterior path.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may
import this module.  The product path (``bayesian_inference_b200``) never does and has no CPU fallback.

It restates, on flat float64 arrays (the ``EmulatorImage.to_flat()`` / ``.npz`` layout), the arithmetic that the
reference delegates to scikit-learn / scipy, following these sources line by line:

* kernel vector        sklearn ``Matern.__call__`` kernels.py:1685-1712, ``RBF.__call__`` :1530-1545,
                       ``ConstantKernel`` :1244-1280, ``WhiteKernel`` :1374-1419 (0 off the diagonal), ``Sum`` :838-873
* GP mean / variance   sklearn ``GaussianProcessRegressor.predict`` _gpr.py:444-500 (forward substitution with ``L_``,
                       ``diag - sum V^2``, clip at 0, sqrt then the reference squares it again: emulation.py:497-499)
* inverse PCA          reference ``emulation.predict_emulation_group`` emulation.py:506-509 (mean, no ``pca.mean_``),
                       :522-526 (S diag(v) S^T), :531-532 (unexplained covariance / rows-in-call), :538-539 (scale)
* unexplained cov      reference ``compute_emulator_group_cov_unexplained`` emulation.py:227-251
* group merge          reference ``SortEmulationGroupObservables.convert`` emulation.py:346-406 + ``nd_block_diag`` :254-270
* log-likelihood       reference ``log_posterior._loglikelihood`` log_posterior.py:104-146 (dpotrf / dpotrs, un-normalised)
* prior + assembly     reference ``log_posterior.log_posterior`` log_posterior.py:42-101 (strict bounds, + diag(y_err^2))

Parity status: PINNED against the reference executed in the build container (tests/golden/make_golden.py imports
/root/reference and stores its outputs under tests/golden/); the reference's own test-suite holds no vectors for
this path.
"""
from __future__ import annotations

from typing import Mapping

import numpy as np
from scipy.linalg import lapack, solve_triangular

SQRT3 = np.sqrt(3.0)
SQRT5 = np.sqrt(5.0)


def n_groups(z: Mapping[str, np.ndarray]) -> int:
    return int(z["n_groups"])


def kernel_vectors(z: Mapping[str, np.ndarray], g: int, theta: np.ndarray) -> np.ndarray:
    """k[w, p, j] = kernel_p(theta_w, X_j): (W, P, N).  Direct differences like scipy ``cdist`` (no norm trick)."""
    X = z[f"g{g}_X"]
    ell = z[f"g{g}_ell"]
    kind = z[f"g{g}_kind"]
    const = z[f"g{g}_const"]
    theta = np.atleast_2d(np.asarray(theta, dtype=np.float64))
    W, P, N = theta.shape[0], ell.shape[0], X.shape[0]
    out = np.empty((W, P, N))
    for p in range(P):
        a = theta / ell[p]            # sklearn scales both operands before cdist
        b = X / ell[p]
        diff = a[:, None, :] - b[None, :, :]
        sq = np.zeros((W, N))
        for k in range(diff.shape[2]):  # sequential accumulation over parameters, as cdist's inner loop
            sq += diff[:, :, k] * diff[:, :, k]
        if kind[p] == 3:              # RBF: exp(-0.5 * sqeuclidean)
            kv = np.exp(-0.5 * sq)
        else:
            r = np.sqrt(sq)
            if kind[p] == 0:
                kv = np.exp(-r)
            elif kind[p] == 1:
                t = r * SQRT3
                kv = (1.0 + t) * np.exp(-t)
            elif kind[p] == 2:
                t = r * SQRT5
                kv = (1.0 + t + t ** 2 / 3.0) * np.exp(-t)
            else:
                raise ValueError(f"unknown kernel kind {kind[p]}")
        out[:, p, :] = kv + const[p]
    return out


def predict_pc(z: Mapping[str, np.ndarray], g: int, theta: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
    """Per-PC GP mean and variance, (W, P) each."""
    k = kernel_vectors(z, g, theta)
    alpha, L = z[f"g{g}_alpha"], z[f"g{g}_L"]
    const, noise = z[f"g{g}_const"], z[f"g{g}_noise"]
    W, P, _ = k.shape
    m = np.empty((W, P))
    v = np.empty((W, P))
    for p in range(P):
        m[:, p] = k[:, p, :] @ alpha[p]
        V = solve_triangular(L[p], k[:, p, :].T, lower=True, check_finite=False)   # (N, W)
        var = (1.0 + const[p] + noise[p]) - np.einsum("ij,ij->j", V, V)
        var[var < 0] = 0.0
        v[:, p] = np.sqrt(var) ** 2        # the reference squares sklearn's std (emulation.py:499)
    return m, v


def cov_unexplained(z: Mapping[str, np.ndarray], g: int) -> np.ndarray:
    Sr, ev = z[f"g{g}_S_rest"], z[f"g{g}_ev_rest"]
    return Sr.T.dot(np.diag(ev).dot(Sr))


def predict_group(z: Mapping[str, np.ndarray], g: int, theta: np.ndarray, n_div: float | None = None):
    """(mean (W,F_g), cov (W,F_g,F_g)) of one emulation group.  ``n_div=None`` -> rows in the call (reference, F3)."""
    theta = np.atleast_2d(np.asarray(theta, dtype=np.float64))
    m, v = predict_pc(z, g, theta)
    S, scale, mean = z[f"g{g}_S"], z[f"g{g}_scale"], z[f"g{g}_mean"]
    W = theta.shape[0]
    if n_div is None:
        n_div = float(W)
    mu = m.dot(S) * scale + mean
    U = cov_unexplained(z, g) / n_div
    St = S.T
    cov = np.empty((W, S.shape[1], S.shape[1]))
    ss = np.outer(scale, scale)
    for w in range(W):
        cov[w] = (St.dot(np.diag(v[w]).dot(S)) + U) * ss
    return mu, cov


def predict(z: Mapping[str, np.ndarray], theta: np.ndarray, n_div: float | None = None):
    """Merged (central_value (W,F), cov (W,F,F)): block-diagonal per observable."""
    theta = np.atleast_2d(np.asarray(theta, dtype=np.float64))
    W, F = theta.shape[0], int(z["n_features"])
    per_group = [predict_group(z, g, theta, n_div) for g in range(n_groups(z))]
    mean = np.zeros((W, F))
    cov = np.zeros((W, F, F))
    for g, so, si, ln in np.asarray(z["blocks"]).reshape(-1, 4):
        mean[:, so:so + ln] = per_group[g][0][:, si:si + ln]
        cov[:, so:so + ln, so:so + ln] = per_group[g][1][:, si:si + ln, si:si + ln]
    return mean, cov


def sigma_exp(z: Mapping[str, np.ndarray]) -> np.ndarray:
    if "sigma_exp" in z:
        return np.asarray(z["sigma_exp"])
    return np.diag(np.asarray(z["y_err"]) ** 2)


def log_likelihood(dy: np.ndarray, cov: np.ndarray) -> float:
    """-1/2 dy^T C^-1 dy - sum log diag chol(C); NaN when C is not positive definite (the reference
    does not detect info>0 and ends in NaN/-inf through ``np.log`` of a bad pivot; the oracle makes that explicit)."""
    Lf, info = lapack.dpotrf(cov, clean=False)
    if info < 0:
        raise ValueError(f"dpotrf: illegal argument {-info}")
    if info > 0:
        return float("nan")
    a, info = lapack.dpotrs(Lf, dy)
    if info != 0:
        raise ValueError(f"dpotrs: illegal argument {-info}")
    return float(-0.5 * np.dot(dy, a) - np.log(Lf.diagonal()).sum())


def log_posterior(z: Mapping[str, np.ndarray], theta: np.ndarray, n_div: float | None = None) -> np.ndarray:
    """(W,) log-posterior.  ``n_div=None`` reproduces a batched reference call (divisor = rows inside the prior box);
    ``n_div=1`` reproduces what emcee evaluates in the reference's production run (one row per call)."""
    theta = np.atleast_2d(np.asarray(theta, dtype=np.float64))
    out = np.zeros(theta.shape[0])
    inside = np.all((theta > z["min"]) & (theta < z["max"]), axis=1)
    out[~inside] = -np.inf
    n_in = int(np.count_nonzero(inside))
    if n_in == 0:
        return out
    mean, cov = predict(z, theta[inside], n_div if n_div is not None else float(n_in))
    dy = mean - z["y"]
    C = cov + sigma_exp(z)
    out[inside] = [log_likelihood(dy[i], C[i]) for i in range(n_in)]
    return out


# ---- structure-exploiting evaluation (valid when Sigma_exp has no entries across observable blocks) -----------
def log_posterior_blockwise(z: Mapping[str, np.ndarray], theta: np.ndarray, n_div: float = 1.0) -> np.ndarray:
    """Same value as ``log_posterior`` computed block by block (structured variant): the merged
    covariance is block diagonal per observable, so log-det and quadratic form are sums over observables."""
    theta = np.atleast_2d(np.asarray(theta, dtype=np.float64))
    out = np.zeros(theta.shape[0])
    inside = np.all((theta > z["min"]) & (theta < z["max"]), axis=1)
    out[~inside] = -np.inf
    idx = np.nonzero(inside)[0]
    if idx.size == 0:
        return out
    Se = sigma_exp(z)
    per_group = [predict_group(z, g, theta[idx], n_div) for g in range(n_groups(z))]
    acc = np.zeros(idx.size)
    for g, so, si, ln in np.asarray(z["blocks"]).reshape(-1, 4):
        mu = per_group[g][0][:, si:si + ln]
        cv = per_group[g][1][:, si:si + ln, si:si + ln] + Se[so:so + ln, so:so + ln]
        dy = mu - z["y"][so:so + ln]
        for i in range(idx.size):
            acc[i] += log_likelihood(dy[i], cv[i])
    out[idx] = acc
    return out
