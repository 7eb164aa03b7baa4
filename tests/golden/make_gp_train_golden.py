"""Generate tests/golden/gp_train.npz by EXECUTING scikit-learn (run in the build container only).

    python tests/golden/make_gp_train_golden.py

The reference trains its emulators with ``GaussianProcessRegressor(kernel, alpha, n_restarts_optimizer).fit(design, y)``
(reference emulation.py:128-172); the objective is scikit-learn's ``log_marginal_likelihood(theta, eval_gradient=True)``
(_gpr.py:541-639).  For every kernel family the reference can configure (Matern nu = 0.5 / 1.5 / 2.5 or RBF, with and
without ConstantKernel / WhiteKernel) this stores, for a handful of hyper-parameter vectors inside the bounds:

    c{i}_X (N, d), c{i}_Y (n_y, N), c{i}_theta (B, n_hyper), c{i}_yidx (B,), c{i}_meta = [kind, has_const, has_white, d]
    c{i}_alpha, c{i}_bounds (n_hyper, 2)
    c{i}_lml (B,), c{i}_grad (B, n_hyper)       <- scikit-learn's values, the pin
    c{i}_fit_theta (n_y, n_hyper), c{i}_fit_lml (n_y,)   <- .fit() with restarts (fit-quality yardstick, cases 0-2 only)
"""
from __future__ import annotations

import os

import numpy as np
from sklearn.gaussian_process import GaussianProcessRegressor
from sklearn.gaussian_process import kernels as sk

HERE = os.path.dirname(os.path.abspath(__file__))

# (kind, nu or None, has_const, has_white, N, d, n_y, B, alpha, fit?)
CASES = [
    (1, 1.5, True, True, 48, 3, 3, 6, 1e-10, True),      # the reference's production kernel (config/*.yaml)
    (2, 2.5, False, True, 70, 6, 2, 4, 1e-10, True),
    (3, None, True, False, 37, 2, 2, 4, 1e-3, True),      # ragged N, RBF + constant, jitter only
    (0, 0.5, False, False, 33, 4, 2, 4, 1e-6, False),
    (1, 1.5, False, False, 100, 5, 2, 3, 1e-6, False),
    (2, 2.5, True, True, 64, 3, 2, 3, 1e-10, False),
    (3, None, False, True, 96, 6, 2, 3, 1e-10, False),
]


def make_kernel(kind, nu, has_const, has_white, lo, hi):
    ell = hi - lo
    bounds = np.outer(ell, (0.1, 10.0))
    k = sk.RBF(length_scale=ell, length_scale_bounds=bounds) if kind == 3 else \
        sk.Matern(length_scale=ell, length_scale_bounds=bounds, nu=nu)
    if has_const:
        k = k + sk.ConstantKernel(constant_value=1.0, constant_value_bounds=(1e-3, 10.0))
    if has_white:
        k = k + sk.WhiteKernel(noise_level=0.25, noise_level_bounds=(1e-4, 1.0))
    return k


def main():
    out = {"n_cases": np.int64(len(CASES))}
    for i, (kind, nu, hc, hw, N, d, n_y, B, alpha, do_fit) in enumerate(CASES):
        rng = np.random.RandomState(100 + i)
        lo, hi = np.zeros(d), rng.uniform(0.5, 3.0, d)
        X = rng.uniform(lo, hi, (N, d))
        # smooth functions of the design + noise, roughly unit variance like PCA scores of standardised data
        Wm = rng.normal(size=(n_y, d))
        Y = np.sin(X / hi @ Wm.T * 2.0).T + 0.3 * rng.normal(size=(n_y, N))
        kernel = make_kernel(kind, nu, hc, hw, lo, hi)
        bounds = kernel.bounds
        theta = rng.uniform(bounds[:, 0], bounds[:, 1], (B, bounds.shape[0]))
        theta[0] = kernel.theta
        yidx = rng.randint(0, n_y, B)
        lml = np.empty(B)
        grad = np.empty_like(theta)
        for b in range(B):
            gpr = GaussianProcessRegressor(kernel=kernel, alpha=alpha, optimizer=None).fit(X, Y[yidx[b]])
            lml[b], grad[b] = gpr.log_marginal_likelihood(theta[b], eval_gradient=True)
        out.update({f"c{i}_X": X, f"c{i}_Y": Y, f"c{i}_theta": theta, f"c{i}_yidx": yidx.astype(np.int64),
                    f"c{i}_meta": np.array([kind, int(hc), int(hw), d], np.int64), f"c{i}_alpha": np.float64(alpha),
                    f"c{i}_bounds": bounds, f"c{i}_theta0": kernel.theta, f"c{i}_lml": lml, f"c{i}_grad": grad})
        if do_fit:
            ft, fl = [], []
            for j in range(n_y):
                gpr = GaussianProcessRegressor(kernel=kernel, alpha=alpha, n_restarts_optimizer=8, random_state=j).fit(X, Y[j])
                ft.append(gpr.kernel_.theta)
                fl.append(gpr.log_marginal_likelihood_value_)
            out[f"c{i}_fit_theta"] = np.array(ft)
            out[f"c{i}_fit_lml"] = np.array(fl)
        print(f"case {i}: kind {kind} N {N} d {d}  lml {lml}")
    np.savez_compressed(os.path.join(HERE, "gp_train.npz"), **out)


if __name__ == "__main__":
    main()
