"""Emulator training on the device: gp_lml_kernel through the C-ABI against scikit-learn's stored
values and the numpy oracle; the fitted state against numpy's Cholesky; the device fit against scikit-learn's fit."""
from __future__ import annotations

import os

import numpy as np
import pytest

from bayesian_inference_b200 import training as tr
from oracle import gp_lml

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "gp_train.npz")
TOL_LML = 1e-9        # relative, the path's tolerance (BASELINE north_star)
TOL_GRAD = 1e-8       # relative to the largest gradient component of the batch


def cases():
    z = np.load(GOLD)
    for i in range(int(z["n_cases"])):
        kind, hc, hw, d = (int(v) for v in z[f"c{i}_meta"])
        yield i, z, tr.KernelSpec(kind, d, bool(hc), bool(hw), z[f"c{i}_theta0"], z[f"c{i}_bounds"])


def test_lml_and_gradient_match_sklearn_golden():
    for i, z, spec in cases():
        with tr.DeviceGPTrainer(z[f"c{i}_X"], z[f"c{i}_Y"], spec, float(z[f"c{i}_alpha"])) as t:
            lml, grad = t.log_marginal_likelihood(z[f"c{i}_theta"], z[f"c{i}_yidx"])
        want, wg = z[f"c{i}_lml"], z[f"c{i}_grad"]
        assert np.max(np.abs(lml - want) / np.abs(want)) < TOL_LML, (i, lml, want)
        assert np.max(np.abs(grad - wg)) < TOL_GRAD * np.max(np.abs(wg)), (i, grad, wg)


@pytest.mark.parametrize("N,d,kind", [(230, 6, 1), (97, 3, 2), (500, 8, 1), (33, 2, 3), (32, 4, 0), (31, 16, 1), (640, 5, 2)])
def test_lml_matches_oracle_at_reference_sizes(N, d, kind):
    """n_design = 230 (C3) and 500 (C5), ragged and exact tile counts; 40 random hyper-parameter vectors inside the
    reference's bounds, more vectors than SMs x slots so that the persistent loop wraps."""
    rng = np.random.RandomState(N + d)
    hi = rng.uniform(0.5, 3.0, d)
    X = rng.uniform(0, hi, (N, d))
    Y = np.sin(X / hi @ rng.normal(size=(d, 4)) * 2.0).T + 0.2 * rng.normal(size=(4, N))
    active = {("rbf" if kind == 3 else "matern"): {"nu": {0: 0.5, 1: 1.5, 2: 2.5, 3: 0}[kind], "length_scale_bounds_factor": [0.1, 10]},
              "constant": {"constant_value": 1.0, "constant_value_bounds": [1e-3, 10]},
              "noise": {"args": {"noise_level": 0.25, "noise_level_bounds": [1e-4, 1]}}}
    spec = tr.kernel_spec(active, np.zeros(d), hi)
    B = 200 if N <= 100 else (40 if N <= 500 else 6)
    theta = rng.uniform(spec.bounds[:, 0], spec.bounds[:, 1], (B, spec.n_hyper))
    yidx = rng.randint(0, 4, B).astype(np.int32)
    with tr.DeviceGPTrainer(X, Y, spec, 1e-10) as t:
        lml, grad = t.log_marginal_likelihood(theta, yidx)
        lml2, L, al = t.factor(theta[:6], yidx[:6])
    nchk = B if N <= 100 else min(B, 12)
    want, wg = gp_lml.log_marginal_likelihood_batch(X, Y, theta[:nchk], kind, True, True, 1e-10, yidx[:nchk])
    assert np.max(np.abs(lml[:nchk] - want) / np.abs(want)) < TOL_LML
    assert np.max(np.abs(grad[:nchk] - wg)) < TOL_GRAD * np.max(np.abs(wg))
    assert np.all(np.isfinite(lml)) and np.all(np.isfinite(grad))
    assert np.max(np.abs(lml2 - lml[:6]) / np.abs(lml[:6])) < 1e-13          # same factorisation, other output path
    for b in range(6):
        K, _ = gp_lml.kernel_matrix(X, theta[b], kind, True, True, eval_gradient=False)
        K[np.diag_indices(N)] += 1e-10
        Lw = np.linalg.cholesky(K)
        assert np.max(np.abs(L[b] - Lw)) < 1e-10 * np.max(np.abs(Lw))
        assert np.all(np.triu(L[b], 1) == 0.0)
        aw = np.linalg.solve(K, Y[yidx[b]])
        assert np.max(np.abs(al[b] - aw)) < 1e-7 * np.max(np.abs(aw))


def test_not_positive_definite_gives_minus_inf_and_zero_gradient():
    """scikit-learn returns (-inf, 0) when the Cholesky fails (_gpr.py:586-589): duplicate design points, no jitter."""
    rng = np.random.RandomState(0)
    X = rng.uniform(0, 1, (40, 2))
    X[7] = X[3]
    Y = rng.normal(size=(1, 40))
    spec = tr.KernelSpec(3, 2, False, False, np.zeros(2), np.array([[-2.0, 2.0]] * 2))
    with tr.DeviceGPTrainer(X, Y, spec, 0.0) as t:
        lml, grad = t.log_marginal_likelihood(np.array([[1.5, 1.5], [0.0, 0.0]]), np.zeros(2, np.int32))
    assert lml[0] == -np.inf and np.all(grad[0] == 0.0)


def test_device_fit_is_not_worse_than_sklearn_fit():
    for i, z, spec in cases():
        if f"c{i}_fit_lml" not in z:
            continue
        fit = tr.fit_gps(z[f"c{i}_X"], z[f"c{i}_Y"], spec, float(z[f"c{i}_alpha"]), n_restarts=8, seed=1)
        want = z[f"c{i}_fit_lml"]
        assert np.all(fit.lml >= want - 1e-6 * np.abs(want)), (i, fit.lml, want)
        assert fit.L.shape == (len(want), z[f"c{i}_X"].shape[0], z[f"c{i}_X"].shape[0])


def test_trained_group_drives_the_log_posterior():
    """fit_emulation_group -> GroupImage -> DeviceEmulator: a group trained on the device predicts like scikit-learn GPs
    refitted (optimizer=None) at the same hyper-parameters -- the reference's predict path on device-trained emulators."""
    from sklearn.gaussian_process import GaussianProcessRegressor
    from sklearn.gaussian_process import kernels as sk

    from bayesian_inference_b200 import image
    from bayesian_inference_b200.device import DeviceEmulator

    rng = np.random.RandomState(5)
    N, d, F, P = 60, 3, 24, 4
    hi = np.array([1.0, 2.0, 0.5])
    X = rng.uniform(0, hi, (N, d))
    Yobs = np.sin(X / hi @ rng.normal(size=(d, F)) * 1.5) + 0.05 * rng.normal(size=(N, F))
    spec = tr.kernel_spec({"matern": {"nu": 1.5, "length_scale_bounds_factor": [0.1, 10]},
                           "noise": {"args": {"noise_level": 0.25, "noise_level_bounds": [1e-4, 1]}}}, np.zeros(d), hi)
    grp, gps = tr.fit_emulation_group("g", X, Yobs, P, spec, 1e-10, n_restarts=4, seed=0)
    theta = rng.uniform(0.1, 0.9, (16, d)) * hi
    img = image.EmulatorImage(d=d, min=np.zeros(d), max=hi, groups=[grp], blocks=np.array([[0, 0, 0, F]], np.int64),
                              n_features=F)
    with DeviceEmulator(img) as dev:
        m, v = dev.predict_pc(theta)
    Ypca = tr.pca_transform(Yobs, P)[0]
    for p in range(P):
        k = sk.Matern(length_scale=gps.ell[p], nu=1.5) + sk.WhiteKernel(gps.noise[p])
        g = GaussianProcessRegressor(kernel=k, alpha=1e-10, optimizer=None).fit(X, Ypca[:, p])
        mw, sw = g.predict(theta, return_std=True)
        assert np.max(np.abs(m[:, p] - mw)) < 1e-8 * max(1.0, np.max(np.abs(mw)))
        assert np.max(np.abs(v[:, p] - sw ** 2)) < 1e-8 * max(1.0, np.max(sw ** 2))


def test_objective_is_deterministic_and_batch_independent():
    """The same hyper-parameter vector gives bit-identical results on every call, alone or inside any batch, in any
    slot of the persistent grid (a data race or a stale scratch tile would show up here)."""
    rng = np.random.RandomState(11)
    N, d = 150, 5
    X = rng.uniform(0, 1, (N, d))
    Y = np.sin(X @ rng.normal(size=(d, 3)) * 2.0).T + 0.2 * rng.normal(size=(3, N))
    spec = tr.kernel_spec({"matern": {"nu": 2.5, "length_scale_bounds_factor": [0.1, 10]},
                           "constant": {"constant_value": 1.0, "constant_value_bounds": [1e-3, 10]},
                           "noise": {"args": {"noise_level": 0.25, "noise_level_bounds": [1e-4, 1]}}}, np.zeros(d), np.ones(d))
    B = 700                                                    # more than two waves of the 296 resident CTAs
    theta = rng.uniform(spec.bounds[:, 0], spec.bounds[:, 1], (B, spec.n_hyper))
    yidx = rng.randint(0, 3, B).astype(np.int32)
    with tr.DeviceGPTrainer(X, Y, spec, 1e-10) as t:
        l1, g1 = t.log_marginal_likelihood(theta, yidx)
        l2, g2 = t.log_marginal_likelihood(theta, yidx)
        perm = rng.permutation(B)
        l3, g3 = t.log_marginal_likelihood(theta[perm], yidx[perm])
        l4, g4 = t.log_marginal_likelihood(theta[:5], yidx[:5])
        f1 = t.factor(theta[:9], yidx[:9])
        f2 = t.factor(theta[:9][::-1].copy(), yidx[:9][::-1].copy())
    assert np.array_equal(l1, l2) and np.array_equal(g1, g2)
    assert np.array_equal(l1[perm], l3) and np.array_equal(g1[perm], g3)
    assert np.array_equal(l1[:5], l4) and np.array_equal(g1[:5], g4)
    assert all(np.array_equal(a, b[::-1]) for a, b in zip(f1, f2))
