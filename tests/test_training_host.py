"""Emulator training, CPU side: the oracle against scikit-learn's stored values, the batched
optimiser and the kernel construction.  No GPU."""
from __future__ import annotations

import os

import numpy as np
import pytest

from bayesian_inference_b200 import training as tr
from oracle import gp_lml

GOLD = os.path.join(os.path.dirname(__file__), "golden", "gp_train.npz")


def cases():
    z = np.load(GOLD)
    for i in range(int(z["n_cases"])):
        kind, hc, hw, d = (int(v) for v in z[f"c{i}_meta"])
        yield i, z, kind, bool(hc), bool(hw), d


def test_oracle_matches_sklearn_log_marginal_likelihood():
    """The numpy restatement against GaussianProcessRegressor.log_marginal_likelihood(theta, eval_gradient=True)
    for every kernel family the reference can configure (fixture made by tests/golden/make_gp_train_golden.py)."""
    for i, z, kind, hc, hw, d in cases():
        lml, grad = gp_lml.log_marginal_likelihood_batch(z[f"c{i}_X"], z[f"c{i}_Y"], z[f"c{i}_theta"], kind, hc, hw,
                                                         float(z[f"c{i}_alpha"]), z[f"c{i}_yidx"])
        assert np.max(np.abs(lml - z[f"c{i}_lml"]) / np.abs(z[f"c{i}_lml"])) < 1e-12, i
        assert np.max(np.abs(grad - z[f"c{i}_grad"])) < 1e-10 * np.max(np.abs(z[f"c{i}_grad"])), i


def test_oracle_gradient_is_the_derivative():
    i, z, kind, hc, hw, d = next(cases())
    X, y, th, al = z["c0_X"], z["c0_Y"][0], z["c0_theta"][2], float(z["c0_alpha"])
    _, g = gp_lml.log_marginal_likelihood(X, y, th, kind, hc, hw, al)
    for m in range(len(th)):
        e = np.zeros_like(th)
        e[m] = 1e-6
        fd = (gp_lml.log_marginal_likelihood(X, y, th + e, kind, hc, hw, al)[0]
              - gp_lml.log_marginal_likelihood(X, y, th - e, kind, hc, hw, al)[0]) / 2e-6
        assert abs(fd - g[m]) < 1e-5 * max(1.0, abs(g[m]))


def test_kernel_spec_mirrors_reference_construction():
    """emulation.py:128-162: length_scale = max - min, bounds = outer(length_scale, factor), + constant, + noise."""
    from sklearn.gaussian_process import kernels as sk
    lo, hi = np.array([0.1, 1.0, 0.0]), np.array([0.5, 4.0, 2.0])
    active = {"matern": {"nu": 1.5, "length_scale_bounds_factor": [0.01, 100]},
              "constant": {"constant_value": 1.0, "constant_value_bounds": [0.001, 10]},
              "noise": {"type": "white", "args": {"noise_level": 0.25, "noise_level_bounds": [0.0001, 1]}}}
    spec = tr.kernel_spec(active, lo, hi)
    k = (sk.Matern(length_scale=hi - lo, length_scale_bounds=np.outer(hi - lo, (0.01, 100)), nu=1.5)
         + sk.ConstantKernel(1.0, (0.001, 10)) + sk.WhiteKernel(0.25, (0.0001, 1)))
    assert (spec.kind, spec.d, spec.has_const, spec.has_white) == (1, 3, True, True)
    assert np.allclose(spec.theta0, k.theta, rtol=0, atol=1e-15) and np.allclose(spec.bounds, k.bounds, rtol=0, atol=1e-15)
    spec = tr.kernel_spec({"rbf": {"length_scale_bounds_factor": [0.1, 10]}}, lo, hi)
    assert (spec.kind, spec.n_hyper) == (3, 3)
    with pytest.raises(ValueError):
        tr.kernel_spec({"constant": {"constant_value": 1.0, "constant_value_bounds": [0.1, 1]}}, lo, hi)


def test_batched_lbfgsb_box_constrained_quadratics():
    """B independent convex quadratics with known box-constrained minimisers, some on the boundary."""
    rng = np.random.RandomState(0)
    B, n = 40, 6
    A = rng.normal(size=(B, n, n))
    H = np.einsum("bij,bkj->bik", A, A) + 0.5 * np.eye(n)
    c = rng.normal(size=(B, n)) * 2
    lo, hi = -np.ones(n), np.ones(n)

    def fun(x, idx):
        r = x - c[idx]
        Hr = np.einsum("bij,bj->bi", H[idx], r)
        return 0.5 * np.einsum("bi,bi->b", r, Hr), Hr

    x, f, nit, nb = tr.batched_lbfgsb(fun, np.zeros((B, n)), lo, hi, pgtol=1e-9, ftol=1e-15, maxiter=500)
    from scipy.optimize import minimize
    for b in range(B):
        ref = minimize(lambda v: tuple(o[0] for o in fun(v[None], np.array([b]))), np.zeros(n), jac=True, method="L-BFGS-B",
                       bounds=list(zip(lo, hi)), options={"ftol": 1e-15, "gtol": 1e-10})
        assert f[b] <= ref.fun + 1e-8 * max(1.0, abs(ref.fun)), b
        assert np.max(np.abs(x[b] - ref.x)) < 1e-4, b
    assert (x >= lo).all() and (x <= hi).all() and (np.abs(x) == 1.0).any()


def test_fit_reaches_sklearn_fit_with_oracle_objective():
    """The lock-step optimiser with restarts, driven by the oracle objective, ends at scikit-learn's fitted LML."""
    i, z, kind, hc, hw, d = next(cases())
    X, Y, al = z["c0_X"], z["c0_Y"], float(z["c0_alpha"])
    spec = tr.KernelSpec(kind, d, hc, hw, z["c0_theta0"], z["c0_bounds"])
    fit = tr.fit_gps(X, Y, spec, al, n_restarts=8, seed=1,
                     objective=lambda th, yi: gp_lml.log_marginal_likelihood_batch(X, Y, th, kind, hc, hw, al, yi))
    assert np.all(fit.lml >= z["c0_fit_lml"] - 1e-6 * np.abs(z["c0_fit_lml"]))
    assert np.all((fit.theta >= spec.bounds[:, 0]) & (fit.theta <= spec.bounds[:, 1]))
    assert fit.n_evaluations > fit.n_batches > 1


def test_pca_transform_matches_sklearn():
    from sklearn.decomposition import PCA
    from sklearn.preprocessing import StandardScaler
    rng = np.random.RandomState(3)
    Y = rng.normal(size=(40, 25)) @ rng.normal(size=(25, 25)) + rng.normal(size=25) * 5
    sc = StandardScaler()
    pca = PCA(n_components=None, svd_solver="full", whiten=False)
    want = pca.fit_transform(sc.fit_transform(Y))
    got, comps, ev, scale, mean = tr.pca_transform(Y, 6)
    assert np.allclose(scale, sc.scale_) and np.allclose(mean, sc.mean_)
    assert np.allclose(ev, pca.explained_variance_)
    sgn = np.sign(np.sum(comps * pca.components_, axis=1))        # a principal axis is defined up to its sign
    assert np.allclose(comps * sgn[:, None], pca.components_, atol=1e-10)
    assert np.allclose(got * sgn[:6], want[:, :6], atol=1e-10)
    # sign rules: 'u' = scikit-learn 1.3.0 (the reference's pin): largest |entry| of each U column positive;
    # 'v' = scikit-learn >= 1.5: largest |entry| of each components_ row positive
    gu = tr.pca_transform(Y, 25, sign_rule="u")[0]
    assert np.all(gu[np.argmax(np.abs(gu), axis=0), np.arange(gu.shape[1])] > 0)
    cv = tr.pca_transform(Y, 6, sign_rule="v")[1]
    assert np.all(cv[np.arange(cv.shape[0]), np.argmax(np.abs(cv), axis=1)] > 0)
    import sklearn
    if tuple(int(v) for v in sklearn.__version__.split(".")[:2]) >= (1, 5):
        assert np.allclose(cv, pca.components_, atol=1e-10)          # same sign, no flip needed
    with pytest.raises(ValueError):
        tr.pca_transform(Y, 6, sign_rule="x")


def test_batched_lbfgsb_reports_unconverged_problems():
    """a start still moving at maxiter is flagged instead of passing silently."""
    A = np.diag([1.0, 1e4])
    fun = lambda X, idx: (0.5 * np.einsum("bi,ij,bj->b", X, A, X), X @ A)     # noqa: E731
    x0 = np.array([[5.0, 5.0], [0.0, 0.0]])
    tr.batched_lbfgsb(fun, x0, -10 * np.ones(2), 10 * np.ones(2), maxiter=1)
    assert list(tr.batched_lbfgsb.last_converged) == [False, True]           # problem 1 starts at its optimum
    tr.batched_lbfgsb(fun, x0, -10 * np.ones(2), 10 * np.ones(2), maxiter=200)
    assert list(tr.batched_lbfgsb.last_converged) == [True, True]


# ---- world_size-2 gloo: the components of a fit are sharded over the ranks and gathered ---------------------------------
def _fit_worker(rank, world, port, q):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        z = np.load(GOLD)
        kind, hc, hw, d = (int(v) for v in z["c0_meta"])
        X, Y, al = z["c0_X"], z["c0_Y"], float(z["c0_alpha"])        # 3 components over 2 ranks: 2 + 1
        spec = tr.KernelSpec(kind, d, bool(hc), bool(hw), z["c0_theta0"], z["c0_bounds"])
        seen = []

        def obj(th, yi):
            seen.extend(np.unique(yi).tolist())
            return gp_lml.log_marginal_likelihood_batch(X, Y, th, kind, hc, hw, al, yi)

        fit = tr.fit_gps(X, Y, spec, al, n_restarts=3, seed=4, objective=obj)
        q.put((rank, fit.theta, fit.lml, sorted(set(seen)), fit.n_evaluations))
    finally:
        dist.destroy_process_group()


def test_fit_is_sharded_by_component_gloo_world2():
    import socket

    import torch.multiprocessing as mp

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_fit_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    res = sorted([q.get(timeout=300) for _ in procs], key=lambda t: t[0])
    [p.join(60) for p in procs]
    z = np.load(GOLD)
    kind, hc, hw, d = (int(v) for v in z["c0_meta"])
    X, Y, al = z["c0_X"], z["c0_Y"], float(z["c0_alpha"])
    spec = tr.KernelSpec(kind, d, bool(hc), bool(hw), z["c0_theta0"], z["c0_bounds"])
    single = tr.fit_gps(X, Y, spec, al, n_restarts=3, seed=4,
                        objective=lambda th, yi: gp_lml.log_marginal_likelihood_batch(X, Y, th, kind, hc, hw, al, yi))
    for rank, theta, lml, seen, nev in res:
        assert np.array_equal(theta, single.theta) and np.array_equal(lml, single.lml)     # same fit on every rank
        assert seen == ([0, 1] if rank == 0 else [2])                                      # each rank fitted only its slice
        assert nev == single.n_evaluations
