"""Emulator training on the device: the GP hyper-parameter fits of ``emulation.fit_emulation_group`` (reference
emulation.py:128-172), all principal components and all optimiser restarts advanced together.

The reference calls ``GaussianProcessRegressor(kernel, alpha, n_restarts_optimizer=50).fit(design, y)`` once per
principal component; scikit-learn then runs L-BFGS-B on ``-log_marginal_likelihood(theta)`` from the kernel's initial
theta and from ``n_restarts`` log-uniform draws, one after the other (_gpr.py:296-366), each objective evaluation an
N x N Cholesky on one core.  Here

* :class:`DeviceGPTrainer` evaluates the objective and its gradient for any number of hyper-parameter vectors in one
  launch of ``gp_lml_kernel`` (csrc/gp_train.cuh) through ``bi_trainer_*`` (include/bi_b200.h);
* :func:`batched_lbfgsb` runs the P x (1 + n_restarts) box-constrained quasi-Newton optimisations in lock-step, one
  batched objective call per line-search trial;
* :func:`fit_gps` returns, per principal component, the best hyper-parameters with ``L_`` and ``alpha_`` -- what
  ``predict`` needs, so a trained group goes straight into an :class:`~.image.GroupImage` without scikit-learn.

There is no CPU fallback: without the CUDA library :class:`DeviceGPTrainer` raises ``BackendUnavailable``.
The optimiser does not reproduce scipy's L-BFGS-B iterates (line searches differ and the reference's restarts come
from an unseeded generator); what is checked is the objective itself (<= 1e-9 against scikit-learn's values) and that
the fitted log-marginal-likelihood is not below scikit-learn's.
"""
from __future__ import annotations

import logging

import ctypes as C
from dataclasses import dataclass
from typing import Any, Callable, Mapping

import numpy as np

from . import _lib
from .image import KIND_MATERN12, KIND_MATERN32, KIND_MATERN52, KIND_RBF, GroupImage

logger = logging.getLogger(__name__)


@dataclass
class KernelSpec:
    """``Matern | RBF [+ ConstantKernel] [+ WhiteKernel]`` with scikit-learn's log-space theta and bounds."""
    kind: int
    d: int
    has_const: bool
    has_white: bool
    theta0: np.ndarray       # (n_hyper,)  kernel.theta of the untrained kernel
    bounds: np.ndarray       # (n_hyper, 2) kernel.bounds (log space)

    @property
    def n_hyper(self) -> int:
        return self.d + int(self.has_const) + int(self.has_white)


def kernel_spec(active_kernels: Mapping[str, Mapping[str, Any]], lo, hi) -> KernelSpec:
    """The kernel the reference builds from ``config.active_kernels`` and the prior box (emulation.py:128-162):
    length_scale = max - min with bounds ``outer(length_scale, length_scale_bounds_factor)``, an optional
    ``constant`` term and an optional ``noise`` (WhiteKernel) term, summed in the order of the mapping."""
    lo, hi = np.asarray(lo, float), np.asarray(hi, float)
    ell = hi - lo
    kind = None
    th, bd = [], []
    has_const = has_white = False
    for name, args in active_kernels.items():
        if name in ("matern", "rbf"):
            kind = KIND_RBF if name == "rbf" else {0.5: KIND_MATERN12, 1.5: KIND_MATERN32, 2.5: KIND_MATERN52}[float(args["nu"])]
            fac = tuple(args["length_scale_bounds_factor"])
            th = [np.log(ell)]
            bd = [np.log(np.outer(ell, fac))]
        elif name == "constant":
            has_const = True
            th.append(np.log([float(args["constant_value"])]))
            bd.append(np.log(np.asarray(args["constant_value_bounds"], float))[None, :])
        elif name == "noise":
            has_white = True
            th.append(np.log([float(args["args"]["noise_level"])]))
            bd.append(np.log(np.asarray(args["args"]["noise_level_bounds"], float))[None, :])
    if kind is None:
        raise ValueError("active_kernels needs a 'matern' or 'rbf' entry")
    return KernelSpec(kind, ell.size, has_const, has_white, np.concatenate(th), np.concatenate(bd, axis=0))


class DeviceGPTrainer:
    """Device-resident design + targets; batched ``log_marginal_likelihood(theta, eval_gradient=True)``."""

    def __init__(self, X, Y, spec: KernelSpec, alpha: float, device: int = 0):
        self._lib = _lib.load()
        self.X = np.ascontiguousarray(X, np.float64)
        self.Y = np.ascontiguousarray(np.atleast_2d(Y), np.float64)          # (n_y, N)
        self.N, self.d = self.X.shape
        if self.Y.shape[1] != self.N or spec.d != self.d:
            raise ValueError("X (N, d), Y (n_y, N) and the kernel's d must agree")
        self.spec, self.alpha = spec, float(alpha)
        self._h = C.c_void_p()
        rc = self._lib.bi_trainer_create(self.X.ctypes.data, self.N, self.d, self.Y.ctypes.data, self.Y.shape[0], spec.kind,
                                         int(spec.has_const), int(spec.has_white), self.alpha, device, C.byref(self._h))
        if rc != 0:
            self._h = None
            raise RuntimeError(f"bi_trainer_create failed with code {rc}")

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.bi_trainer_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    @property
    def launches(self) -> int:
        return int(self._lib.bi_trainer_launches(self._h))

    def _args(self, theta, y_index):
        theta = np.ascontiguousarray(np.atleast_2d(theta), np.float64)
        if theta.shape[1] != self.spec.n_hyper:
            raise ValueError(f"theta must have {self.spec.n_hyper} columns")
        B = theta.shape[0]
        idx = None if y_index is None else np.ascontiguousarray(y_index, np.int32)
        if idx is not None and idx.shape != (B,):
            raise ValueError("y_index must have one entry per theta row")
        return theta, idx, B

    def _check(self, rc: int, what: str) -> None:
        if rc != 0:
            raise RuntimeError(f"{what}: {self._lib.bi_trainer_last_error(self._h).decode()} (code {rc})")

    def log_marginal_likelihood(self, theta, y_index=None):
        """(lml (B,), grad (B, n_hyper)) -- scikit-learn _gpr.py:541-639 for every row of theta against Y[y_index]."""
        theta, idx, B = self._args(theta, y_index)
        lml = np.empty(B)
        grad = np.empty_like(theta)
        self._check(self._lib.bi_trainer_lml_host(self._h, theta.ctypes.data, None if idx is None else idx.ctypes.data, B,
                                                  lml.ctypes.data, grad.ctypes.data), "bi_trainer_lml_host")
        return lml, grad

    def log_marginal_likelihood_tensor(self, theta, y_index=None):
        """Same with torch CUDA tensors in and out, enqueued on torch's current stream (no synchronisation)."""
        import torch
        B = theta.shape[0]
        lml = torch.empty(B, dtype=torch.float64, device=theta.device)
        grad = torch.empty_like(theta)
        self._check(self._lib.bi_trainer_lml(self._h, theta.data_ptr(), None if y_index is None else y_index.data_ptr(), B,
                                             lml.data_ptr(), grad.data_ptr(), torch.cuda.current_stream().cuda_stream),
                    "bi_trainer_lml")
        return lml, grad

    def factor(self, theta, y_index=None):
        """(lml (B,), L (B, N, N), alpha (B, N)): the state ``fit`` keeps for ``predict`` (_gpr.py:335-366)."""
        theta, idx, B = self._args(theta, y_index)
        lml = np.empty(B)
        L = np.empty((B, self.N, self.N))
        al = np.empty((B, self.N))
        self._check(self._lib.bi_trainer_factor_host(self._h, theta.ctypes.data, None if idx is None else idx.ctypes.data, B,
                                                     lml.ctypes.data, L.ctypes.data, al.ctypes.data), "bi_trainer_factor_host")
        return lml, L, al


def batched_lbfgsb(fun: Callable[[np.ndarray, np.ndarray], tuple[np.ndarray, np.ndarray]], x0, lo, hi, *, m: int = 10,
                   maxiter: int = 200, pgtol: float = 1e-5, ftol: float = 2.220446049250313e-09, maxls: int = 20):
    """Minimise B independent box-constrained problems in lock-step with a projected L-BFGS method.

    ``fun(X (b, n), idx (b,)) -> (f (b,), g (b, n))`` evaluates the problems ``idx`` at the rows of ``X``; every call
    is one batch.  Stopping rules are scipy's L-BFGS-B defaults as scikit-learn uses them (``pgtol`` on the projected
    gradient, ``ftol = factr * eps`` on the relative decrease).  Returns (x, f, n_iterations, n_batches); after the call
    ``batched_lbfgsb.last_converged`` (B,) tells which problems met a stopping rule (False: still moving at ``maxiter`` --
    scipy would have carried on up to its 15000 iterations -- or no acceptable step within ``maxls`` trials)."""
    x = np.clip(np.array(x0, float), lo, hi)
    B, n = x.shape
    lo = np.broadcast_to(np.asarray(lo, float), (B, n))
    hi = np.broadcast_to(np.asarray(hi, float), (B, n))
    allidx = np.arange(B)
    f, g = fun(x, allidx)
    nbatch = 1
    S = np.zeros((B, m, n))
    Yh = np.zeros((B, m, n))
    rho = np.zeros((B, m))
    active = np.isfinite(f)
    nit = np.zeros(B, int)
    gave_up = np.zeros(B, bool)
    for it in range(maxiter):
        pg = np.where(((x <= lo) & (g > 0)) | ((x >= hi) & (g < 0)), 0.0, g)
        active &= np.max(np.abs(pg), axis=1) > pgtol
        if not active.any():
            break
        # two-loop recursion, newest pair at slot 0; empty slots have rho = 0 and drop out
        q = pg.copy()
        a = np.zeros((B, m))
        for i in range(m):
            a[:, i] = rho[:, i] * np.einsum("bn,bn->b", S[:, i], q)
            q -= a[:, i, None] * Yh[:, i]
        yy = np.einsum("bn,bn->b", Yh[:, 0], Yh[:, 0])
        gamma = np.where(rho[:, 0] > 0, 1.0 / np.maximum(rho[:, 0] * yy, 1e-300), 1.0)
        r = gamma[:, None] * q
        for i in range(m - 1, -1, -1):
            bcoef = rho[:, i] * np.einsum("bn,bn->b", Yh[:, i], r)
            r += S[:, i] * (a[:, i] - bcoef)[:, None]
        dvec = np.where(pg != 0.0, -r, 0.0)
        bad = np.einsum("bn,bn->b", dvec, g) >= 0
        dvec[bad] = -pg[bad]
        t = np.ones(B)
        first = rho[:, 0] == 0
        t[first] = np.minimum(1.0, 1.0 / np.maximum(np.linalg.norm(pg[first], axis=1), 1e-300))
        searching = active.copy()
        xn, fn, gn = x.copy(), f.copy(), g.copy()
        for _ in range(maxls):
            sel = np.flatnonzero(searching)
            if sel.size == 0:
                break
            xt = np.clip(x[sel] + t[sel, None] * dvec[sel], lo[sel], hi[sel])
            ft, gt = fun(xt, sel)
            nbatch += 1
            ok = np.isfinite(ft) & (ft <= f[sel] + 1e-4 * np.einsum("bn,bn->b", g[sel], xt - x[sel]))
            acc = sel[ok]
            xn[acc], fn[acc], gn[acc] = xt[ok], ft[ok], gt[ok]
            searching[acc] = False
            t[sel[~ok]] *= 0.5
        stalled = searching                      # no acceptable step within maxls trials: stop that problem
        moved = active & ~stalled
        s = xn - x
        y = gn - g
        sy = np.einsum("bn,bn->b", s, y)
        upd = moved & (sy > 1e-10 * np.linalg.norm(s, axis=1) * np.linalg.norm(y, axis=1))
        ui = np.flatnonzero(upd)
        S[ui] = np.roll(S[ui], 1, axis=1)
        Yh[ui] = np.roll(Yh[ui], 1, axis=1)
        rho[ui] = np.roll(rho[ui], 1, axis=1)
        S[ui, 0], Yh[ui, 0], rho[ui, 0] = s[ui], y[ui], 1.0 / sy[ui]
        small = (f - fn) <= ftol * np.maximum(np.maximum(np.abs(f), np.abs(fn)), 1.0)
        nit[moved] += 1
        x[moved], f[moved], g[moved] = xn[moved], fn[moved], gn[moved]
        gave_up |= active & stalled
        active &= ~stalled & ~small
    batched_lbfgsb.last_converged = ~active & ~gave_up & np.isfinite(f)
    return x, f, nit, nbatch


@dataclass
class FittedGPs:
    """Per principal component: what ``GaussianProcessRegressor.fit`` leaves behind."""
    spec: KernelSpec
    X: np.ndarray            # (N, d)
    theta: np.ndarray        # (P, n_hyper)  kernel_.theta
    lml: np.ndarray          # (P,)          log_marginal_likelihood_value_
    L: np.ndarray            # (P, N, N)     L_
    alpha: np.ndarray        # (P, N)        alpha_
    n_batches: int = 0       # batched objective launches the optimiser needed
    n_evaluations: int = 0   # hyper-parameter vectors evaluated in total
    n_unconverged: int = 0   # winning starts that had not met a stopping rule when the optimiser stopped (maxiter / line search)

    @property
    def ell(self) -> np.ndarray:
        return np.exp(self.theta[:, :self.spec.d])

    @property
    def const(self) -> np.ndarray:
        return np.exp(self.theta[:, self.spec.d]) if self.spec.has_const else np.zeros(len(self.theta))

    @property
    def noise(self) -> np.ndarray:
        k = self.spec.d + int(self.spec.has_const)
        return np.exp(self.theta[:, k]) if self.spec.has_white else np.zeros(len(self.theta))

    def group_image(self, name: str, S, S_rest, ev_rest, scale, mean) -> GroupImage:
        """Join the fitted GPs with the group's inverse-PCA constants into the device image's group record."""
        P = len(self.theta)
        return GroupImage(name=name, X=self.X, kind=np.full(P, self.spec.kind, np.int32), ell=self.ell, const=self.const,
                          noise=self.noise, alpha=self.alpha, L=self.L, S=np.asarray(S, float), S_rest=np.asarray(S_rest, float),
                          ev_rest=np.asarray(ev_rest, float), scale=np.asarray(scale, float), mean=np.asarray(mean, float))


def fit_gps(design, Y_pca_T, spec: KernelSpec, alpha: float, n_restarts: int = 0, *, seed: int | None = None,
            device: int = 0, maxiter: int = 200, objective=None) -> FittedGPs:
    """Fit one GP per row of ``Y_pca_T`` (P, N) -- the reference's ``[GPR(...).fit(design, y) for y in Y_pca_truncated.T]``
    (emulation.py:169-172) -- with all P x (1 + n_restarts) optimisations advanced together on the device.

    Start 0 of every component is the kernel's initial theta, the others are log-uniform draws from the bounds, as in
    scikit-learn (_gpr.py:319-333).  ``objective`` replaces the device objective in host-logic tests only."""
    X = np.ascontiguousarray(design, np.float64)
    Y = np.ascontiguousarray(np.atleast_2d(Y_pca_T), np.float64)
    P = Y.shape[0]
    R = 1 + int(n_restarts)
    rng = np.random.RandomState(seed)
    lo, hi = spec.bounds[:, 0], spec.bounds[:, 1]
    starts = np.empty((P, R, spec.n_hyper))
    starts[:, 0] = spec.theta0
    if R > 1:
        starts[:, 1:] = rng.uniform(lo, hi, (P, R - 1, spec.n_hyper))
    # one process per GPU: the components are independent, so rank r fits the slice P_r and the results are gathered
    # (no exchange on the data path); every rank draws the same starts, so the fit does not depend on the world size
    rank, world = _dist_rank_world()
    per = -(-P // world)
    p0, p1 = min(P, rank * per), min(P, (rank + 1) * per)
    Pl = p1 - p0
    yidx = np.repeat(np.arange(p0, p1, dtype=np.int32), R)
    trainer = None
    if objective is None:
        trainer = DeviceGPTrainer(X, Y, spec, alpha, device)
        objective = trainer.log_marginal_likelihood
    nev = 0

    def neg(theta, idx):
        nonlocal nev
        nev += len(idx)
        lml, grad = objective(theta, yidx[idx])
        return -lml, -grad

    try:
        if Pl > 0:
            xs, fs, _, nbatch = batched_lbfgsb(neg, starts[p0:p1].reshape(Pl * R, -1), lo, hi, maxiter=maxiter)
            conv = batched_lbfgsb.last_converged.reshape(Pl, R)
            fs = np.where(np.isfinite(fs), fs, np.inf).reshape(Pl, R)
            best = np.argmin(fs, axis=1)
            n_unconv = int(np.count_nonzero(~conv[np.arange(Pl), best]))
            if n_unconv:
                logger.warning(f"fit_gps: the best start of {n_unconv} of {Pl} components had not converged after "
                               f"{maxiter} iterations (scipy's L-BFGS-B would continue up to 15000); raise maxiter")
            theta = xs.reshape(Pl, R, -1)[np.arange(Pl), best]
        else:
            theta, fs, best, nbatch, n_unconv = np.empty((0, spec.n_hyper)), np.empty((0, R)), np.empty(0, int), 0, 0
        if trainer is not None and Pl > 0:
            lml, L, al = trainer.factor(theta, np.arange(p0, p1, dtype=np.int32))
        else:                                      # host-logic tests: the caller's objective has no factor output
            lml, L, al = -fs[np.arange(Pl), best], np.empty((Pl, 0, 0)), np.empty((Pl, 0))
    finally:
        if trainer is not None:
            trainer.close()
    if world > 1:
        import torch.distributed as dist
        parts: list[Any] = [None] * world
        dist.all_gather_object(parts, (theta, lml, L, al, nbatch, nev, n_unconv))
        theta, lml, L, al = (np.concatenate([p[i] for p in parts if len(p[0])], axis=0) for i in range(4))
        nbatch, nev, n_unconv = max(p[4] for p in parts), sum(p[5] for p in parts), sum(p[6] for p in parts)
    return FittedGPs(spec, X, theta, lml, L, al, n_batches=nbatch, n_evaluations=nev, n_unconverged=n_unconv)


def _dist_rank_world() -> tuple[int, int]:
    try:
        import torch.distributed as dist
    except ImportError:          # torch is plumbing only; a single-process fit does not need it
        return 0, 1
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def pca_transform(Y, n_pc: int, max_n_components: int | None = None, sign_rule: str = "u"):
    """StandardScaler + PCA(svd_solver='full', whiten=False) as emulation.py:107-122 applies them, by one SVD.
    Returns (Y_pca_truncated (N, n_pc), components (Q, F), explained_variance (Q,), scale (F,), mean (F,)).

    ``sign_rule``: how the sign of each component is fixed.  'u' (default) is scikit-learn 1.3.0's -- the version the
    reference pins (pdm.lock:1998) -- ``svd_flip(U, Vt)`` with ``u_based_decision=True``: the entry of largest magnitude in
    every column of U is made positive.  'v' is scikit-learn >= 1.5's rule (largest entry of every row of Vt).  The final
    mean and covariance do not depend on it; per-PC quantities (targets, components) compared against pickles do."""
    Y = np.asarray(Y, float)
    mean = Y.mean(axis=0)
    scale = Y.std(axis=0)
    scale = np.where(scale == 0.0, 1.0, scale)
    Z = (Y - mean) / scale
    Zc = Z - Z.mean(axis=0)
    U, sv, Vt = np.linalg.svd(Zc, full_matrices=False)
    if sign_rule == "u":
        sign = np.sign(U[np.argmax(np.abs(U), axis=0), np.arange(U.shape[1])])     # sklearn 1.3.0 svd_flip, u-based
    elif sign_rule == "v":
        sign = np.sign(Vt[np.arange(Vt.shape[0]), np.argmax(np.abs(Vt), axis=1)])  # sklearn >= 1.5 svd_flip, v-based
    else:
        raise ValueError("sign_rule must be 'u' or 'v'")
    sign[sign == 0] = 1.0
    U, Vt = U * sign, Vt * sign[:, None]
    Q = len(sv) if max_n_components is None else min(len(sv), int(max_n_components))
    ev = sv[:Q] ** 2 / (Y.shape[0] - 1)
    return (U[:, :n_pc] * sv[:n_pc]), Vt[:Q], ev, scale, mean


def fit_emulation_group(name: str, design, Y, n_pc: int, spec: KernelSpec, alpha: float, n_restarts: int, *,
                        max_n_components: int | None = None, seed: int | None = None, device: int = 0) -> tuple[GroupImage, FittedGPs]:
    """``emulation.fit_emulation_group`` (emulation.py:38-192) on arrays: scale + PCA on the host (one SVD), the P GP
    fits on the device, result as the group record of the device image."""
    Ypca, comps, ev, scale, mean = pca_transform(Y, n_pc, max_n_components)
    gps = fit_gps(design, Ypca.T, spec, alpha, n_restarts, seed=seed, device=device)
    return gps.group_image(name, comps[:n_pc], comps[n_pc:], ev[n_pc:], scale, mean), gps
