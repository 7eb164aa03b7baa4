"""The emulator image: every constant the log-posterior path needs, as flat float64/int arrays.

This is the host-side source of what ``bi_create`` uploads to HBM once.  It is built either from
the reference's own fitted objects -- the dict ``{'PCA': {'pca', 'scaler'}, 'emulators': [GaussianProcessRegressor]}``
per emulation group written by ``emulation.write_emulators`` (emulation.py:181-211) -- or from a flat ``.npz`` file
(`save` / `load`), so that a GPU job never needs to unpickle scikit-learn objects.

Kernel tree accepted (what ``emulation.fit_emulator_group`` can build, emulation.py:132-162):
``Matern(nu in {0.5, 1.5, 2.5, inf}) | RBF``, optionally ``+ ConstantKernel``, optionally ``+ WhiteKernel``, summed in any
order.  Anything else raises -- there is no fallback.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Mapping, Sequence

import numpy as np

from .observable_map import ObservableBlockMap

# kernel kind codes shared with csrc/ (bi_b200.h: BI_KERNEL_*)
KIND_MATERN12, KIND_MATERN32, KIND_MATERN52, KIND_RBF = 0, 1, 2, 3


@dataclass
class GroupImage:
    name: str
    X: np.ndarray        # (N, d)   shared design matrix X_train_
    kind: np.ndarray     # (P,)     int32 kernel kind per principal component
    ell: np.ndarray      # (P, d)   anisotropic length scales
    const: np.ndarray    # (P,)     ConstantKernel value (0 when absent)
    noise: np.ndarray    # (P,)     WhiteKernel level (0 when absent)
    alpha: np.ndarray    # (P, N)   K^-1 y  (gpr.alpha_)
    L: np.ndarray        # (P, N, N) lower Cholesky factor of K + alpha I (gpr.L_)
    S: np.ndarray        # (P, F_g) retained principal axes  pca.components_[:P]
    S_rest: np.ndarray   # (Q-P, F_g) truncated principal axes
    ev_rest: np.ndarray  # (Q-P,)   their explained variances
    scale: np.ndarray    # (F_g,)   scaler.scale_
    mean: np.ndarray     # (F_g,)   scaler.mean_

    @property
    def n_design(self) -> int:
        return int(self.X.shape[0])

    @property
    def n_pc(self) -> int:
        return int(self.S.shape[0])

    @property
    def n_features(self) -> int:
        return int(self.S.shape[1])

    def cov_unexplained(self) -> np.ndarray:
        """S_u diag(explained_variance_[P:]) S_u^T  (emulation.py:227-251)."""
        return self.S_rest.T @ (self.ev_rest[:, None] * self.S_rest)


@dataclass
class EmulatorImage:
    d: int
    min: np.ndarray                 # (d,)
    max: np.ndarray                 # (d,)
    groups: list[GroupImage]
    blocks: np.ndarray              # (B, 4) int64 [group, out_start, in_start, len] sorted by out_start
    n_features: int                 # F, width of the merged matrices
    y: np.ndarray | None = None     # (F,) experimental central values
    y_err: np.ndarray | None = None  # (F,) experimental uncertainties -> Sigma_exp = diag(y_err^2)
    sigma_exp: np.ndarray | None = None  # (F, F) dense experimental covariance (overrides y_err when given)
    meta: dict[str, Any] = field(default_factory=dict)

    # ---- flat dict / npz round trip (the on-disk format either side of the path) -----------------
    def to_flat(self) -> dict[str, np.ndarray]:
        out: dict[str, np.ndarray] = {
            "d": np.int64(self.d), "min": self.min, "max": self.max, "blocks": self.blocks,
            "n_features": np.int64(self.n_features), "n_groups": np.int64(len(self.groups)),
            "group_names": np.array([g.name for g in self.groups]),
        }
        for nm in ("y", "y_err", "sigma_exp"):
            if getattr(self, nm) is not None:
                out[nm] = getattr(self, nm)
        for i, g in enumerate(self.groups):
            for nm in ("X", "kind", "ell", "const", "noise", "alpha", "L", "S", "S_rest", "ev_rest", "scale", "mean"):
                out[f"g{i}_{nm}"] = getattr(g, nm)
        return out

    @classmethod
    def from_flat(cls, z: Mapping[str, np.ndarray]) -> "EmulatorImage":
        names = [str(s) for s in np.asarray(z["group_names"]).tolist()]
        groups = []
        for i, nm in enumerate(names):
            kw = {k: np.asarray(z[f"g{i}_{k}"]) for k in
                  ("X", "kind", "ell", "const", "noise", "alpha", "L", "S", "S_rest", "ev_rest", "scale", "mean")}
            kw["kind"] = kw["kind"].astype(np.int32)
            groups.append(GroupImage(name=nm, **kw))
        opt = {k: (np.asarray(z[k], dtype=np.float64) if k in z else None) for k in ("y", "y_err", "sigma_exp")}
        return cls(d=int(z["d"]), min=np.asarray(z["min"], dtype=np.float64), max=np.asarray(z["max"], dtype=np.float64),
                   groups=groups, blocks=np.asarray(z["blocks"], dtype=np.int64).reshape(-1, 4),
                   n_features=int(z["n_features"]), **opt)

    def save(self, path: str) -> None:
        np.savez_compressed(path, **self.to_flat())

    @classmethod
    def load(cls, path: str) -> "EmulatorImage":
        with np.load(path, allow_pickle=False) as z:
            return cls.from_flat({k: z[k] for k in z.files})

    @property
    def group_names(self) -> list[str]:
        return [g.name for g in self.groups]

    def experimental_cov(self) -> np.ndarray:
        """Dense Sigma_exp (reference: ``np.diag(data_y_err**2)``, log_posterior.py:94)."""
        if self.sigma_exp is not None:
            return np.asarray(self.sigma_exp, dtype=np.float64)
        if self.y_err is None:
            raise ValueError("image carries no experimental uncertainties")
        return np.diag(np.asarray(self.y_err, dtype=np.float64) ** 2)


# ---------------------------------------------------------------------------------------------------------------
def _flatten_sum(kernel) -> list[Any]:
    """Leaves of a tree of sklearn ``Sum`` kernels (kernels.py:838-873)."""
    if type(kernel).__name__ == "Sum":
        return _flatten_sum(kernel.k1) + _flatten_sum(kernel.k2)
    return [kernel]


def kernel_constants(kernel, d: int) -> tuple[int, np.ndarray, float, float]:
    """(kind, ell[d], constant_value, noise_level) of one fitted ``gpr.kernel_``.

    Matern/RBF: kernels.py:1685-1712 / :1530-1545; ConstantKernel :1244-1280; WhiteKernel :1374-1441.
    """
    kind = None
    ell = None
    const = 0.0
    noise = 0.0
    for leaf in _flatten_sum(kernel):
        nm = type(leaf).__name__
        if nm in ("Matern", "RBF"):
            if kind is not None:
                raise ValueError(f"more than one stationary kernel in {kernel}")
            ell = np.broadcast_to(np.asarray(leaf.length_scale, dtype=np.float64), (d,)).copy()
            if nm == "RBF":
                kind = KIND_RBF
            else:
                nu = float(leaf.nu)
                table = {0.5: KIND_MATERN12, 1.5: KIND_MATERN32, 2.5: KIND_MATERN52, float("inf"): KIND_RBF}
                if nu not in table:
                    raise ValueError(f"Matern nu={nu} is not supported on the device path (needs Bessel K_nu)")
                kind = table[nu]
        elif nm == "ConstantKernel":
            const += float(leaf.constant_value)
        elif nm == "WhiteKernel":
            noise += float(leaf.noise_level)
        else:
            raise ValueError(f"unsupported kernel component {nm} in {kernel}")
    if kind is None:
        raise ValueError(f"no Matern/RBF component in {kernel}")
    return kind, ell, const, noise


def factor_cov_unexplained(U: np.ndarray, n_features: int) -> tuple[np.ndarray, np.ndarray]:
    """(S_rest, ev_rest) with ``S_rest.T @ diag(ev_rest) @ S_rest == U``: how a caller-supplied truncated-PC covariance
    (``emulator_cov_unexplained[group]``, emulation.py:445-448, :487-488) enters the device image, which carries that
    term in factored form.  Symmetric eigendecomposition; eigenvalues of either sign are kept as they are."""
    U = np.asarray(U, dtype=np.float64)
    if U.shape != (n_features, n_features):
        raise ValueError(f"emulator_cov_unexplained has shape {U.shape}, expected ({n_features}, {n_features})")
    scale = float(np.max(np.abs(U))) if U.size else 0.0
    if not np.allclose(U, U.T, rtol=0.0, atol=1e-12 * max(scale, 1e-300)):
        raise ValueError("emulator_cov_unexplained must be symmetric (it is a covariance)")
    w, V = np.linalg.eigh(0.5 * (U + U.T))
    keep = np.abs(w) > 0.0
    return np.ascontiguousarray(V[:, keep].T), np.ascontiguousarray(w[keep])


def pack_group(name: str, result: Mapping[str, Any], n_pc: int, cov_unexplained: np.ndarray | None = None) -> GroupImage:
    """Flatten one emulation group's fitted objects (emulation.py:181-190).  ``cov_unexplained``: use this (F_g, F_g)
    matrix as the truncated-PC covariance instead of the one the PCA object implies (emulation.py:487-488)."""
    emulators: Sequence[Any] = result["emulators"]
    if len(emulators) < n_pc:
        raise ValueError(f"group {name}: {len(emulators)} emulators for n_pc={n_pc}")
    emulators = list(emulators)[:n_pc] if len(emulators) > n_pc else list(emulators)
    pca, scaler = result["PCA"]["pca"], result["PCA"]["scaler"]
    X = np.ascontiguousarray(emulators[0].X_train_, dtype=np.float64)
    N, d = X.shape
    P = len(emulators)
    kind = np.zeros(P, np.int32)
    ell = np.zeros((P, d))
    const = np.zeros(P)
    noise = np.zeros(P)
    alpha = np.zeros((P, N))
    L = np.zeros((P, N, N))
    for p, gpr in enumerate(emulators):
        if not np.array_equal(np.asarray(gpr.X_train_), X):
            raise ValueError(f"group {name}: emulator {p} was trained on a different design")
        if float(np.ravel(gpr._y_train_std)[0]) != 1.0 or float(np.ravel(gpr._y_train_mean)[0]) != 0.0:
            raise ValueError("normalize_y=True emulators are not produced by the reference (emulation.py:169-172)")
        kind[p], ell[p], const[p], noise[p] = kernel_constants(gpr.kernel_, d)
        alpha[p] = np.ravel(gpr.alpha_)
        L[p] = np.tril(gpr.L_)
    comp = np.asarray(pca.components_, dtype=np.float64)
    ev = np.asarray(pca.explained_variance_, dtype=np.float64)
    S_rest, ev_rest = np.ascontiguousarray(comp[P:]), np.ascontiguousarray(ev[P:])
    if cov_unexplained is not None:
        S_rest, ev_rest = factor_cov_unexplained(cov_unexplained, comp.shape[1])
    return GroupImage(name=name, X=X, kind=kind, ell=ell, const=const, noise=noise, alpha=alpha, L=L,
                      S=np.ascontiguousarray(comp[:P]), S_rest=S_rest, ev_rest=ev_rest,
                      scale=np.asarray(scaler.scale_, dtype=np.float64), mean=np.asarray(scaler.mean_, dtype=np.float64))


def pack(emulation_config, emulation_results: Mapping[str, Mapping[str, Any]], *, min, max,
         experimental_results: Mapping[str, np.ndarray] | None = None,
         sigma_exp: np.ndarray | None = None,
         emulator_cov_unexplained: Mapping[str, np.ndarray] | None = None) -> EmulatorImage:
    """Build the image from what ``mcmc.run_mcmc`` hands to ``initialize_pool_variables`` (mcmc.py:48-78).

    ``emulation_config`` may be the reference's ``EmulationConfig`` or anything exposing
    ``emulation_groups_config[g].n_pc`` and ``sort_observables_in_matrix``.  A non-empty ``emulator_cov_unexplained``
    ({group: (F_g, F_g)}) replaces the truncated-PC covariance of every group, as in the reference (emulation.py:445-448:
    a non-empty dictionary must hold every group).
    """
    names = list(emulation_config.emulation_groups_config.keys())
    ucov = emulator_cov_unexplained if emulator_cov_unexplained else {}
    if ucov and any(g not in ucov for g in names):
        raise KeyError(f"emulator_cov_unexplained lacks group(s) {[g for g in names if g not in ucov]}")
    groups = [pack_group(g, emulation_results[g], int(emulation_config.emulation_groups_config[g].n_pc),
                         cov_unexplained=ucov.get(g)) for g in names]
    bmap = ObservableBlockMap.from_reference(emulation_config.sort_observables_in_matrix)
    blocks = bmap.block_table(names)
    lo = np.asarray(min, dtype=np.float64).ravel()
    hi = np.asarray(max, dtype=np.float64).ravel()
    d = groups[0].X.shape[1]
    if lo.shape != (d,) or hi.shape != (d,):
        raise ValueError(f"prior bounds have shape {lo.shape}/{hi.shape}, design has d={d}")
    img = EmulatorImage(d=d, min=lo, max=hi, groups=groups, blocks=blocks, n_features=bmap.n_features)
    if experimental_results is not None:
        img.y = np.asarray(experimental_results["y"], dtype=np.float64).ravel()
        img.y_err = np.asarray(experimental_results["y_err"], dtype=np.float64).ravel()
    if sigma_exp is not None:
        img.sigma_exp = np.ascontiguousarray(sigma_exp, dtype=np.float64)
    validate(img)
    return img


def validate(img: EmulatorImage) -> None:
    F = img.n_features
    cover = np.zeros(F, dtype=np.int64)
    for g, so, si, ln in img.blocks:
        if not (0 <= g < len(img.groups)) or ln <= 0:
            raise ValueError(f"bad block row {(g, so, si, ln)}")
        if si + ln > img.groups[g].n_features or so + ln > F:
            raise ValueError(f"block {(g, so, si, ln)} exceeds its matrix")
        cover[so:so + ln] += 1
    if not np.all(cover == 1):
        raise ValueError("observable blocks must tile the merged feature axis exactly once")
    for nm in ("y", "y_err"):
        v = getattr(img, nm)
        if v is not None and v.shape != (F,):
            raise ValueError(f"{nm} has shape {v.shape}, expected ({F},)")
    if img.sigma_exp is not None and img.sigma_exp.shape != (F, F):
        raise ValueError(f"sigma_exp has shape {img.sigma_exp.shape}, expected ({F},{F})")
    for g in img.groups:
        if g.X.shape[1] != img.d:
            raise ValueError("all groups must share the parameter dimension")
