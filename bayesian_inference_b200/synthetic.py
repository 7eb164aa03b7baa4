"""Synthetic analyses of the shapes BASELINE.json names (SURVEY §8 d).

There is no network and no experimental data in this image, so every benchmark and parity test runs on a seeded
synthetic analysis: a random design, smooth latent observables with a decaying PCA spectrum, a genuine
``StandardScaler`` + ``PCA(svd_solver='full')`` fit and one genuine ``GaussianProcessRegressor`` per principal
component -- i.e. exactly the object graph the reference pickles in ``emulation.write_emulators``
(emulation.py:181-211) and reads back in ``mcmc.run_mcmc`` (mcmc.py:54-60).  The hot path never sees the
generator, only the fitted objects it produces.

The same generator feeds the GPU path, the CPU oracle and (in the build container) the reference itself.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from types import SimpleNamespace
from typing import Any

import numpy as np

from .observable_map import ObservableBlockMap, ObservableSelection, sort_labels


@dataclass
class GroupSpec:
    name: str
    n_pc: int
    n_obs: int            # number of observables in this group
    n_bins: int           # bins per observable
    kernel: str = "matern"  # 'matern' | 'rbf'
    nu: float = 1.5
    constant: bool = False
    noise: bool = True


@dataclass
class ProblemSpec:
    name: str
    d: int
    n_design: int
    groups: list[GroupSpec]
    n_walkers: int
    seed: int = 0
    optimise: bool = False       # run sklearn's hyper-parameter optimiser (C1) instead of fixed random hyper-parameters
    dense_sigma_exp: bool = False  # Sigma_exp = diag(y_err^2) + 0.3 y_err y_err^T inside each observable block
    oob_fraction: float = 0.0    # fraction of walker rows pushed outside the prior box (correctness runs only)

    @property
    def n_features(self) -> int:
        return sum(g.n_obs * g.n_bins for g in self.groups)

    @property
    def n_pc_total(self) -> int:
        return sum(g.n_pc for g in self.groups)


# ---- the five BASELINE.json configurations, concretely ------------------------------------------
def config(name: str, **over: Any) -> ProblemSpec:
    n = name.upper()
    if n == "C1":      # jet_substructure.yaml-shaped CPU run
        s = ProblemSpec("C1", d=4, n_design=230, groups=[GroupSpec("g0", 30, 20, 10)], n_walkers=100, seed=42)
    elif n == "C2":    # predict-only, 1e5 theta
        s = ProblemSpec("C2", d=6, n_design=230, groups=[GroupSpec("g0", 30, 30, 20)], n_walkers=100_000, seed=2)
    elif n in ("C3", "C3A"):  # 30 observables x 20 bins
        s = ProblemSpec("C3a", d=6, n_design=230, groups=[GroupSpec("g0", 30, 30, 20)], n_walkers=1024, seed=3)
    elif n == "C3B":   # 1 observable x 600 bins: genuinely dense covariance
        s = ProblemSpec("C3b", d=6, n_design=230, groups=[GroupSpec("g0", 30, 1, 600)], n_walkers=1024, seed=3)
    elif n == "C4":    # rehlers.yaml-style: 4 groups, all kernel kinds exercised
        s = ProblemSpec("C4", d=6, n_design=230, n_walkers=8192, seed=4, groups=[
            GroupSpec("high_pt_hadron", 10, 15, 10),
            GroupSpec("low_pt_hadron", 10, 25, 10, constant=True),
            GroupSpec("high_pt_jet", 10, 10, 10, kernel="rbf"),
            GroupSpec("low_pt_jet", 10, 10, 10, nu=2.5),
        ])
    elif n in ("C5", "C5A"):
        s = ProblemSpec("C5a", d=6, n_design=500, groups=[GroupSpec("g0", 50, 100, 20)], n_walkers=65536, seed=5)
    elif n == "C5B":
        s = ProblemSpec("C5b", d=6, n_design=500, groups=[GroupSpec("g0", 50, 1, 2000)], n_walkers=65536, seed=5)
    elif n == "TOY":   # seconds on one core; used by CPU tests and golden fixtures
        s = ProblemSpec("toy", d=3, n_design=40, n_walkers=24, seed=7, groups=[
            GroupSpec("alpha", 4, 3, 5, constant=True),
            GroupSpec("beta", 3, 2, 4, kernel="rbf"),
        ])
    else:
        raise KeyError(f"unknown synthetic configuration {name!r}")
    for k, v in over.items():
        setattr(s, k, v)
    return s


@dataclass
class SyntheticAnalysis:
    """Everything ``mcmc.run_mcmc`` would have loaded from disk, plus walkers."""

    spec: ProblemSpec
    names: list[str]
    min: np.ndarray
    max: np.ndarray
    design: np.ndarray                              # (N, d)
    observables: dict[str, Any]                     # in-memory stand-in for observables.h5
    emulation_results: dict[str, dict[str, Any]]    # {group: {'PCA': {'pca','scaler'}, 'emulators': [GPR]*P}}
    emulation_config: Any                           # duck-typed EmulationConfig (see make_config)
    experimental_results: dict[str, np.ndarray]     # {'y': (F,), 'y_err': (F,)}
    sigma_exp: np.ndarray | None                    # dense (F,F) experimental covariance, or None for diag(y_err^2)
    theta: np.ndarray                               # (W, d) walker positions
    labels_by_group: dict[str, list[str]] = field(default_factory=dict)


_TYPES = ("hadron", "jet", "dijet")


def _labels(spec: ProblemSpec) -> dict[str, list[str]]:
    """Observable labels ``sqrts__system__type__observable__sub__centrality`` (data_IO.py:488-506).

    Observable types are rotated so that the merged (sorted) order interleaves the groups, which is what makes
    the block map of row a7 non-trivial.
    """
    out: dict[str, list[str]] = {}
    for gi, g in enumerate(spec.groups):
        out[g.name] = [
            f"5020__PbPb__{_TYPES[(gi + k) % len(_TYPES)]}__obs{k:03d}_{g.name}__R0.{2 + k % 3}__{10 * (k % 2)}-{10 * (k % 2) + 10}"
            for k in range(g.n_obs)
        ]
    return out


def make_config(spec: ProblemSpec, n_bins: dict[str, int]) -> Any:
    """A stand-in for the reference's ``EmulationConfig`` exposing exactly what the hot path touches:
    ``emulation_groups_config[g].n_pc / .observable_filter`` (emulation.py:434, :315), ``output_dir`` (:304) and
    ``sort_observables_in_matrix`` (:702-709)."""
    groups = {
        g.name: SimpleNamespace(n_pc=g.n_pc, observable_filter=ObservableSelection(include_list=[f"_{g.name}__"]),
                                emulation_outputfile=f"<memory>/emulation_group_{g.name}.pkl")
        for g in spec.groups
    }
    cfg = SimpleNamespace(emulation_groups_config=groups, output_dir="<memory>", analysis_name=spec.name,
                          parameterization="synthetic")
    cfg.sort_observables_in_matrix = ObservableBlockMap.learn(
        n_bins, {name: gc.observable_filter for name, gc in groups.items()})
    return cfg


def generate(spec: ProblemSpec | str, **over: Any) -> SyntheticAnalysis:
    from sklearn.decomposition import PCA
    from sklearn.gaussian_process import GaussianProcessRegressor
    from sklearn.gaussian_process import kernels as sk
    from sklearn.preprocessing import StandardScaler

    if isinstance(spec, str):
        spec = config(spec, **over)
    rng = np.random.default_rng(1000 + spec.seed)
    d, N = spec.d, spec.n_design
    lo, hi = np.zeros(d), np.ones(d)
    design = rng.uniform(lo, hi, size=(N, d))
    labels = _labels(spec)

    prediction: dict[str, dict[str, np.ndarray]] = {}
    results: dict[str, dict[str, Any]] = {}
    for g in spec.groups:
        Fg = g.n_obs * g.n_bins
        W1 = rng.normal(size=(d, 32))
        phase = rng.uniform(0, 2 * np.pi, size=32)
        decay = 1.0 / (1.0 + np.arange(32)) ** 1.0
        W2 = rng.normal(size=(32, Fg)) * decay[:, None]
        Z = np.sin(design @ W1 + phase) @ W2
        Y = 2.0 + 0.3 * Z + 0.02 * rng.normal(size=(N, Fg))
        for k, lab in enumerate(labels[g.name]):
            blk = Y[:, k * g.n_bins:(k + 1) * g.n_bins]
            prediction[lab] = {"y": blk.T.copy(), "y_err": 0.01 * np.abs(blk.T)}

    # the group matrix is assembled in the group's own sorted-label order, as predictions_matrix_from_h5 does
    for g in spec.groups:
        order = sort_labels(labels[g.name])
        Yg = np.concatenate([prediction[lab]["y"].T for lab in order], axis=1)
        scaler = StandardScaler()
        pca = PCA(n_components=None, svd_solver="full", whiten=False)
        Ypca = pca.fit_transform(scaler.fit_transform(Yg))
        emulators = []
        for p in range(g.n_pc):
            ell = rng.uniform(0.5, 3.0, size=d)
            if g.kernel == "rbf":
                kern = sk.RBF(length_scale=ell, length_scale_bounds=np.outer(hi - lo, (0.01, 100)))
            else:
                kern = sk.Matern(length_scale=ell, length_scale_bounds=np.outer(hi - lo, (0.01, 100)), nu=g.nu)
            if g.constant:
                kern = kern + sk.ConstantKernel(constant_value=rng.uniform(0.5, 2.0), constant_value_bounds=(1e-3, 10))
            if g.noise:
                eta = float(np.exp(rng.uniform(np.log(1e-3), np.log(0.3))))
                kern = kern + sk.WhiteKernel(noise_level=eta, noise_level_bounds=(1e-4, 1))
            gpr = GaussianProcessRegressor(kernel=kern, alpha=1e-10, copy_X_train=False,
                                           optimizer="fmin_l_bfgs_b" if spec.optimise else None,
                                           n_restarts_optimizer=0, random_state=spec.seed)
            emulators.append(gpr.fit(design, Ypca[:, p]))
        results[g.name] = {"PCA": {"pca": pca, "scaler": scaler}, "emulators": emulators}

    n_bins = {lab: v["y"].shape[0] for lab, v in prediction.items()}
    cfg = make_config(spec, n_bins)
    merged = list(cfg.sort_observables_in_matrix.entries)

    data: dict[str, dict[str, np.ndarray]] = {}
    for lab in merged:
        yv = prediction[lab]["y"][:, 0] + 0.1
        data[lab] = {"y": yv, "y_err": 0.1 * np.abs(yv), "xmin": np.arange(len(yv), dtype=float),
                     "xmax": np.arange(len(yv), dtype=float) + 1.0}
    y = np.concatenate([data[lab]["y"] for lab in merged])
    y_err = np.concatenate([data[lab]["y_err"] for lab in merged])

    sigma_exp = None
    if spec.dense_sigma_exp:
        sigma_exp = np.diag(y_err ** 2)
        for (_, so, _) in cfg.sort_observables_in_matrix.entries.values():
            e = y_err[so]
            sigma_exp[so, so] += 0.3 * np.outer(e, e)

    wrng = np.random.default_rng(2000 + spec.seed)
    theta = wrng.uniform(0.05, 0.95, size=(spec.n_walkers, d))
    if spec.oob_fraction > 0:
        n_bad = max(1, int(round(spec.oob_fraction * spec.n_walkers)))
        rows = wrng.choice(spec.n_walkers, size=n_bad, replace=False)
        cols = wrng.integers(0, d, size=n_bad)
        theta[rows, cols] = np.where(wrng.random(n_bad) < 0.5, -0.01, 1.01)
        # one row exactly on the boundary: the prior is strict (log_posterior.py:63)
        theta[rows[0], cols[0]] = 1.0

    observables = {"Prediction": prediction, "Data": data, "Design": design}
    return SyntheticAnalysis(spec=spec, names=[f"p{i}" for i in range(d)], min=lo, max=hi, design=design,
                             observables=observables, emulation_results=results, emulation_config=cfg,
                             experimental_results={"y": y, "y_err": y_err}, sigma_exp=sigma_exp, theta=theta,
                             labels_by_group=labels)
