"""ORACLE (test infrastructure) -- run the *unmodified* reference hot path in the build container.

``/root/reference`` is a pure-Python package; its hot path (``emulation.predict``, ``log_posterior.log_posterior``)
imports and runs here once ``silx.io.dictdump`` -- used only for HDF5 I/O (data_IO.py:32, :232, :251) -- is stubbed
and ``observables.h5`` is served from memory.  This module is how the numpy restatement (closed_form.py) and the
golden fixtures under tests/golden/ are pinned.  It cannot travel to the GPU box (no /root/reference there), so
nothing in ``-m gpu`` tests, ``smoke()`` or ``bench.py`` imports it.
"""
from __future__ import annotations

import os
import sys
import types
from typing import Any

import numpy as np

REFERENCE_SRC = os.environ.get("BI_REFERENCE_SRC", "/root/reference/src")
BASELINE_REF = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref")
_H5_STORE: dict[str, Any] = {}


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_SRC, "bayesian_inference"))


def use_baseline_ref() -> bool:
    """Point the harness at ``baseline/_ref`` (the copy oracle/make_ref.py makes; it travels to the GPU box, where
    /root/reference does not exist).  bench.py's CPU arm calls this and never reads /root/reference."""
    global REFERENCE_SRC
    if not os.path.isfile(os.path.join(BASELINE_REF, "bayesian_inference", "log_posterior.py")):
        return False
    REFERENCE_SRC = BASELINE_REF
    return True


def _install_silx_stub() -> None:
    if "silx.io.dictdump" in sys.modules:
        return
    silx = types.ModuleType("silx")
    io = types.ModuleType("silx.io")
    dd = types.ModuleType("silx.io.dictdump")

    def h5todict(path, *a, **k):
        return _H5_STORE[os.path.normpath(path)]

    def dicttoh5(d, path, *a, **k):
        _H5_STORE[os.path.normpath(path)] = d

    dd.h5todict, dd.dicttoh5 = h5todict, dicttoh5
    silx.io, io.dictdump = io, dd
    sys.modules.update({"silx": silx, "silx.io": io, "silx.io.dictdump": dd})


def import_reference():
    """Returns the reference's (emulation, log_posterior, data_IO) modules."""
    if not available():
        raise RuntimeError(f"reference sources not found under {REFERENCE_SRC}")
    _install_silx_stub()
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    from bayesian_inference import data_IO, emulation, log_posterior  # type: ignore

    return emulation, log_posterior, data_IO


class ReferenceConfig:
    """Duck-typed ``EmulationConfig`` driving the reference's own ``learn_mapping`` (emulation.py:289-344)."""

    def __init__(self, analysis, tag: str):
        emulation, _, data_IO = import_reference()
        self.output_dir = f"/memory/{tag}"
        _H5_STORE[os.path.normpath(os.path.join(self.output_dir, "observables.h5"))] = analysis.observables
        self.emulation_groups_config = {}
        for name, gc in analysis.emulation_config.emulation_groups_config.items():
            self.emulation_groups_config[name] = types.SimpleNamespace(
                n_pc=gc.n_pc,
                observable_filter=data_IO.ObservableFilter(include_list=list(gc.observable_filter.include_list),
                                                           exclude_list=list(gc.observable_filter.exclude_list)),
            )
        self._sorter = None
        self._emulation = emulation

    @property
    def sort_observables_in_matrix(self):
        if self._sorter is None:
            self._sorter = self._emulation.SortEmulationGroupObservables.learn_mapping(self)
        return self._sorter


def reference_predict(analysis, theta: np.ndarray, merge: bool = True):
    emulation, _, _ = import_reference()
    cfg = ReferenceConfig(analysis, analysis.spec.name)
    return emulation.predict(np.asarray(theta, dtype=np.float64), cfg, merge_predictions_over_groups=merge,
                             emulation_group_results=analysis.emulation_results)


def reference_log_posterior(analysis, theta: np.ndarray, per_row: bool = False) -> np.ndarray:
    """The reference's ``log_posterior`` on ``theta``; ``per_row`` calls it once per row as emcee does (mcmc.py:83-85)."""
    _, lp, _ = import_reference()
    cfg = ReferenceConfig(analysis, analysis.spec.name)
    lp.initialize_pool_variables(analysis.min, analysis.max, cfg, analysis.emulation_results,
                                 analysis.experimental_results, None)
    theta = np.asarray(theta, dtype=np.float64)
    if per_row:
        return np.array([lp.log_posterior(np.ascontiguousarray(row))[0] for row in theta])
    return lp.log_posterior(theta)
