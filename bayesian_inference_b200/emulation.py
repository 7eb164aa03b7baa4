"""Drop-in for the hot half of the reference's ``bayesian_inference.emulation`` (emulation.py:214-548).

Same names, arguments and return values as the reference:

    predict(parameters, emulation_config, merge_predictions_over_groups=True,
            emulation_group_results=None, emulator_cov_unexplained=None)          emulation.py:410-462
    predict_emulation_group(parameters, results, emulation_group_config,
                            emulator_group_cov_unexplained=None)                  emulation.py:466-548
    compute_emulator_cov_unexplained(emulation_config, emulation_results)         emulation.py:214-224
    compute_emulator_group_cov_unexplained(emulation_group_config, result)        emulation.py:227-251
    nd_block_diag(arrays)                                                         emulation.py:254-270

Everything per-theta runs in the CUDA library through a ``DeviceEmulator``; there is no numpy implementation behind
these functions.  Training, configuration parsing and I/O (emulation.py:38-211, :551-709) are out of scope: keep
importing those from the reference package -- the objects they produce are exactly what these functions consume.

Faithfulness notes: the truncated-PC covariance is divided by the number of rows in the call (F3) -- use
``n_div=1`` (keyword, extension) for what an emcee run evaluates; the merged covariance is block diagonal per
observable (F2); ``pca.mean_`` is not added (F8).
"""
from __future__ import annotations

import pickle
from typing import Any, Mapping, Optional

import numpy as np

from . import image as _image
from .device import DeviceEmulator
from .observable_map import ObservableBlockMap

_CACHE: dict[tuple, tuple[Any, DeviceEmulator]] = {}
_CACHE_MAX = 8
_DEVICE = 0


def set_device(device: int) -> None:
    """CUDA device used for emulators built from now on (one process per GPU: call with LOCAL_RANK)."""
    global _DEVICE
    _DEVICE = int(device)


def clear_cache() -> None:
    for _, dev in _CACHE.values():
        dev.close()
    _CACHE.clear()


def _remember(key: tuple, anchor: Any, build) -> DeviceEmulator:
    hit = _CACHE.get(key)
    if hit is not None and hit[0] is anchor:
        return hit[1]
    if len(_CACHE) >= _CACHE_MAX:
        old = next(iter(_CACHE))
        _CACHE.pop(old)[1].close()
    dev = build()
    _CACHE[key] = (anchor, dev)
    return dev


def read_emulators(emulation_group_config) -> dict[str, Any]:
    """emulation.py:196-202: the pickled {'PCA': ..., 'emulators': [...]} of one group."""
    with open(emulation_group_config.emulation_outputfile, "rb") as f:
        return pickle.load(f)


def _group_results(emulation_config, emulation_group_results) -> dict[str, Any]:
    if emulation_group_results is None:
        emulation_group_results = {}
    out = {}
    for name, gc in emulation_config.emulation_groups_config.items():
        res = emulation_group_results.get(name)
        out[name] = res if res is not None else read_emulators(gc)   # emulation.py:441-442
    return out


def _same_as_image(passed: Optional[Mapping[str, Any]], emulation_config, results: Mapping[str, Any]) -> bool:
    """True when a caller-supplied ``emulator_cov_unexplained`` is what the PCA objects imply anyway (the usual case:
    callers pass the output of ``compute_emulator_cov_unexplained``), so the default image serves it."""
    if not passed:
        return True
    for name, gc in emulation_config.emulation_groups_config.items():
        if name not in passed:
            raise KeyError(name)            # the reference indexes the dictionary with every group name (emulation.py:446)
        own = compute_emulator_group_cov_unexplained(gc, results[name])
        if not _close(passed[name], own):
            return False
    return True


def _close(a, b) -> bool:
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return a.shape == b.shape and bool(np.allclose(a, b, rtol=1e-13, atol=1e-14 * max(float(np.max(np.abs(b))) if b.size else 0.0, 1e-300)))


def device_emulator(emulation_config, emulation_group_results: Mapping[str, Any], *, min=None, max=None,
                    experimental_results=None, sigma_exp=None, emulator_cov_unexplained=None) -> DeviceEmulator:
    """Build (or fetch from the cache) the device image of a whole analysis.  A non-empty ``emulator_cov_unexplained``
    that differs from what the PCA objects imply gets an image of its own (emulation.py:445-448)."""
    first = next(iter(emulation_group_results.values()))
    d = int(np.asarray(first["emulators"][0].X_train_).shape[1])
    lo = np.full(d, -np.inf) if min is None else np.asarray(min, dtype=np.float64)
    hi = np.full(d, np.inf) if max is None else np.asarray(max, dtype=np.float64)
    override = None if _same_as_image(emulator_cov_unexplained, emulation_config, emulation_group_results) \
        else emulator_cov_unexplained
    key = ("analysis", id(emulation_config), id(emulation_group_results), _DEVICE, experimental_results is not None,
           None if override is None else id(override))

    def build():
        img = _image.pack(emulation_config, emulation_group_results, min=lo, max=hi,
                          experimental_results=experimental_results, sigma_exp=sigma_exp,
                          emulator_cov_unexplained=override)
        return DeviceEmulator(img, device=_DEVICE)

    if experimental_results is not None:
        return build()   # owned by the caller (log_posterior module), not cached here
    return _remember(key, emulation_group_results if override is None else override, build)


def predict(parameters: np.ndarray, emulation_config, merge_predictions_over_groups: bool = True,
            emulation_group_results: Optional[dict[str, dict[str, Any]]] = None,
            emulator_cov_unexplained: Optional[dict[str, Any]] = None, *, n_div: Optional[float] = None
            ) -> dict[str, Any]:
    """Emulator central values and covariance at ``parameters`` (n_samples, n_parameters).

    merged:     {'central_value': (n, F), 'cov': (n, F, F)}
    not merged: {group: {'central_value': (n, F_g), 'cov': (n, F_g, F_g)}}
    ``emulator_cov_unexplained``: as in the reference (emulation.py:445-448) a non-empty ``{group: (F_g, F_g)}`` is used
    instead of the truncated-PC covariance the PCA objects imply; ``None`` / empty -> computed from the PCA objects (once,
    into the device image; the reference recomputes it per call).
    """
    parameters = np.atleast_2d(np.asarray(parameters, dtype=np.float64))
    results = emulation_group_results if emulation_group_results else _group_results(emulation_config, emulation_group_results)
    dev = device_emulator(emulation_config, results, emulator_cov_unexplained=emulator_cov_unexplained)
    if not merge_predictions_over_groups:
        out = {}
        for gi, name in enumerate(dev.group_names):
            mean, cov = dev.predict_group(gi, parameters, n_div=n_div)
            out[name] = {"central_value": mean, "cov": cov}
        return out
    mean, cov = dev.predict(parameters, n_div=n_div)
    return {"central_value": mean, "cov": cov}


def predict_emulation_group(parameters: np.ndarray, results: dict[str, Any], emulation_group_config,
                            emulator_group_cov_unexplained: Optional[np.ndarray] = None, *,
                            n_div: Optional[float] = None) -> dict[str, np.ndarray]:
    """One emulation group in its own feature order (what plot_emulation.py:312,357 calls on the whole design)."""
    parameters = np.atleast_2d(np.asarray(parameters, dtype=np.float64))
    override = emulator_group_cov_unexplained
    if override is not None and _close(override, compute_emulator_group_cov_unexplained(emulation_group_config, results)):
        override = None                    # what the PCA object implies anyway: the default image serves it
    key = ("group", id(results), int(emulation_group_config.n_pc), _DEVICE, None if override is None else id(override))

    def build():
        g = _image.pack_group("group", results, int(emulation_group_config.n_pc), cov_unexplained=override)
        Fg, d = g.n_features, g.X.shape[1]
        img = _image.EmulatorImage(d=d, min=np.full(d, -np.inf), max=np.full(d, np.inf), groups=[g],
                                   blocks=np.array([[0, 0, 0, Fg]], dtype=np.int64), n_features=Fg)
        return DeviceEmulator(img, device=_DEVICE)

    dev = _remember(key, results if override is None else override, build)
    mean, cov = dev.predict_group(0, parameters, n_div=n_div)
    return {"central_value": mean, "cov": cov}


def compute_emulator_group_cov_unexplained(emulation_group_config, emulation_group_result) -> np.ndarray:
    """S_u diag(explained_variance_[P:]) S_u^T (emulation.py:227-251).  Theta independent: host numpy, run once."""
    pca = emulation_group_result["PCA"]["pca"]
    P = int(emulation_group_config.n_pc)
    S = np.asarray(pca.components_, dtype=np.float64).T[:, P:]
    return S.dot(np.diag(np.asarray(pca.explained_variance_, dtype=np.float64)[P:]).dot(S.T))


def compute_emulator_cov_unexplained(emulation_config, emulation_results) -> dict[str, np.ndarray]:
    """{group: unexplained covariance}.  The reference's version forgets to return it (emulation.py:214-224, F4)."""
    return {name: compute_emulator_group_cov_unexplained(gc, emulation_results[name])
            for name, gc in emulation_config.emulation_groups_config.items()}


def nd_block_diag(arrays: list[np.ndarray]) -> np.ndarray:
    """Block-diagonal stack over the last two axes (emulation.py:254-270); used by tests and callers, not the GPU path."""
    shapes = np.array([a.shape for a in arrays])
    lead = arrays[0].shape[:-2]
    out = np.zeros(lead + tuple(shapes[:, -2:].sum(axis=0)), dtype=arrays[0].dtype)
    r = c = 0
    for a in arrays:
        out[..., r:r + a.shape[-2], c:c + a.shape[-1]] = a
        r += a.shape[-2]
        c += a.shape[-1]
    return out


SortEmulationGroupObservables = ObservableBlockMap
