"""Drop-in for the reference's ``bayesian_inference.mcmc.run_mcmc`` (mcmc.py:34-134) on the B200 log-posterior path.

Same sequence as the reference: random start in the prior box -> first half of burn-in -> re-seed the walkers at the
``n_walkers`` best distinct log-probabilities of the chain (mcmc.py:99) -> second half of burn-in -> production ->
``mcmc.h5`` with the keys ``chain, acceptance_fraction, log_prob, autocorrelation_time[, design_point,
experimental_pseudodata]`` (mcmc.py:111-125).

Differences, all forced by what is (not) installed here or by the move to the GPU:
* no ``multiprocessing`` pool: ``initialize_pool_variables`` loads the emulators into HBM once and the sampler
  evaluates all proposals of a half-step in one call (``StretchEnsemble`` / ``DeviceStretchEnsemble``, restatements
  of emcee's stretch move; emcee itself is not installed in this image); by default the ensemble itself lives on the
  GPU and only the chain-independent random numbers come from the host;
* with ``torch.distributed`` initialised (one process per GPU) the proposals are sharded over the ranks and one
  all-gather of log-probabilities closes each half-step;
* ``seed`` makes the run reproducible (the reference never seeds, mcmc.py:88);
* configuration / emulator / data loading is the reference's own code (out of scope): pass the loaded
  objects in, or leave them ``None`` to have them loaded through the reference package if it is importable;
* ``mcmc.h5`` is written by the reference's writer (silx) when importable, else by an h5py restatement with the same
  keys and nesting; with neither library installed the same dictionary goes to ``mcmc.npz``;
* the sampler object pickled to ``config.sampler_outputfile`` exposes ``get_chain / get_log_prob /
  acceptance_fraction / flatchain`` like the emcee sampler the reference pickles (mcmc.py:127-132).
"""
from __future__ import annotations

import logging
import os
import pickle
from typing import Any, Optional

import numpy as np

from . import log_posterior
from .ensemble import BatchedClosureEnsemble, DeviceStretchEnsemble, ShardedLogProb, StretchEnsemble

logger = logging.getLogger(__name__)


def write_dict_to_h5(results: dict[str, Any], output_dir: str, filename: str) -> str:
    """The reference's ``data_IO.write_dict_to_h5`` (data_IO.py:217-236: silx ``dicttoh5(results, path,
    update_mode='modify')``) restated on h5py: nested dictionaries become groups, arrays and scalars datasets under their
    key, existing entries of the same name are replaced, ``None`` (``autocorrelation_time`` of a short chain, mcmc.py:118)
    is not written.  ``silx.io.dictdump.h5todict`` -- what ``plot_mcmc.py`` reads with -- returns the same dictionary."""
    import h5py

    os.makedirs(output_dir, exist_ok=True)
    path = os.path.join(output_dir, filename)

    def put(group, d):
        for k, v in d.items():
            if v is None:
                continue
            if k in group:
                del group[k]
            if isinstance(v, dict):
                put(group.create_group(k), v)
            else:
                a = np.asarray(v)
                group.create_dataset(k, data=a.astype("S") if a.dtype.kind == "U" else a)

    with h5py.File(path, "a") as f:
        put(f, results)
    return path


def read_dict_from_h5(input_dir: str, filename: str) -> dict[str, Any]:
    """Inverse of ``write_dict_to_h5`` (the reference's ``data_IO.read_dict_from_h5``, data_IO.py:239-257)."""
    import h5py

    def get(group):
        return {k: (get(v) if isinstance(v, h5py.Group) else v[()]) for k, v in group.items()}

    with h5py.File(os.path.join(input_dir, filename), "r") as f:
        return get(f)


def _write_results(output_dict: dict[str, Any], config) -> str:
    """``mcmc.h5`` with the reference's keys (mcmc.py:111-125) through, in this order: the reference's own writer when the
    reference package and silx are importable; the h5py restatement above; and only when neither HDF5 library is installed
    (the case in this image: ``import h5py`` / ``import silx`` fail here and on the GPU box) the same dictionary as
    ``mcmc.npz``.  Only a missing module selects the next writer -- I/O errors propagate."""
    out_dir = getattr(config, "mcmc_output_dir", None) or getattr(config, "output_dir", ".")
    os.makedirs(out_dir, exist_ok=True)
    try:
        from bayesian_inference import data_IO  # the reference's writer, when its dependencies exist
    except ImportError:
        data_IO = None
    if data_IO is not None:
        data_IO.write_dict_to_h5(output_dict, out_dir, "mcmc.h5", verbose=True)
        return os.path.join(out_dir, "mcmc.h5")
    try:
        import h5py  # noqa: F401
    except ImportError:
        h5py = None
    if h5py is not None:
        return write_dict_to_h5(output_dict, out_dir, "mcmc.h5")
    path = os.path.join(out_dir, "mcmc.npz")
    flat = {}
    for k, v in output_dict.items():
        if isinstance(v, dict):
            for kk, vv in v.items():
                flat[f"{k}/{kk}"] = np.asarray(vv)
        elif v is not None:
            flat[k] = np.asarray(v)
    np.savez(path, **flat)
    return path


def run_mcmc(config, closure_index: int = -1, *, emulation_config=None, emulation_results=None,
             experimental_results=None, design_point=None, seed: Optional[int] = None, mode: str = "dense",
             sigma_exp: Optional[np.ndarray] = None, write: bool = True, sampler: str = "device") -> dict[str, Any]:
    """Run burn-in + production and return (and write) the reference's output dictionary.

    ``sampler='device'`` keeps the ensemble on the GPU (``DeviceStretchEnsemble``: propose / accept kernels around the
    log-posterior kernels, no host round trip per half-step); ``sampler='host'`` runs the numpy stretch move and calls
    the log-posterior once per half-step.  Both consume the random stream identically and produce the same chain."""
    par = config.analysis_config["parameterization"][config.parameterization]
    names, lo, hi = par["names"], np.asarray(par["min"], dtype=np.float64), np.asarray(par["max"], dtype=np.float64)
    ndim = len(names)

    if emulation_config is None or emulation_results is None or experimental_results is None:
        from bayesian_inference import data_IO, emulation as ref_emulation   # reference package (training / I/O side)

        emulation_config = ref_emulation.EmulationConfig.from_config_file(
            analysis_name=config.analysis_name, parameterization=config.parameterization,
            analysis_config=config.analysis_config, config_file=config.config_file)
        emulation_results = emulation_config.read_all_emulator_groups()
        experimental_results = data_IO.data_array_from_h5(config.output_dir, "observables.h5", pseudodata_index=closure_index,
                                                          observable_filter=emulation_config.observable_filter)
        if closure_index >= 0:
            design_point = data_IO.design_array_from_h5(config.output_dir, filename="observables.h5", validation_set=True)[closure_index]

    try:
        import torch.distributed as dist

        distributed = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
    except Exception:
        distributed = False

    log_posterior.initialize_pool_variables(lo, hi, emulation_config, emulation_results, experimental_results, None,
                                            sigma_exp=sigma_exp, mode=mode)
    rng = np.random.RandomState(seed)
    if sampler in ("device", "device_graph"):      # device_graph: blocks of steps replayed as one CUDA graph, same chain
        sampler = DeviceStretchEnsemble(config.n_walkers, ndim, log_posterior.device(), mode=mode, random_state=rng,
                                        graph=sampler.endswith("graph"))
    elif sampler in ("device_philox", "device_philox_graph"):      # device-side random numbers: not emcee's stream
        from .ensemble import DevicePhiloxEnsemble
        sampler = DevicePhiloxEnsemble(config.n_walkers, ndim, log_posterior.device(), mode=mode,
                                       seed=int(rng.randint(0, 2**31 - 1)), graph=sampler.endswith("graph"))
    elif sampler == "host":
        fn = log_posterior.log_posterior_walkers
        if distributed:
            fn = ShardedLogProb(fn)
        sampler = StretchEnsemble(config.n_walkers, ndim, fn, random_state=rng)
    else:
        raise ValueError("sampler must be 'device', 'device_graph', 'host', 'device_philox' or 'device_philox_graph'")

    logger.info("Initializing sampler...")
    random_pos = rng.uniform(lo, hi, (config.n_walkers, ndim))                       # mcmc.py:88
    nburn0 = config.n_burn_steps // 2
    logger.info("Starting initial burn-in...")
    sampler.run_mcmc(random_pos, nburn0, n_logging_steps=config.n_logging_steps, logger=logger)
    logger.info("Resampling walker positions...")
    X0 = sampler.flatchain[np.unique(sampler.flatlnprobability, return_index=True)[1][-config.n_walkers:]]   # mcmc.py:99
    sampler.reset()
    X0 = sampler.run_mcmc(X0, config.n_burn_steps - nburn0, n_logging_steps=config.n_logging_steps, logger=logger)[0]
    sampler.reset()
    logger.info("Burn-in complete.")
    logger.info("Starting production...")
    sampler.run_mcmc(X0, config.n_sampling_steps, n_logging_steps=config.n_logging_steps, logger=logger)

    output_dict: dict[str, Any] = {
        "chain": sampler.get_chain(),
        "acceptance_fraction": sampler.acceptance_fraction,
        "log_prob": sampler.get_log_prob(),
    }
    try:
        output_dict["autocorrelation_time"] = sampler.get_autocorr_time()
    except Exception as e:   # as the reference: keep going (mcmc.py:116-119)
        output_dict["autocorrelation_time"] = None
        logger.info(f"Could not compute autocorrelation time: {str(e)}")
    if closure_index >= 0:
        output_dict["design_point"] = design_point
        output_dict["experimental_pseudodata"] = experimental_results
    rank0 = (not distributed) or dist.get_rank() == 0
    if write and rank0:
        path = _write_results(output_dict, config)
        logger.info(f"Wrote {path}")
        sampler_file = getattr(config, "sampler_outputfile", None)
        if sampler_file:
            os.makedirs(os.path.dirname(sampler_file) or ".", exist_ok=True)
            with open(sampler_file, "wb") as f:      # mcmc.py:127-132: the sampler object (get_chain, get_log_prob, ...)
                pickle.dump(sampler, f)
    output_dict["n_log_prob_rows"] = sampler.n_log_prob_rows
    logger.info("Done.")
    return output_dict


def run_mcmc_closures(config, pseudodata, *, emulation_config, emulation_results, design_points=None, seeds=None,
                      mode: str = "dense", sigma_exp: Optional[np.ndarray] = None) -> list[dict[str, Any]]:
    """All closure tests of an analysis at once.

    The reference loops ``for i in validation_indices: mcmc.run_mcmc(config, closure_index=i)`` (steer_analysis.py:168-183):
    one full MCMC per validation design point against that point's pseudo-data (``data_IO.data_array_from_h5(...,
    pseudodata_index=i)``, data_IO.py:362-372), always with the same emulators.  Here the pseudo-data sets are loaded side
    by side (``bi_set_data_sets``) and the chains advance in lock-step, one batched log-posterior call per half-step.
    ``pseudodata``: list of ``{'y': (F,), 'y_err': (F,)}``; ``seeds[i]`` seeds chain i, which is then exactly the chain
    ``run_mcmc(config, closure_index=i, seed=seeds[i])`` produces.  Returns one reference-style output dictionary per
    closure test (nothing is written).
    """
    if sigma_exp is not None:
        # the data-set path carries one diag(y_err**2) per pseudo-data set (bi_set_data_sets); a shared dense Sigma_exp
        # would silently be dropped there, and chain i would no longer be run_mcmc(closure_index=i, sigma_exp=...)
        raise ValueError("run_mcmc_closures: a dense sigma_exp is not supported on the batched data-set path; "
                         "run the closure tests one by one with run_mcmc(..., sigma_exp=...)")
    par = config.analysis_config["parameterization"][config.parameterization]
    lo, hi = np.asarray(par["min"], dtype=np.float64), np.asarray(par["max"], dtype=np.float64)
    ndim, W, C_ = len(par["names"]), config.n_walkers, len(pseudodata)
    seeds = list(range(C_)) if seeds is None else list(seeds)
    log_posterior.initialize_pool_variables(lo, hi, emulation_config, emulation_results, pseudodata[0], None,
                                            sigma_exp=sigma_exp, mode=mode)
    dev = log_posterior.device()
    dev.set_data_sets(np.stack([np.asarray(p["y"], dtype=np.float64) for p in pseudodata]),
                      np.stack([np.asarray(p["y_err"], dtype=np.float64) for p in pseudodata]))
    rngs = [np.random.RandomState(sd) for sd in seeds]
    ens = BatchedClosureEnsemble(W, ndim, dev, rngs, mode=mode)
    start = np.stack([rng.uniform(lo, hi, (W, ndim)) for rng in rngs])                               # mcmc.py:88, per chain
    nburn0 = config.n_burn_steps // 2
    ens.run_mcmc(start, nburn0)
    X0 = np.stack([c.flatchain[np.unique(c.flatlnprobability, return_index=True)[1][-W:]] for c in ens.chains])   # mcmc.py:99
    ens.reset()
    X0 = ens.run_mcmc(X0, config.n_burn_steps - nburn0)
    ens.reset()
    ens.run_mcmc(X0, config.n_sampling_steps)
    out = []
    for i, c in enumerate(ens.chains):
        d: dict[str, Any] = {"chain": c.get_chain(), "acceptance_fraction": c.acceptance_fraction, "log_prob": c.get_log_prob()}
        try:
            d["autocorrelation_time"] = c.get_autocorr_time()
        except Exception as e:
            d["autocorrelation_time"] = None
            logger.info(f"Could not compute autocorrelation time: {str(e)}")
        d["design_point"] = None if design_points is None else design_points[i]
        d["experimental_pseudodata"] = pseudodata[i]
        out.append(d)
    return out
