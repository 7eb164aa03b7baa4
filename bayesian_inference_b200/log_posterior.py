"""Drop-in for the reference's ``bayesian_inference.log_posterior`` (log_posterior.py:1-146).

    initialize_pool_variables(local_min, local_max, local_emulation_config, local_emulation_results,
                              local_experimental_results, local_emulator_cov_unexplained)     log_posterior.py:26-38
    log_posterior(X)                                                                           log_posterior.py:42-101
    _loglikelihood(y, cov)                                                                     log_posterior.py:104-146

As in the reference the state lives in module globals, because that is what emcee's ``pool``/callable protocol and
``mcmc.run_mcmc`` expect.  ``initialize_pool_variables`` uploads the fitted emulators, the experimental data and the
prior box into HBM once; ``log_posterior`` is then a single C-ABI call per batch of walkers.  No numpy/sklearn
evaluation exists behind these functions: without the CUDA library they raise.

Divisor of the truncated-PC covariance: ``log_posterior(X)`` reproduces the reference literally -- the
divisor is the number of rows of ``X`` inside the prior box, i.e. 1 when emcee calls it per walker.
``log_posterior_walkers(X)`` evaluates every row as its own call (divisor 1); that is the function to hand to
``emcee.EnsembleSampler(..., vectorize=True)`` to get the reference's chain from one batched call per half-step.
"""
from __future__ import annotations

import logging
from typing import Optional

import numpy as np

from . import emulation
from .device import DeviceEmulator

logger = logging.getLogger(__name__)

min = None
max = None
emulation_config = None
emulation_results = None
experimental_results = None
emulator_cov_unexplained = None

_device: Optional[DeviceEmulator] = None
_mode = "dense"


def initialize_pool_variables(local_min, local_max, local_emulation_config, local_emulation_results,
                              local_experimental_results, local_emulator_cov_unexplained, *,
                              sigma_exp: Optional[np.ndarray] = None, mode: str = "dense") -> None:
    """Install the analysis (same positional arguments as the reference) and load it into HBM.

    ``sigma_exp`` (keyword, extension): dense (F, F) experimental covariance replacing ``diag(y_err**2)`` -- the
    reference's TODO at log_posterior.py:90.  ``mode``: 'dense' (reference arithmetic: full F x F Cholesky) or
    'structured' (skips the structurally-zero tiles of the block-diagonal-per-observable covariance; same value).
    """
    global min, max, emulation_config, emulation_results, experimental_results, emulator_cov_unexplained
    global _device, _mode
    min = local_min
    max = local_max
    emulation_config = local_emulation_config
    emulation_results = local_emulation_results
    experimental_results = local_experimental_results
    emulator_cov_unexplained = local_emulator_cov_unexplained
    if _device is not None:
        _device.close()
    _mode = mode
    _device = emulation.device_emulator(local_emulation_config, local_emulation_results, min=local_min, max=local_max,
                                        experimental_results=local_experimental_results, sigma_exp=sigma_exp,
                                        emulator_cov_unexplained=local_emulator_cov_unexplained)


def device() -> DeviceEmulator:
    if _device is None:
        raise RuntimeError("initialize_pool_variables has not been called")
    return _device


def close() -> None:
    global _device
    if _device is not None:
        _device.close()
        _device = None


def _as_2d(X) -> np.ndarray:
    X = np.asarray(X, dtype=np.float64)   # the reference's np.array(X, copy=False, ndmin=2) (numpy>=2 safe, F13)
    if X.ndim < 2:
        X = X.reshape(1, -1)
    return X


def log_posterior(X) -> np.ndarray:
    """(n,) log-posterior of ``X`` (d,) or (n, d); -inf outside the strict prior box (log_posterior.py:63-64)."""
    return device().log_posterior(_as_2d(X), n_div=None, mode=_mode)


def log_posterior_walkers(X) -> np.ndarray:
    """Every row evaluated as the reference evaluates a single walker (divisor 1): for ``vectorize=True`` samplers."""
    return device().log_posterior(_as_2d(X), n_div=1.0, mode=_mode)


def _loglikelihood(y, cov) -> float:
    """-1/2 y^T C^-1 y - sum log diag chol(C), un-normalised (log_posterior.py:104-146).

    A non positive-definite ``cov`` gives NaN (the reference falls through its mistyped ``elif info < 0`` and also
    ends in NaN)."""
    return float(device().loglikelihood(np.asarray(y, dtype=np.float64), np.asarray(cov, dtype=np.float64))[0])
