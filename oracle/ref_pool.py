"""ORACLE / CPU baseline (test infrastructure, not product code) -- the reference's own log-posterior on host cores.

Drives the UNMODIFIED reference modules in ``baseline/_ref`` (oracle/make_ref.py) the way ``mcmc.run_mcmc`` does
(mcmc.py:72-85): a ``multiprocessing`` **spawn** pool whose workers are initialised with the reference's
``log_posterior.initialize_pool_variables`` and then evaluate ``log_posterior.log_posterior(theta_row)`` one walker per
task, BLAS pinned to one thread per worker.  The only thing added is that each worker builds the initialiser's arguments
itself (the synthetic analysis is unpickled from a temporary file and the duck-typed config is made there), because the
harness objects hold module references that do not pickle.

Used by ``bench.py`` (``cpu_baseline`` / ``--impl reference``) and tests; never by the product path.
"""
from __future__ import annotations

import os
import pickle
import tempfile

import numpy as np

_LP = None


def _init(path: str, ref_src: str) -> None:
    global _LP
    for k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[k] = "1"
    try:
        from threadpoolctl import threadpool_limits

        threadpool_limits(1)
    except Exception:
        pass
    import logging

    logging.disable(logging.INFO)
    from oracle import ref_harness as rh

    rh.REFERENCE_SRC = ref_src
    with open(path, "rb") as f:
        analysis = pickle.load(f)
    _, lp, _ = rh.import_reference()
    cfg = rh.ReferenceConfig(analysis, f"pool{os.getpid()}")
    # the reference's own initialiser, with the arguments mcmc.py:77-78 passes
    lp.initialize_pool_variables(analysis.min, analysis.max, cfg, analysis.emulation_results,
                                 analysis.experimental_results, None)
    _LP = lp


def _eval(row: np.ndarray) -> float:
    return float(_LP.log_posterior(np.ascontiguousarray(row))[0])        # one walker per call, as emcee drives it


def _eval_many(rows: np.ndarray) -> list:
    return [_eval(r) for r in rows]


class ReferencePool:
    def __init__(self, analysis, processes: int | None = None):
        import multiprocessing as mp

        from oracle import ref_harness as rh

        if not os.path.isfile(os.path.join(rh.REFERENCE_SRC, "bayesian_inference", "log_posterior.py")):
            raise RuntimeError(f"no reference modules under {rh.REFERENCE_SRC}")
        if analysis.sigma_exp is not None:
            raise ValueError("the reference's log_posterior only knows diag(y_err**2) (log_posterior.py:90-94)")
        self.n = int(processes or os.cpu_count() or 1)
        tmp = tempfile.NamedTemporaryFile(suffix=".pkl", delete=False)
        pickle.dump(analysis, tmp)
        tmp.close()
        self._path = tmp.name
        self._pool = mp.get_context("spawn").Pool(self.n, initializer=_init, initargs=(tmp.name, rh.REFERENCE_SRC))

    def map_rows(self, rows, chunksize: int = 1) -> np.ndarray:
        return np.asarray(self._pool.map(_eval, list(rows), chunksize=chunksize))

    def one_chunk(self, rows) -> np.ndarray:
        """All rows on ONE worker (the per-core rate)."""
        return np.asarray(self._pool.apply(_eval_many, (np.asarray(rows),)))

    def close(self) -> None:
        self._pool.close()
        self._pool.join()
        try:
            os.unlink(self._path)
        except OSError:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
