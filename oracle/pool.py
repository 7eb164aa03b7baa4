"""ORACLE (test infrastructure, not product code) -- the numpy oracle spread over host processes.

``OraclePool(z).log_posterior(X)`` evaluates ``closed_form.log_posterior(z, X[i:i+1], n_div=1.0)`` row by row -- one
theta per task, single-threaded workers, i.e. how the reference farms walkers to its ``multiprocessing`` pool
(mcmc.py:77-85) -- so that a 100-walker x 1000-step chain driven by the oracle finishes in minutes.  Used only by
tests/ and tools/chain_trace.py (the checker side of the chain-identity runs).
"""
from __future__ import annotations

import os
import tempfile

import numpy as np

_Z = None


def _init(path: str) -> None:
    global _Z
    for k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[k] = "1"
    try:
        from threadpoolctl import threadpool_limits

        threadpool_limits(1)
    except Exception:
        pass
    with np.load(path, allow_pickle=False) as z:
        _Z = {k: z[k] for k in z.files}


def _rows(X: np.ndarray) -> np.ndarray:
    from oracle import closed_form as cf

    return np.array([cf.log_posterior(_Z, X[i:i + 1], n_div=1.0)[0] for i in range(X.shape[0])])


class OraclePool:
    def __init__(self, z, processes: int | None = None):
        import multiprocessing as mp

        self.n = int(processes or os.cpu_count() or 1)
        tmp = tempfile.NamedTemporaryFile(suffix=".npz", delete=False)
        tmp.close()
        np.savez(tmp.name, **{k: np.asarray(v) for k, v in z.items()})
        self._path = tmp.name
        # spawn: the parent may hold a CUDA context (pytest -m gpu), which must not be forked
        self._pool = mp.get_context("spawn").Pool(self.n, initializer=_init, initargs=(tmp.name,))

    def log_posterior(self, X: np.ndarray) -> np.ndarray:
        X = np.ascontiguousarray(np.atleast_2d(X), dtype=np.float64)
        parts = [p for p in np.array_split(X, min(self.n, max(1, X.shape[0]))) if p.shape[0]]
        return np.concatenate(self._pool.map(_rows, parts, chunksize=1)) if parts else np.empty(0)

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
