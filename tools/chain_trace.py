"""Accepted-chain identity at BASELINE configs[0] (north_star: "the same seed produces an identical accepted-chain trace
for at least the first 1,000 steps"): the C1 shape (jet_substructure.yaml-shaped: d=4, n_design=230, n_pc=30, n_obs=200,
100 walkers, seed 42; reference driver mcmc.py:83-107), one chain driven by the CUDA log-posterior, one by the numpy
oracle (row by row over a pool of host processes), both under the same ``StretchEnsemble`` and RandomState.

    python tools/chain_trace.py [--steps 1000] [--workload C1] [--walkers 100] [--seed 42] > profiles/r02_chain_trace_c1.json
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C1")
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--walkers", type=int, default=0)
    ap.add_argument("--seed", type=int, default=42)
    ap.add_argument("--modes", default="dense,structured")
    a = ap.parse_args()

    from bayesian_inference_b200 import image, synthetic
    from bayesian_inference_b200.device import DeviceEmulator
    from bayesian_inference_b200.ensemble import DeviceStretchEnsemble, StretchEnsemble
    from oracle.pool import OraclePool

    spec = synthetic.config(a.workload)
    if a.walkers:
        spec.n_walkers = a.walkers
    an = synthetic.generate(spec)
    img = image.pack(an.emulation_config, an.emulation_results, min=an.min, max=an.max,
                     experimental_results=an.experimental_results, sigma_exp=an.sigma_exp)
    z = img.to_flat()
    W, d = spec.n_walkers, spec.d
    x0 = np.random.RandomState(a.seed).uniform(an.min, an.max, (W, d))          # mcmc.py:88: random start in the prior box
    out = {"workload": spec.name, "walkers": W, "steps": a.steps, "seed": a.seed, "d": d, "n_obs": img.n_features,
           "n_design": spec.n_design, "n_pc": spec.n_pc_total, "cores": os.cpu_count()}
    t0 = time.perf_counter()
    with OraclePool(z) as pool:
        cpu = StretchEnsemble(W, d, pool.log_posterior, seed=a.seed)
        cpu.run_mcmc(x0, a.steps)
    out["oracle_seconds"] = time.perf_counter() - t0
    ref_chain, ref_lp = cpu.get_chain(), cpu.get_log_prob()
    with DeviceEmulator(img) as dev:
        for mode in a.modes.split(","):
            for name, mk in (("host_sampler", lambda: StretchEnsemble(W, d, lambda X: dev.log_posterior(X, n_div=1.0, mode=mode), seed=a.seed)),
                             ("device_sampler", lambda: DeviceStretchEnsemble(W, d, dev, mode=mode, seed=a.seed))):
                t0 = time.perf_counter()
                s = mk()
                s.run_mcmc(x0, a.steps)
                dt = time.perf_counter() - t0
                same = np.all(s.get_chain() == ref_chain, axis=(1, 2))
                fin = np.isfinite(ref_lp) & np.isfinite(s.get_log_prob())
                rec = {"seconds": dt, "identical_steps": int(a.steps if same.all() else np.argmin(same)),
                       "identical": bool(same.all()),
                       "first_divergent_step": None if same.all() else int(np.argmin(same)),
                       "max_rel_dlogp": float(np.max(np.abs(s.get_log_prob()[fin] - ref_lp[fin]) / np.abs(ref_lp[fin]))),
                       "acceptance_identical": bool(np.array_equal(s.acceptance_fraction, cpu.acceptance_fraction)),
                       "acceptance_mean": float(s.acceptance_fraction.mean())}
                out[f"{mode}_{name}"] = rec
    print(json.dumps(out))


if __name__ == "__main__":
    main()
