"""Emulator-training benchmark: batched log-marginal-likelihood + gradient evaluations per
second on the device against the roofline, and a whole multi-restart fit, next to scikit-learn / scipy on the host.

    python tools/train_bench.py [--n-design 230] [--d 6] [--n-pc 30] [--restarts 50] [--cpu-seconds 10] [--json out.json]

Useful FP64 work per evaluation (what ``frac`` counts): Cholesky N^3/3 + inverse factor N^3/3 + K^-1 = L^-T L^-1 N^3/3
= N^3 flops (the kernel-matrix build and the gradient contraction are O(N^2 d) and not counted).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from bayesian_inference_b200 import training as tr  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n-design", type=int, default=230)
    ap.add_argument("--d", type=int, default=6)
    ap.add_argument("--n-pc", type=int, default=30)
    ap.add_argument("--restarts", type=int, default=50)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--no-fit", action="store_true")
    ap.add_argument("--json", default=None)
    a = ap.parse_args()
    import torch

    N, d, P, R = a.n_design, a.d, a.n_pc, a.restarts + 1
    rng = np.random.RandomState(0)
    hi = rng.uniform(0.5, 3.0, d)
    X = rng.uniform(0, hi, (N, d))
    Y = np.sin(X / hi @ rng.normal(size=(d, P)) * 2.0).T * np.linspace(3, 0.3, P)[:, None] + 0.2 * rng.normal(size=(P, N))
    spec = tr.kernel_spec({"matern": {"nu": 1.5, "length_scale_bounds_factor": [0.01, 100]},
                           "constant": {"constant_value": 1.0, "constant_value_bounds": [1e-3, 10]},
                           "noise": {"args": {"noise_level": 0.25, "noise_level_bounds": [1e-4, 1]}}}, np.zeros(d), hi)
    B = P * R
    theta = rng.uniform(spec.bounds[:, 0], spec.bounds[:, 1], (B, spec.n_hyper))
    yidx = np.repeat(np.arange(P, dtype=np.int32), R)
    out = {"workload": f"N={N} d={d} P={P} starts={R} (B={B} hyper-parameter vectors per batch)"}
    with tr.DeviceGPTrainer(X, Y, spec, 1e-10) as t:
        th_t = torch.tensor(theta, device="cuda")
        yi_t = torch.tensor(yidx, device="cuda")
        for _ in range(3):
            t.log_marginal_likelihood_tensor(th_t, yi_t)
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps + 1)]
        ev[0].record()
        for i in range(a.steps):
            lml, grad = t.log_marginal_likelihood_tensor(th_t, yi_t)
            ev[i + 1].record()
        torch.cuda.synchronize()
        ms = ev[0].elapsed_time(ev[-1]) / a.steps
        peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))) \
            if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else {}
        peak = 37.1        # TFLOP/s, FP64 DMMA measured (profiles/r01_fp64_peak_microbench.txt)
        flops = float(N) ** 3 * B
        out.update({"gpu_ms_per_batch": ms, "gpu_evals_per_s": B / ms * 1e3,
                    "roofline": {"bound": "tensor", "achieved": flops / ms / 1e9, "peak": peak, "unit": "TFLOP/s",
                                 "frac": flops / ms / 1e9 / peak}, "finite": int(torch.isfinite(lml).sum())})
        # host: the oracle restatement of scikit-learn's objective (scipy LAPACK), one evaluation after another
        from oracle import gp_lml
        t0 = time.time()
        n = 0
        while time.time() - t0 < a.cpu_seconds:
            gp_lml.log_marginal_likelihood(X, Y[yidx[n % B]], theta[n % B], spec.kind, spec.has_const, spec.has_white, 1e-10)
            n += 1
        cpu = n / (time.time() - t0)
        out["cpu_baseline"] = {"value": cpu, "unit": "evals/s", "cores": int(os.environ.get("OMP_NUM_THREADS", os.cpu_count() or 1)),
                               "kind": "port", "sample": f"{n} evaluations"}
        out["speedup_objective"] = out["gpu_evals_per_s"] / cpu
    if not a.no_fit:
        t0 = time.time()
        fit = tr.fit_gps(X, Y, spec, 1e-10, n_restarts=a.restarts, seed=0)
        dt = time.time() - t0
        out["fit"] = {"seconds": dt, "batches": fit.n_batches, "evaluations": fit.n_evaluations,
                      "lml_sum": float(fit.lml.sum()), "cpu_seconds_estimate": fit.n_evaluations / cpu}
    print(json.dumps(out))
    if a.json:
        with open(a.json, "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
