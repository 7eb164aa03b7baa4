"""Chain throughput of mcmc-style sampling: host-side stretch move calling the CUDA log-posterior once
per half-step vs the device-resident sampler.  python tools/chain_bench.py [--workload C1] [--steps 500] [--mode dense]"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C1")
    ap.add_argument("--steps", type=int, default=500)
    ap.add_argument("--walkers", type=int, default=0)
    ap.add_argument("--mode", default="dense")
    ap.add_argument("--block", type=int, default=64)
    ap.add_argument("--closures", type=int, default=0, help="also run N closure chains batched (SURVEY f3)")
    a = ap.parse_args()
    import torch

    from bayesian_inference_b200 import image, synthetic
    from bayesian_inference_b200.device import DeviceEmulator
    from bayesian_inference_b200.ensemble import DevicePhiloxEnsemble, DeviceStretchEnsemble, StretchEnsemble

    spec = synthetic.config(a.workload)
    if a.walkers:
        spec.n_walkers = a.walkers
    an = synthetic.generate(spec)
    img = image.pack(an.emulation_config, an.emulation_results, min=an.min, max=an.max,
                     experimental_results=an.experimental_results, sigma_exp=an.sigma_exp)
    W, d = spec.n_walkers, spec.d
    x0 = np.random.RandomState(1).uniform(0.3, 0.7, (W, d))
    out = {"workload": spec.name, "walkers": W, "steps": a.steps, "mode": a.mode, "n_obs": img.n_features}
    with DeviceEmulator(img, max_batch=max(W, a.closures * W)) as dev:
        for name, mk in (("host", lambda: StretchEnsemble(W, d, lambda X: dev.log_posterior(X, n_div=1.0, mode=a.mode), seed=3)),
                         ("device", lambda: DeviceStretchEnsemble(W, d, dev, mode=a.mode, seed=3, block=a.block)),
                         ("device_philox", lambda: DevicePhiloxEnsemble(W, d, dev, mode=a.mode, seed=3, block=a.block)),
                         ("device_philox_graph", lambda: DevicePhiloxEnsemble(W, d, dev, mode=a.mode, seed=3, block=a.block, graph=True))):
            s = mk()
            s.run_mcmc(x0, max(10, a.block) if "graph" in name else 10)      # warm-up (captures the block graph)
            if "graph" in name:
                s.reset(rewind=True)       # keeps buffers and the captured graph: steady-state replay is what is timed
            else:
                s = mk()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            s.run_mcmc(x0, a.steps)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            out[name] = {"seconds": dt, "steps_per_s": a.steps / dt, "evals_per_s": (a.steps + 1) * W / dt,
                         "acceptance": float(s.acceptance_fraction.mean())}
            out[name + "_chain_sum"] = float(np.sum(s.get_chain()))
        if a.closures:
            from bayesian_inference_b200.ensemble import BatchedClosureEnsemble

            C_ = a.closures
            rng = np.random.default_rng(9)
            y, ye = an.experimental_results["y"], an.experimental_results["y_err"]
            dev.set_data_sets(np.stack([y * (1.0 + 0.03 * rng.normal(size=y.shape)) for _ in range(C_)]), np.stack([ye] * C_))
            starts = np.stack([np.random.RandomState(100 + c).uniform(0.3, 0.7, (W, d)) for c in range(C_)])
            BatchedClosureEnsemble(W, d, dev, [np.random.RandomState(c) for c in range(C_)], mode=a.mode, block=a.block).run_mcmc(starts, 5)
            ens = BatchedClosureEnsemble(W, d, dev, [np.random.RandomState(c) for c in range(C_)], mode=a.mode, block=a.block)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ens.run_mcmc(starts, a.steps)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            out["closures_batched"] = {"chains": C_, "seconds": dt, "chain_steps_per_s": C_ * a.steps / dt,
                                       "evals_per_s": ens.n_log_prob_rows / dt,
                                       "speedup_vs_serial_device_chains": (C_ * a.steps / dt) / out["device"]["steps_per_s"]}
    out["identical_chain"] = out.pop("host_chain_sum") == out.pop("device_chain_sum")
    out["identical_chain_philox_graph"] = out.pop("device_philox_chain_sum") == out.pop("device_philox_graph_chain_sum")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
