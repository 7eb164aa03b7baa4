"""Small target for compute-sanitizer (one tool per gpurun call, B200_PROFILING.md): the smoke call plus one dense /
structured / Woodbury log-posterior at a 3-tile F, the predict API, the sampler kernels, the data-set path and one
trainer call.  Every kernel of the library is launched at least once; results are checked against the oracle so a
silent corruption fails the run as well.

    compute-sanitizer --tool memcheck python tools/sanitize_target.py
"""
from __future__ import annotations

import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main() -> None:
    import torch

    import __graft_entry__ as entry
    from bayesian_inference_b200 import image, synthetic
    from bayesian_inference_b200.device import DeviceEmulator
    from bayesian_inference_b200.ensemble import DevicePhiloxEnsemble, DeviceStretchEnsemble
    from oracle import closed_form as cf

    entry.smoke()
    # a 3-tile problem (F = 70: tiles of 32 -> Fp = 96), ragged observable blocks, a few walkers outside the prior
    spec = synthetic.config("TOY", n_walkers=24, n_design=40)
    spec.groups = [synthetic.GroupSpec("alpha", 4, 5, 7, constant=True), synthetic.GroupSpec("beta", 3, 5, 7, kernel="rbf")]   # F = 70
    a = synthetic.generate(spec, oob_fraction=0.1)
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                     experimental_results=a.experimental_results)
    flat = img.to_flat()
    want = cf.log_posterior(flat, a.theta, n_div=1.0)
    fin = np.isfinite(want)
    with DeviceEmulator(img, device=0) as dev:
        for mode in ("dense", "structured", "woodbury"):
            got = dev.log_posterior(a.theta, n_div=1.0, mode=mode)
            err = float(np.max(np.abs(got[fin] - want[fin]) / np.abs(want[fin])))
            assert err < 1e-9, (mode, err)
        dev.predict(a.theta[:5])
        dev.predict_group(0, a.theta[:5])
        dev.kernel_vectors(0, a.theta[:5])
        F = img.n_features
        rng = np.random.RandomState(0)
        A = rng.normal(size=(2, F, F))
        cov = A @ A.transpose(0, 2, 1) + F * np.eye(F)
        dev.loglikelihood(rng.normal(size=(2, F)), cov)
        y, ye = a.experimental_results["y"], a.experimental_results["y_err"]
        dev.set_data_sets(np.stack([y, y * 1.01]), np.stack([ye, ye]))
        dev.log_posterior_sets(a.theta[:8], np.arange(8) % 2, mode="dense")
        W, d = 16, img.d
        x0 = np.random.RandomState(1).uniform(0.3, 0.7, (W, d))
        DeviceStretchEnsemble(W, d, dev, seed=3, block=4).run_mcmc(x0, 6)
        DevicePhiloxEnsemble(W, d, dev, seed=3, block=4).run_mcmc(x0, 6)
    torch.cuda.synchronize()
    print("sanitize target ok")


if __name__ == "__main__":
    main()
