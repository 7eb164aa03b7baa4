"""Generate the golden fixtures under tests/golden/ by EXECUTING THE REFERENCE (run in the build container only).

    python tests/golden/make_golden.py

For each small synthetic analysis it stores, in one compressed .npz:
  * the flat emulator image (``EmulatorImage.to_flat()``: what bi_create uploads), the walker positions ``theta``;
  * outputs of the unmodified reference functions imported from /root/reference/src (oracle/ref_harness.py):
      ref_mean, ref_cov            emulation.predict(theta, merged)                      emulation.py:410-462
      ref_g{i}_mean, ref_g{i}_cov  emulation.predict(theta, merge=False)[group i]        emulation.py:458-459
      ref_logp_batch               log_posterior.log_posterior(theta)   (one call)       log_posterior.py:42-101
      ref_logp_rows                log_posterior.log_posterior(theta[i]) per row (emcee) mcmc.py:83-85
      ref_pc_mean, ref_pc_var      sklearn GaussianProcessRegressor.predict per PC       emulation.py:496-499
      ref_kvec_g{i}                gpr.kernel_(theta, X_train_) per PC                   sklearn _gpr.py:446
The reference has no golden vectors of its own for this path; these are the pin.
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.abspath(os.path.join(HERE, "..", "..")))

from bayesian_inference_b200 import image, synthetic  # noqa: E402
from bayesian_inference_b200.synthetic import GroupSpec, ProblemSpec  # noqa: E402
from oracle import ref_harness as rh  # noqa: E402

CASES = {
    # name: (spec factory, number of walkers stored)
    "toy": (lambda: synthetic.config("TOY", oob_fraction=0.1), 24),
    "toy_dense_sigma": (lambda: synthetic.config("TOY", dense_sigma_exp=True), 12),
    "mid": (lambda: ProblemSpec("mid", d=4, n_design=64, n_walkers=8, seed=11, oob_fraction=0.13, groups=[
        GroupSpec("wide", 6, 4, 7, nu=2.5),
        GroupSpec("tall", 5, 3, 40, nu=0.5, constant=True, noise=False),
    ]), 8),
}


def make(name: str) -> dict[str, np.ndarray]:
    factory, W = CASES[name]
    a = synthetic.generate(factory())
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                     experimental_results=a.experimental_results, sigma_exp=a.sigma_exp)
    th = a.theta[:W]
    out = {f"img_{k}": v for k, v in img.to_flat().items()}
    out["theta"] = th
    # the reference's log_posterior only knows diag(y_err^2): for the dense-Sigma_exp case the golden log-posterior is
    # the reference's own _loglikelihood applied to (reference prediction - y, reference cov + Sigma_exp)
    pred = rh.reference_predict(a, th)
    out["ref_mean"], out["ref_cov"] = pred["central_value"], pred["cov"]
    per_group = rh.reference_predict(a, th, merge=False)
    for gi, gname in enumerate(img.group_names):
        out[f"ref_g{gi}_mean"] = per_group[gname]["central_value"]
        out[f"ref_g{gi}_cov"] = per_group[gname]["cov"]
    _, lp, _ = rh.import_reference()
    if a.sigma_exp is None:
        out["ref_logp_batch"] = rh.reference_log_posterior(a, th)
        out["ref_logp_rows"] = rh.reference_log_posterior(a, th, per_row=True)
    else:
        inside = np.all((th > a.min) & (th < a.max), axis=1)
        assert inside.all()
        dy = pred["central_value"] - a.experimental_results["y"]
        out["ref_logp_batch"] = np.array([lp._loglikelihood(dy[i], pred["cov"][i] + a.sigma_exp) for i in range(W)])
        rows = []
        for i in range(W):
            p1 = rh.reference_predict(a, th[i:i + 1])
            rows.append(lp._loglikelihood(p1["central_value"][0] - a.experimental_results["y"],
                                          p1["cov"][0] + a.sigma_exp))
        out["ref_logp_rows"] = np.array(rows)
    # per-PC sklearn outputs
    pcm, pcv = [], []
    for gi, gname in enumerate(img.group_names):
        ems = a.emulation_results[gname]["emulators"]
        kv = np.stack([gpr.kernel_(th, gpr.X_train_) for gpr in ems], axis=1)
        out[f"ref_kvec_g{gi}"] = kv
        for gpr in ems:
            mu, sd = gpr.predict(th, return_std=True)
            pcm.append(mu)
            pcv.append(sd ** 2)
    out["ref_pc_mean"] = np.stack(pcm, axis=1)
    out["ref_pc_var"] = np.stack(pcv, axis=1)
    return out


def main() -> None:
    if not rh.available():
        raise SystemExit("the reference checkout is not available: golden fixtures can only be made in the build container")
    for name in CASES:
        data = make(name)
        path = os.path.join(HERE, f"{name}.npz")
        np.savez_compressed(path, **data)
        print(f"{name}: {os.path.getsize(path) / 1024:.0f} KiB, logp rows {data['ref_logp_rows'][:3]}")


if __name__ == "__main__":
    main()
