"""Developer check (run under gpurun): every stage of the CUDA path against the numpy oracle, plus rough timings.

    python tools/gpu_check.py [TOY C1 C3A ...]
"""
from __future__ import annotations

import sys
import time

import numpy as np

sys.path.insert(0, ".")
from bayesian_inference_b200 import image, synthetic  # noqa: E402
from bayesian_inference_b200.device import DeviceEmulator  # noqa: E402
from oracle import closed_form as cf  # noqa: E402


def rel(a, b):
    a, b = np.asarray(a), np.asarray(b)
    den = np.maximum(np.abs(b), 1e-300)
    return float(np.max(np.abs(a - b) / den)) if a.size else 0.0


def check(name: str, n_check: int = 24, **over):
    t0 = time.time()
    a = synthetic.generate(name, **over)
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                     experimental_results=a.experimental_results, sigma_exp=a.sigma_exp)
    z = img.to_flat()
    print(f"[{name}] generated in {time.time() - t0:.1f}s: d={img.d} F={img.n_features} "
          f"groups={[(g.n_design, g.n_pc, g.n_features) for g in img.groups]} B={len(img.blocks)}", flush=True)
    dev = DeviceEmulator(img)
    th = a.theta[:n_check]
    # K1
    for g in range(len(img.groups)):
        k = dev.kernel_vectors(g, th)
        print(f"  K1 g{g} kvec rel err      {rel(k, cf.kernel_vectors(z, g, th)):.3e}")
    # K2
    m, v = dev.predict_pc(th)
    mo = np.concatenate([cf.predict_pc(z, g, th)[0] for g in range(len(img.groups))], axis=1)
    vo = np.concatenate([cf.predict_pc(z, g, th)[1] for g in range(len(img.groups))], axis=1)
    print(f"  K2 mean abs err / scale  {np.max(np.abs(m - mo)) / np.max(np.abs(mo)):.3e}")
    print(f"  K2 var  abs err          {np.max(np.abs(v - vo)):.3e}   (min var {vo.min():.3e})")
    # K3
    mean, cov = dev.predict(th)
    mo_, co_ = cf.predict(z, th)
    print(f"  K3 mean rel (fro)        {np.linalg.norm(mean - mo_) / np.linalg.norm(mo_):.3e}")
    print(f"  K3 cov  rel (fro)        {np.linalg.norm(cov - co_) / np.linalg.norm(co_):.3e}")
    for g in range(len(img.groups)):
        mg, cg = dev.predict_group(g, th)
        mgo, cgo = cf.predict_group(z, g, th)
        print(f"  K3 g{g} mean/cov rel      {np.linalg.norm(mg - mgo) / np.linalg.norm(mgo):.3e} "
              f"{np.linalg.norm(cg - cgo) / np.linalg.norm(cgo):.3e}")
    # K4
    ref = cf.log_posterior(z, th, n_div=1.0)
    for mode in ("dense", "structured", "woodbury"):
        if mode != "dense" and a.sigma_exp is not None:
            continue
        lp, info = dev.log_posterior(th, n_div=1.0, mode=mode, return_info=True)
        fin = np.isfinite(ref)
        print(f"  K4 {mode:10s} logp rel err {rel(lp[fin], ref[fin]):.3e}  inf match {np.array_equal(np.isinf(lp), np.isinf(ref))} "
              f"info {np.unique(info)}")
    refb = cf.log_posterior(z, th)
    lpb = dev.log_posterior(th, n_div=None)
    fin = np.isfinite(refb)
    print(f"  K4 batched-n_div logp rel err {rel(lpb[fin], refb[fin]):.3e}")
    print(f"  logp sample {ref[:3]}")
    # timing
    W = a.theta.shape[0]
    for mode in ("dense", "structured", "woodbury"):
        if mode != "dense" and a.sigma_exp is not None:
            continue
        dev.log_posterior(a.theta[:min(W, 512)], mode=mode)
        t0 = time.time()
        dev.log_posterior(a.theta, mode=mode)
        dt = time.time() - t0
        print(f"  time {mode:10s} W={W}: {dt * 1e3:.2f} ms -> {W / dt:.0f} evals/s (host API)")
    t0 = time.time()
    dev.predict_pc(a.theta)
    dt = time.time() - t0
    print(f"  time predict_pc W={W}: {dt * 1e3:.2f} ms -> {W / dt:.0f} evals/s")
    print(f"  chol CTAs {dev.query(5)}  scratch {dev.query(6) / 2**20:.1f} MiB  launches {dev.launches}")
    dev.close()


if __name__ == "__main__":
    names = sys.argv[1:] or ["TOY"]
    for n in names:
        if n.upper() == "TOY":
            check("TOY", oob_fraction=0.1)
            check("TOY", dense_sigma_exp=True)
        else:
            check(n)
