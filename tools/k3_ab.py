"""Developer A/B helper for K3 (predict() covariance assembly): time bi_predict's covariance launch alone and compare the
result with a stored one.

    BI_B200_LIB=build/lib_x.so python tools/k3_ab.py --workload C3B [--save ref.npy | --check ref.npy]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C3B")
    ap.add_argument("--reps", type=int, default=6)
    ap.add_argument("--save", default=None)
    ap.add_argument("--check", default=None)
    ap.add_argument("--walkers", type=int, default=0)
    a = ap.parse_args()
    import torch

    from bayesian_inference_b200 import image, synthetic
    from bayesian_inference_b200.device import DeviceEmulator

    an = synthetic.generate(synthetic.config(a.workload))
    img = image.pack(an.emulation_config, an.emulation_results, min=an.min, max=an.max,
                     experimental_results=an.experimental_results, sigma_exp=an.sigma_exp)
    F, d = img.n_features, img.d
    dev = DeviceEmulator(img, max_batch=8192)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rng = np.random.default_rng(1)
    W = a.walkers or max(64, min(2048, int((4 << 30) / (8 * F * F))))
    th = torch.from_numpy(np.ascontiguousarray(rng.uniform(0.05, 0.95, size=(W, d)))).cuda()
    mean = torch.empty((W, F), dtype=torch.float64, device="cuda")
    cov = torch.empty((W, F, F), dtype=torch.float64, device="cuda")

    def timed(fn):
        fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(a.reps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        return float(np.median(ts))

    ms_gp = timed(lambda: dev.predict_pc_tensor(th))
    ms = timed(lambda: dev.predict_tensor(th, n_div=1.0, mean=mean, cov=cov))
    k3 = ms - ms_gp
    byt = 8.0 * W * (F + F * F)
    rec = {"lib": os.environ.get("BI_B200_LIB", ""), "workload": a.workload, "walkers": W, "k3_ms": round(k3, 4),
           "gbs": round(byt / k3 * 1e-6, 1)}
    head = cov[:4].cpu().numpy()
    if a.save:
        np.save(a.save, head)
    if a.check:
        ref = np.load(a.check)
        rec["max_rel_diff"] = float(np.max(np.abs(head - ref)) / np.max(np.abs(ref)))
        rec["zero_pattern_equal"] = bool(np.array_equal(head == 0.0, ref == 0.0))
    print(json.dumps(rec))


if __name__ == "__main__":
    main()
